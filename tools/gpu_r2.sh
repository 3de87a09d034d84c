#!/bin/bash
# Round-2 iteration visit: selected parity tests, then a short bench.  Usage: bash tools/gpu_r2.sh TAG "pytest -k expr" [bench args]
TAG=${1:-r02}; KEXPR=${2:-""}; shift 2
mkdir -p gpurun_out
if [ -n "$KEXPR" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q --timeout 180 -k "$KEXPR" > gpurun_out/${TAG}_tests.log 2>&1
  echo "pytest exit $?"; tail -15 gpurun_out/${TAG}_tests.log
fi
timeout 300 python bench.py "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"; python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_bench.json"))
    r = d["roofline"]
    print("frames/s", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms/step", round(d["ms_per_step"], 3), "clocks", d.get("clocks"))
    print("roofline", {k: r[k] for k in r if k in ("achieved", "frac", "frac_of_burst")}, r.get("ms_per_step_by_class"))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/${TAG}_bench.err").read()[-2000:])
PY
