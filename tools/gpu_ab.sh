#!/bin/bash
# A/B of one environment switch on the same box: bash tools/gpu_ab.sh TAG VAR "v1 v2 ..." [bench args]
TAG=$1; VAR=$2; VALS=$3; shift 3
mkdir -p gpurun_out
for v in $VALS; do
  env $VAR=$v timeout 300 python bench.py "$@" --no-cpu-baseline --no-hbm-kernels > gpurun_out/${TAG}_${VAR}_$v.json 2> gpurun_out/${TAG}_${VAR}_$v.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_${VAR}_$v.json")); r = d["roofline"]
    print("$VAR=$v frames/s", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), "clk", d["clocks"]["sm_mhz"], "conv ms", round(r["ms_per_step_by_class"]["conv3x3"], 3), "glue", round(r["ms_per_step_by_class"]["net_glue"], 3))
except Exception as e:
    print("$VAR=$v failed", e); print(open("gpurun_out/${TAG}_${VAR}_$v.err").read()[-1500:])
PY
done
