#!/bin/bash
# bench only, for several values of an env knob:  bash tools/gpu_sweep.sh TAG VAR v1 v2 ...
TAG=$1; VAR=$2; shift 2
mkdir -p gpurun_out
for v in "$@"; do
  env $VAR=$v python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-hbm-kernels > gpurun_out/${TAG}_$v.json 2> gpurun_out/${TAG}_$v.err
  python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_$v.json"))
print("$VAR=$v", round(d["value"],1), {k:round(x,3) for k,x in d["roofline"]["ms_per_step_by_class"].items()})
PY
done
