#!/bin/bash
# per-kernel times (ncu launch list) for several values of an env knob: bash tools/gpu_sweep_launches.sh TAG VAR v1 v2 ...
TAG=$1; VAR=$2; shift 2
mkdir -p gpurun_out
for v in "$@"; do
  env $VAR=$v timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv \
      --log-file gpurun_out/${TAG}_$v.csv python tools/profile_step.py 1 64 > gpurun_out/${TAG}_$v.log 2>&1
  echo "== $VAR=$v"; python tools/launches.py gpurun_out/${TAG}_$v.csv | grep -E "conv_rows|conv_tc|total"
done
