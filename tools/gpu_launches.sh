#!/bin/bash
# ncu launch list of two pipeline steps (batch 64): bash tools/gpu_launches.sh TAG
TAG=${1:-r02}
mkdir -p gpurun_out
timeout 120 python tools/profile_step.py 2 64 > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python tools/profile_step.py 2 64 > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu launches exit $?"; tail -2 gpurun_out/${TAG}_plain.log
