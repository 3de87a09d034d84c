#!/bin/bash
# Full ncu capture of kernels matching a regex in the profile-step workload.
# Usage: bash tools/gpu_ncu.sh TAG REGEX SKIP COUNT [steps] [batch]
TAG=$1; RE=$2; SKIP=$3; CNT=$4; STEPS=${5:-2}; BATCH=${6:-64}
mkdir -p gpurun_out
python tools/profile_step.py $STEPS $BATCH > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$RE -s $SKIP -c $CNT \
    -o gpurun_out/${TAG} python tools/profile_step.py $STEPS $BATCH > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/${TAG}_ncu.log
