#!/bin/bash
# ncu --set full of the k_conv_rows launches of the second step: bash tools/gpu_ncu_rows.sh TAG [regex] [skip] [count]
TAG=${1:-r02}; RX=${2:-k_conv_rows}; SK=${3:-8}; CN=${4:-8}
mkdir -p gpurun_out
timeout 120 python tools/profile_step.py 2 64 > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$RX -s $SK -c $CN \
    -o gpurun_out/${TAG}_ncu python tools/profile_step.py 2 64 > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu full exit $?"; tail -2 gpurun_out/${TAG}_ncu2.log
