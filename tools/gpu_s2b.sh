#!/bin/bash
TAG=${1:-s2b}
mkdir -p gpurun_out
bash tools/gpu_sweep.sh edge UB_EDGE_FUSED 0 1 0 1
python tools/bench_stage2.py 512 2 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_edge_fused -s 3 -c 2 -o gpurun_out/${TAG}_edge \
    python tools/bench_stage2.py 512 2 > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/${TAG}_ncu.log
