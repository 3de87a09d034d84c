#!/bin/bash
# Full ncu capture of the HBM-side kernels in the stage-2 operator bench.
TAG=${1:-s2ncu}
mkdir -p gpurun_out
python tools/bench_stage2.py 512 1 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_overlay|k_gray|k_morph_words|k_edge_fused|k_burr_morph' -c 24 -o gpurun_out/${TAG} \
    python tools/bench_stage2.py 512 1 > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/${TAG}_ncu.log | cut -c1-200
