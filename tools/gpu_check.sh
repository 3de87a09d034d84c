#!/bin/bash
# One GPU-box visit: parity tests, bench, ncu launch list + full capture of the conv kernels.
# Usage (from repo root, under gpurun): bash tools/gpu_check.sh [tag]
TAG=${1:-r01}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/${TAG}_gpu_tests.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_gpu_tests.log
tail -5 gpurun_out/${TAG}_gpu_tests.log
timeout 240 python bench.py --steps 50 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench_err.log
echo "bench exit $?"; cat gpurun_out/${TAG}_bench.json
timeout 120 python tools/profile_step.py 2 64 > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python tools/profile_step.py 2 64 > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on -k regex:k_conv -s 18 -c 18 \
    -o gpurun_out/${TAG}_conv python tools/profile_step.py 2 64 > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu full exit $?"
