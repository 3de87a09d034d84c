#!/bin/bash
# Full GPU suite, then the launch list (ncu, durations only) of one pass of tools/bench_stage2.py: which kernels make up
# each stage-2 entry point's time.   Usage: bash tools/gpu_s2_launches.sh TAG
TAG=${1:-s2l}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/${TAG}_gpu_tests.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/${TAG}_gpu_tests.log
timeout 200 python tools/bench_stage2.py 512 1 > /dev/null 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/${TAG}_stage2_launches.csv \
    python tools/bench_stage2.py 512 1 > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; wc -l gpurun_out/${TAG}_stage2_launches.csv
