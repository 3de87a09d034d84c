#!/bin/bash
TAG=${1:-s2q}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_stage2.py -m gpu -x -q --timeout 120 > gpurun_out/${TAG}_tests.log 2>&1
echo "pytest exit $?"; tail -5 gpurun_out/${TAG}_tests.log
timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2_ops.json 2> gpurun_out/${TAG}_stage2_ops.log; echo "stage2 bench exit $?"; tail -9 gpurun_out/${TAG}_stage2_ops.log
