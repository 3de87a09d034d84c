#!/bin/bash
TAG=${1:-geom}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stage2.py tests/test_gpu_configs.py -m gpu -x -q --timeout 300 > gpurun_out/${TAG}_tests.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/${TAG}_tests.log
timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2.json 2> gpurun_out/${TAG}_stage2.log; echo "stage2 exit $?"; tail -10 gpurun_out/${TAG}_stage2.log
