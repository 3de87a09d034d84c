#!/bin/bash
TAG=${1:-wide}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_wide.py tests/test_gpu_annotate.py -m gpu -q --timeout 300 > gpurun_out/${TAG}_tests.log 2>&1
echo "pytest exit $?"; tail -25 gpurun_out/${TAG}_tests.log
