#!/bin/bash
# enhance_image visit: its parity tests + step timings
TAG=${1:-enh}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_enhance.py -m gpu -q --timeout 120 > gpurun_out/${TAG}_tests.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_tests.log
tail -30 gpurun_out/${TAG}_tests.log
timeout 200 python tools/time_enhance.py 16 > gpurun_out/${TAG}_time.log 2>&1; cat gpurun_out/${TAG}_time.log
