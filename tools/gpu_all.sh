#!/bin/bash
# Full GPU parity suite + stage-2 operator bench + pipeline bench.  Usage: bash tools/gpu_all.sh TAG
TAG=${1:-all}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/${TAG}_gpu_tests.log 2>&1
echo "pytest exit $?"; tail -6 gpurun_out/${TAG}_gpu_tests.log
timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2_ops.json 2> gpurun_out/${TAG}_stage2_ops.log; echo "stage2 bench exit $?"; tail -9 gpurun_out/${TAG}_stage2_ops.log
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; cut -c1-900 gpurun_out/${TAG}_bench.json
