#!/bin/bash
# Quick round-end style check: smoke() + the full GPU parity suite + one short bench line.
TAG=${1:-check}
mkdir -p gpurun_out
timeout 400 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/${TAG}_gpu_tests.log 2>&1
echo "pytest exit $?"; tail -2 gpurun_out/${TAG}_gpu_tests.log
timeout 300 python bench.py --steps 30 --warmup 5 --no-hbm-kernels > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; cut -c1-220 gpurun_out/${TAG}_bench.json
