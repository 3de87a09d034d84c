#!/bin/bash
# Quick GPU visit: parity tests + bench + launch list (no full ncu capture).
TAG=${1:-q}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/${TAG}_gpu_tests.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_gpu_tests.log
tail -15 gpurun_out/${TAG}_gpu_tests.log
timeout 240 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-hbm-kernels > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench_err.log
echo "bench exit $?"; cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench_err.log
timeout 120 python tools/profile_step.py 2 64 > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python tools/profile_step.py 2 64 > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu launches exit $?"
