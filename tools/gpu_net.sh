#!/bin/bash
# Risky-kernel visit: the net parity tests under a hard timeout, then the bench + launch list.
TAG=${1:-net}
mkdir -p gpurun_out
timeout -s KILL 240 python -m pytest tests/test_gpu_net.py -m gpu -x -q --timeout 100 > gpurun_out/${TAG}_tests.log 2>&1
RC=$?
echo "pytest exit $RC" >> gpurun_out/${TAG}_tests.log
tail -12 gpurun_out/${TAG}_tests.log
if [ $RC -ne 0 ]; then nvidia-smi --query-gpu=name,memory.used --format=csv; exit 0; fi
timeout -s KILL 240 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench_err.log
echo "bench exit $?"; cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench_err.log
timeout -s KILL 120 python tools/profile_step.py 2 64 > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout -s KILL 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python tools/profile_step.py 2 64 > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu launches exit $?"
