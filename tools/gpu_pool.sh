#!/bin/bash
# fused-pool check: parity tests, then A/B bench of UB_FUSE_POOL=0/1 on the same box, then a launch list
TAG=${1:-pool}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/${TAG}_gpu_tests.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_gpu_tests.log
tail -15 gpurun_out/${TAG}_gpu_tests.log
for f in 0 1 0 1; do
  UB_FUSE_POOL=$f timeout 240 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_f$f.json 2> gpurun_out/${TAG}_bench_err.log
  echo "fuse=$f exit $?"; python -c "import json;d=json.load(open('gpurun_out/${TAG}_bench_f$f.json'));print(d['value'], d['e2e']['value'], d['roofline']['ms_per_step_by_class'])"; tail -3 gpurun_out/${TAG}_bench_err.log
done
timeout 120 python tools/profile_step.py 2 64 > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python tools/profile_step.py 2 64 > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu launches exit $?"
