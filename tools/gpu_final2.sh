#!/bin/bash
# Round-end style visit without the conv metrics pass (conv sources unchanged since r02i): smoke(), full GPU parity suite,
# default bench (all keys), reference arm, launch list of one step, stage-2 launch list.
TAG=${1:-final}
mkdir -p gpurun_out
timeout 400 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/${TAG}_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 300 -s > gpurun_out/${TAG}_gpu_tests.log 2>&1
echo "pytest exit $?"; tail -2 gpurun_out/${TAG}_gpu_tests.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; wc -l gpurun_out/${TAG}_bench.json; cut -c1-300 gpurun_out/${TAG}_bench.json
timeout 400 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref exit $?"; cut -c1-200 gpurun_out/${TAG}_bench_ref.json
timeout 120 python tools/profile_step.py 2 64 > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python tools/profile_step.py 2 64 > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu launches exit $?"
timeout 200 python tools/bench_stage2.py 512 1 > /dev/null 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/${TAG}_stage2_launches.csv \
    python tools/bench_stage2.py 512 1 > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu stage2 exit $?"
