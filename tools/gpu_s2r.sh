#!/bin/bash
TAG=${1:-s2r}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_stage2.py tests/test_gpu_cli.py -m gpu -x -q --timeout 120 > gpurun_out/${TAG}_tests.log 2>&1
echo "pytest exit $?"; tail -4 gpurun_out/${TAG}_tests.log
timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2_ops.json 2> gpurun_out/${TAG}_stage2_ops.log; echo "stage2 bench exit $?"; tail -11 gpurun_out/${TAG}_stage2_ops.log | cut -c1-200
for i in 1 2; do python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-hbm-kernels 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['e2e']['value'],1), d['roofline']['ms_per_step_by_class'])"; done
