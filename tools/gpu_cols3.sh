#!/bin/bash
# Third visit: full GPU suite, per-operator bench, and a same-box A/B/C of the pipeline bench
# (default = flood hysteresis issued after the net | UB_HYST_LATE=0 = flood beside the net | UB_HYST_FRAME=0 = union-find).
TAG=${1:-cols3}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/${TAG}_gpu_tests.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/${TAG}_gpu_tests.log
timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2.json 2> gpurun_out/${TAG}_stage2.log; echo "stage2 exit $?"; tail -10 gpurun_out/${TAG}_stage2.log
for i in 1 2; do
  for V in "X=0" "UB_HYST_LATE=0" "UB_HYST_FRAME=0" "UB_HYST_FRAME=0 UB_BURR_COLS=0"; do
    N=$(echo $V | tr ' =' '__')
    env $V timeout 300 python bench.py --steps 40 --warmup 5 --no-hbm-kernels > gpurun_out/${TAG}_bench_${N}_$i.json 2> gpurun_out/${TAG}_bench_${N}_$i.err
    echo "$V run $i: $(python -c "import json,sys; d=json.loads(open('gpurun_out/${TAG}_bench_${N}_$i.json').read().strip().splitlines()[-1]); print(round(d['value']), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'])")"
  done
done
