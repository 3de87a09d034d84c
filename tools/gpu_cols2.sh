#!/bin/bash
# Column-walking / flood stage-2 kernels, second visit: parity tests, per-operator bench (new default, then with the
# listed env overrides), pipeline bench, ncu --set full of one launch of each new kernel.
# Usage: bash tools/gpu_cols2.sh TAG "ENV=.. ENV=.." [ncu]
TAG=${1:-cols2}
OLD=${2:-UB_HYST_FRAME=0 UB_MORPH_TABLES=0}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stage2.py tests/test_gpu_configs.py tests/test_gpu_cli.py -m gpu -x -q --timeout 300 > gpurun_out/${TAG}_tests.log 2>&1
echo "pytest exit $?"; tail -12 gpurun_out/${TAG}_tests.log
timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2_new.json 2> gpurun_out/${TAG}_stage2_new.log; echo "new exit $?"; tail -10 gpurun_out/${TAG}_stage2_new.log
env $OLD timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2_old.json 2> gpurun_out/${TAG}_stage2_old.log; echo "old ($OLD) exit $?"; tail -10 gpurun_out/${TAG}_stage2_old.log
timeout 300 python bench.py --steps 30 --warmup 5 --no-hbm-kernels > gpurun_out/${TAG}_bench_new.json 2> gpurun_out/${TAG}_bench_new.err; echo "bench new exit $?"; cut -c1-200 gpurun_out/${TAG}_bench_new.json
if [ "$3" = "ncu" ]; then
  for K in k_hyst_frame k_morph_cols k_burr_chain_cols k_edge_fused; do
    timeout 300 ncu --set full --clock-control none --import-source on -k regex:$K -s 4 -c 2 -o gpurun_out/${TAG}_$K \
        python tools/bench_stage2.py 512 1 > gpurun_out/${TAG}_ncu_$K.log 2>&1
    echo "ncu $K exit $?"
    ncu -i gpurun_out/${TAG}_$K.ncu-rep --page raw --csv > gpurun_out/${TAG}_${K}_raw.csv 2>/dev/null
  done
fi
