#!/bin/bash
# Column-walking stage-2 kernels: parity tests, per-operator bench A/B (new kernels on / off), pipeline bench A/B,
# one ncu --set full capture of the new kernels.   Usage: bash tools/gpu_cols.sh TAG [ncu]
TAG=${1:-cols}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stage2.py tests/test_gpu_configs.py tests/test_gpu_cli.py -m gpu -x -q --timeout 300 > gpurun_out/${TAG}_tests.log 2>&1
echo "pytest exit $?"; tail -12 gpurun_out/${TAG}_tests.log
timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2_new.json 2> gpurun_out/${TAG}_stage2_new.log; echo "new exit $?"; tail -10 gpurun_out/${TAG}_stage2_new.log
UB_BLUR_COLS=0 UB_MORPH_WORDS=1 UB_BURR_COLS=0 UB_EDGE_COLS=0 timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2_old.json 2> gpurun_out/${TAG}_stage2_old.log; echo "old exit $?"; tail -10 gpurun_out/${TAG}_stage2_old.log
timeout 300 python bench.py --steps 30 --warmup 5 --no-hbm-kernels > gpurun_out/${TAG}_bench_new.json 2> gpurun_out/${TAG}_bench_new.err; echo "bench new exit $?"; cut -c1-200 gpurun_out/${TAG}_bench_new.json
UB_BURR_COLS=0 UB_EDGE_COLS=0 timeout 300 python bench.py --steps 30 --warmup 5 --no-hbm-kernels > gpurun_out/${TAG}_bench_old.json 2> gpurun_out/${TAG}_bench_old.err; echo "bench old exit $?"; cut -c1-200 gpurun_out/${TAG}_bench_old.json
if [ "$2" = "ncu" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_blur5_cols|k_morph_cols|k_burr_chain_cols|k_edge_cols' -c 8 -o gpurun_out/${TAG}_cols \
      python tools/bench_stage2.py 512 1 > gpurun_out/${TAG}_ncu.log 2>&1
  echo "ncu exit $?"; tail -3 gpurun_out/${TAG}_ncu.log
  ncu -i gpurun_out/${TAG}_cols.ncu-rep --page raw --csv > gpurun_out/${TAG}_cols_raw.csv 2>/dev/null
fi
