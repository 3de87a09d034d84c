#!/bin/bash
# k_nlm rows per walk: ncu --set full at RT = 32 and at the default (20) -- the stall reasons tell why 32 ran at half speed --
# then the full GPU suite and the default bench line (its extra.workloads.enhanced leg is BASELINE config 4).
TAG=${1:-nlm}
mkdir -p gpurun_out
for R in 32 20; do
  UB_NLM_RT=$R timeout 300 ncu --set full --clock-control none -k regex:k_nlm -s 2 -c 2 -o gpurun_out/${TAG}_nlm_rt$R python tools/time_enhance.py 16 > gpurun_out/${TAG}_ncu_rt$R.log 2>&1
  echo "ncu RT=$R exit $?"
  ncu -i gpurun_out/${TAG}_nlm_rt$R.ncu-rep --page raw --csv > gpurun_out/${TAG}_nlm_rt${R}_raw.csv 2>/dev/null
done
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/${TAG}_gpu_tests.log 2>&1
echo "pytest exit $?"; tail -2 gpurun_out/${TAG}_gpu_tests.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; cut -c1-200 gpurun_out/${TAG}_bench.json
