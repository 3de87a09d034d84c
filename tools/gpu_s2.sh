#!/bin/bash
# Stage-2 visit: stage-2 parity tests, the per-operator stage-2 bench with and without the fused edge kernel,
# optionally a full ncu capture of the fused kernel.   Usage: bash tools/gpu_s2.sh TAG [ncu]
TAG=${1:-s2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_stage2.py tests/test_gpu_cli.py -m gpu -x -q --timeout 120 > gpurun_out/${TAG}_tests.log 2>&1
echo "pytest exit $?"; tail -12 gpurun_out/${TAG}_tests.log
timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2_fused.json 2> gpurun_out/${TAG}_stage2_fused.log; echo "fused exit $?"; cat gpurun_out/${TAG}_stage2_fused.log | tail -8
UB_EDGE_FUSED=0 timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2_unfused.json 2> gpurun_out/${TAG}_stage2_unfused.log; echo "unfused exit $?"; tail -8 gpurun_out/${TAG}_stage2_unfused.log
if [ "$2" = "ncu" ]; then
  ncu --set full --clock-control none --import-source on -k regex:k_edge_fused -s 4 -c 3 -o gpurun_out/${TAG}_edge \
      python tools/bench_stage2.py 512 2 > gpurun_out/${TAG}_ncu.log 2>&1
  echo "ncu exit $?"; tail -3 gpurun_out/${TAG}_ncu.log
fi
