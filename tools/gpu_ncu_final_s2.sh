#!/bin/bash
# ncu --set full of the final stage-2 kernels (one process per kernel; tools/bench_stage2.py 512 1 as the workload).
TAG=${1:-ncus2}
mkdir -p gpurun_out
timeout 200 python tools/bench_stage2.py 512 1 > /dev/null 2>&1 || exit 1
for SPEC in "k_blur5_cols 1" "k_hyst_frame 3" "k_burr_chain_cols 1" "k_morph_cols 5" "k_gray 2" "k_exterior_frame 0" "k_contour_paint 0"; do
  set -- $SPEC; K=$1; SKIP=$2
  if [ "$K" = "k_exterior_frame" ] || [ "$K" = "k_contour_paint" ]; then
    CMD="python tools/time_contours.py"
  else
    CMD="python tools/bench_stage2.py 512 1"
  fi
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -o gpurun_out/${TAG}_$K $CMD > gpurun_out/${TAG}_ncu_$K.log 2>&1
  echo "ncu $K exit $?"
  ncu -i gpurun_out/${TAG}_$K.ncu-rep --page raw --csv > gpurun_out/${TAG}_${K}_raw.csv 2>/dev/null
done
