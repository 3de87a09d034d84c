#!/bin/bash
# Metrics-only ncu pass over the conv launches of two steps; the summary keeps the second step's (the capture bench.py's roofline.traffic comes from).
# Usage: bash tools/gpu_conv_metrics.sh TAG     then here: python tools/ncu_conv_summary.py gpurun_out/TAG_conv_metrics.csv TAG
TAG=${1:-r02}
mkdir -p gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.avg.per_second
timeout 120 python tools/profile_step.py 2 64 > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 400 ncu --metrics $M --clock-control none -k regex:k_conv -c 60 --csv --log-file gpurun_out/${TAG}_conv_metrics.csv \
    python tools/profile_step.py 2 64 > gpurun_out/${TAG}_ncu_metrics.log 2>&1
echo "ncu conv metrics exit $?"; tail -2 gpurun_out/${TAG}_ncu_metrics.log
