#!/bin/bash
# Round-2 check of the current build: smoke, full GPU parity suite, default bench (new roofline leg / side workloads / batch
# sweep), reference arm, video I/O probe, MMA microbenchmarks, metrics-only ncu of the four conv launches the r02a capture missed.
TAG=${1:-r02b}
mkdir -p gpurun_out
timeout 400 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/${TAG}_smoke.log
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 -s > gpurun_out/${TAG}_gpu_tests.log 2>&1
echo "pytest exit $?"; tail -4 gpurun_out/${TAG}_gpu_tests.log; grep -E "status agreement|config [45] frame" gpurun_out/${TAG}_gpu_tests.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; cut -c1-600 gpurun_out/${TAG}_bench.json; tail -5 gpurun_out/${TAG}_bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref exit $?"; cut -c1-300 gpurun_out/${TAG}_bench_ref.json
timeout 300 python tools/video_io_probe.py 256 > gpurun_out/${TAG}_video_io.json 2> gpurun_out/${TAG}_video_io.err; echo "video probe exit $?"; tail -25 gpurun_out/${TAG}_video_io.json
timeout 120 ./build/ubench_mma > gpurun_out/${TAG}_ubench_mma.log 2>&1; timeout 120 ./build/ubench_mma2 > gpurun_out/${TAG}_ubench_mma2.log 2>&1; cat gpurun_out/${TAG}_ubench_mma2.log
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.avg.per_second
timeout 300 ncu --metrics $M --clock-control none -k regex:k_conv_rows -s 12 -c 4 --csv --log-file gpurun_out/${TAG}_rows_tail_metrics.csv \
    python tools/profile_step.py 2 64 > gpurun_out/${TAG}_ncu3.log 2>&1
echo "ncu metrics exit $?"; tail -3 gpurun_out/${TAG}_ncu3.log
