#!/bin/bash
# k_preprocess check: parity suite (staged and unstaged taps), then its launches alone under ncu (duration + DRAM bytes).
TAG=${1:-pre}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/${TAG}_gpu_tests.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/${TAG}_gpu_tests.log
UB_PRE_STAGED=0 timeout 300 python -m pytest tests/test_gpu_stage2.py tests/test_gpu_net.py -m gpu -x -q --timeout 120 -k "preprocess or pipeline or module" > gpurun_out/${TAG}_gpu_tests_unstaged.log 2>&1
echo "pytest (UB_PRE_STAGED=0) exit $?"; tail -2 gpurun_out/${TAG}_gpu_tests_unstaged.log
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none \
    -k regex:k_preprocess -c 4 --csv --log-file gpurun_out/${TAG}_pre.csv python tools/profile_step.py 2 64 > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; grep -E "gpu__time|dram__|inst_executed" gpurun_out/${TAG}_pre.csv | awk -F'","' '{print $(NF-2), $NF}'
