#!/bin/bash
# pipeline traces of the conv_rows launches of one step (needs tools/build_trace_lib.sh run before gpurun)
mkdir -p gpurun_out
UNETB200_LIB=$PWD/build/libunetb200_trace.so UB_ROWS_TRACE_DIR=gpurun_out timeout 300 python tools/profile_step.py 1 64 > gpurun_out/trace_run.log 2>&1
echo "trace exit $?"; tail -2 gpurun_out/trace_run.log
python tools/rows_trace.py gpurun_out/rows_trace_*.bin > gpurun_out/rows_trace.txt 2>&1
cat gpurun_out/rows_trace.txt
