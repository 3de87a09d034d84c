#!/bin/bash
# Debug build of the library with the conv_rows pipeline trace compiled in (-DUB_ROWS_TRACE):
#   tools/build_trace_lib.sh && UNETB200_LIB=build/libunetb200_trace.so UB_ROWS_TRACE_DIR=gpurun_out python tools/profile_step.py 1 64
set -e
cd "$(dirname "$0")/../unet-_b200/csrc"
mkdir -p ../../build/trace
for f in api conv_tc conv_rows conv_block tmap conv_simt net_glue preprocess stage2 enhance; do
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC \
    -DUB_ROWS_TRACE -c $f.cu -o ../../build/trace/$f.o &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../build/libunetb200_trace.so ../../build/trace/*.o -lcudart
echo built build/libunetb200_trace.so
