#!/bin/bash
# Builds a variant of the library with extra compiler flags into build/libunetb200_<name>.so (the shipped library is
# built by unet-_b200/csrc/Makefile).  Usage: tools/build_variant.sh bf16 -DUB_ACT_BF16
#   UNETB200_LIB=build/libunetb200_bf16.so python ...   selects it.
set -e
NAME=$1; shift
cd "$(dirname "$0")/../unet-_b200/csrc"
OUT=../../build/$NAME
mkdir -p $OUT
for f in api conv_tc conv_rows conv_block tmap conv_simt net_glue preprocess stage2 enhance annotate; do
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC \
    "$@" -c $f.cu -o $OUT/$f.o &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../build/libunetb200_$NAME.so $OUT/*.o -lcudart
echo built build/libunetb200_$NAME.so
