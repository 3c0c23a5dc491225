#!/bin/bash
# Timing-experiment builds of the library: tools/build_variant.sh NAME -DFLAG... -> rwkv_cppy_b200/build/librwkv_NAME.so
# (only gemm_tc.cu is recompiled with the flags; the other objects come from the regular build)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
B=rwkv_cppy_b200/build
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -DRWKV_BUILD -DRWKV_SHARED "$@" -x cu -c ${SRC:-rwkv_cppy_b200/csrc/gemm_tc.cu} -o $B/gemm_tc_$name.o
objs=$(ls $B/*.o | grep -v "gemm_tc")
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o $B/librwkv_$name.so $objs $B/gemm_tc_$name.o -lpthread -ldl -lrt
echo $B/librwkv_$name.so
