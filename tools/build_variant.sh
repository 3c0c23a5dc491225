#!/bin/bash
# Timing-experiment builds of the library:
#   tools/build_variant.sh NAME SOURCE.cu [-DFLAG ...]  ->  rwkv_cppy_b200/build/variants/librwkv_NAME.so
# Only SOURCE (a file of rwkv_cppy_b200/csrc, or a path elsewhere, e.g. an older revision exported to /tmp) is recompiled
# with the flags; every other object comes from the regular build (python -m rwkv_cppy_b200.build first).
set -e
cd "$(dirname "$0")/.."
name=$1; src=$2; shift 2
B=rwkv_cppy_b200/build
mkdir -p $B/variants
base=$(basename "$src")
[ -f "$src" ] || src=rwkv_cppy_b200/csrc/$base
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden \
    -DRWKV_BUILD -DRWKV_SHARED "$@" -x cu -c "$src" -o $B/variants/${base}_$name.o
objs=$(ls $B/*.o | grep -v "/$base.o")
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o $B/variants/librwkv_$name.so $objs $B/variants/${base}_$name.o -lpthread -ldl -lrt
echo $B/variants/librwkv_$name.so
