#!/bin/bash
# developer: link a probe variant of the library with one source recompiled under extra defines
#   tools/build_probe.sh <tag> <source.cu> -DGB_PROF -DGB_PROBE=2   ->  iaas_sglang_b200/csrc/libb200sgl_<tag>.so
set -e
cd "$(dirname "$0")/../iaas_sglang_b200/csrc"
tag=$1; src=$2; shift 2
obj=build/probe-$tag-${src%.cu}.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -c $src -o $obj
objs=$(ls build/*.o | grep -v "build/probe-" | grep -v "build/${src%.cu}.o")
/usr/local/cuda/bin/nvcc -shared -o libb200sgl_$tag.so $objs $obj -gencode arch=compute_100a,code=sm_100a -cudart shared
echo libb200sgl_$tag.so
