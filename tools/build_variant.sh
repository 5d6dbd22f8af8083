#!/bin/bash
# Builds raytracing_gpu_b200/librt_b200_<name>.so: the product library with wavefront.cu compiled under extra -D flags
# (same-box A/B of compile-time variants; select one with RT_B200_LIB=<path>).  usage: tools/build_variant.sh <name> -DX=1 ...
set -e
name=$1; shift
cd "$(dirname "$0")/../raytracing_gpu_b200/csrc"
make -s -j8 >/dev/null
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -fmad=false -Xptxas -v -Xcompiler -fPIC,-ffp-contract=off,-fvisibility=default \
    -cudart static "$@" -c wavefront.cu -o build/wavefront_$name.o 2> build/wavefront_$name.ptxas.log
objs=$(ls build/*.o | grep -v "wavefront" | tr '\n' ' ')
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o ../librt_b200_$name.so $objs build/wavefront_$name.o
grep -A2 "7k_shadeE" build/wavefront_$name.ptxas.log | grep -E "spill|Used" | tr '\n' ' '; echo " -> librt_b200_$name.so"
