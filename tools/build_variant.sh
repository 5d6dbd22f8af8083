#!/bin/bash
# Builds raytracing_gpu_b200/librt_b200_<name>.so: the product library with its .cu files compiled under extra -D flags
# (same-box A/B of compile-time variants; select one with RT_B200_LIB=<path>).  usage: tools/build_variant.sh <name> -DX=1 ...
set -e
name=$1; shift
cd "$(dirname "$0")/../raytracing_gpu_b200/csrc"
make -s -j8 >/dev/null
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
mkdir -p build/$name
for f in device_api bvh_build wavefront multi_gpu; do
  $NVCC -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -fmad=false -Xptxas -v -Xcompiler -fPIC,-ffp-contract=off,-fvisibility=default \
      -cudart static "$@" -c $f.cu -o build/$name/$f.o 2> build/$name/$f.ptxas.log &
done
wait
objs=$(ls build/*.o | grep -v -E "device_api|bvh_build|wavefront|multi_gpu" | tr '\n' ' ')
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o ../librt_b200_$name.so $objs build/$name/*.o
echo "-> librt_b200_$name.so"
