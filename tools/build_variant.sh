#!/bin/bash
# usage: tools/build_variant.sh NAME "-DFLAG ..."  -> mxmoe_b200/lib/libmxmoe_b200_NAME.so (load it with MXB_LIB=...)
# Another build of the SAME sources with extra -D flags, for A/B runs of a kernel variant on one gpurun call.
set -e
cd "$(dirname "$0")/../mxmoe_b200/csrc"
name=$1; shift
b=../../build/var_$name
mkdir -p $b ../lib
for f in gg_kernel gg_pair_kernel moe_ops hadamard w4a16_stream api; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC "$@" -c $f.cu -o $b/$f.o &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../lib/libmxmoe_b200_$name.so $b/*.o -lcudart
echo built mxmoe_b200/lib/libmxmoe_b200_$name.so
