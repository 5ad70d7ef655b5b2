#!/bin/bash
# Development aid: builds a kernel variant of the library into variants/librt_<name>.so (git-ignored, travels with gpurun).
# usage: scripts/build_variant.sh <name> [-DFLAG ...]        run with RT_B200_LIB=variants/librt_<name>.so
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p variants/obj_$name
for f in rt_api coupling_kernels lu_kernels controller_kernels misc_kernels fused_kernels noalloc_api; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC "$@" \
    -c radiative_transfer_b200/csrc/$f.cu -o variants/obj_$name/$f.o &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o variants/librt_$name.so variants/obj_$name/*.o -lcudart
echo variants/librt_$name.so
