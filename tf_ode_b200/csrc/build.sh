#!/bin/bash
# Builds the C-ABI shared library for sm_100a, in-tree (the .so travels to the GPU box with the snapshot).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
# -fmad=false: the kernels place every fused multiply-add explicitly (include/tfode_numerics.h), nvcc must not add its own.
$NVCC -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false \
  -Xcompiler -fPIC -Xcompiler -ffp-contract=off -shared -cudart static ${TFODE_NVCC_EXTRA} \
  -o libtfode_b200.so tfode_capi.cu -lz
echo "built $(pwd)/libtfode_b200.so"
