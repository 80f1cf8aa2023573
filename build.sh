#!/bin/bash
# Build liboevae.so (C ABI + sm_100a kernels) in-tree.  nvcc cross-compiles without a GPU.
set -e
cd "$(dirname "$0")"
OUT=atlas_vae_b200/liboevae.so
SRC="atlas_vae_b200/csrc/oevae_abi.cu"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
  -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -shared -cudart static \
  ${PTXAS_V:+-Xptxas -v} -o $OUT $SRC -lcuda
echo "built $OUT"
