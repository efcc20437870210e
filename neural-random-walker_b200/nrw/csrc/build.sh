#!/bin/bash
# Build libnrw.so for sm_100a (B200) in-tree.  nvcc cross-compiles without a GPU.
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
      -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -shared \
      ${NRW_NVCC_EXTRA:-} \
      -o ${NRW_LIB_OUT:-libnrw.so} nrw_capi.cu -lcudart
echo "built $(pwd)/${NRW_LIB_OUT:-libnrw.so}"
