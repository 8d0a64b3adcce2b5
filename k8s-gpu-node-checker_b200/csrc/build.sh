#!/bin/bash
# Build libgnc.so for sm_100a in-tree (nvcc cross-compiles without a GPU).
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
  -Xcompiler -fPIC,-Wall,-Wno-unused-function -Xptxas -v -shared \
  -o libgnc.so gnc.cu -ldl 2>&1 | tee build.log
