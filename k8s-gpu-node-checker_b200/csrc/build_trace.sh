#!/bin/bash
# Debug build with in-kernel phase timestamps (tools/trace_phases.py); not used by the product.
set -euo pipefail
cd "$(dirname "$0")"
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -DGNC_TRACE \
  -Xcompiler -fPIC,-Wall,-Wno-unused-function -shared -o libgnc_trace.so gnc.cu -ldl
