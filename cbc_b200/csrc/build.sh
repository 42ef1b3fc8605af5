#!/bin/bash
# Builds cbc_b200/lib/libcbp.so for sm_100a (nvcc cross-compiles without a GPU).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../lib"
mkdir -p "$OUT"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
# the environment may export CC/CXX pointing at a wrapper; pin the host compiler
"$NVCC" -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
  -fmad=false -Xcompiler -fPIC,-fvisibility=hidden,-O2 -shared \
  ${CBP_PTXAS_V:+-Xptxas -v} ${CBP_DEFS} \
  -o "${CBP_OUT:-$OUT/libcbp.so}" "$HERE/cbp_engine.cu" "$HERE/cbp_noalloc.cu" -lcudart
echo "built $OUT/libcbp.so"
