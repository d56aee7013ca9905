#!/bin/bash
# Builds libgentun_b200.so in-tree for sm_100a (B200).  -fmad=false: the split arithmetic mirrors the
# reference's float/double expression order, so no FMA contraction.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="${GT_OUT:-$HERE/../libgentun_b200.so}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false \
  -ccbin /usr/bin/g++ -Xcompiler -fPIC -Xcompiler -O2 -shared ${GT_PTXAS_V:+-Xptxas -v} ${GT_DEFS} \
  -o "$OUT" "$HERE/sketch.cu" "$HERE/boost.cu" "$HERE/small.cu" "$HERE/engine.cu" 2>&1
echo "built $OUT"
