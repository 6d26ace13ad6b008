#!/bin/sh
# nn_bvh_b200/csrc/build.sh -- compile libnnbvh.so in-tree for sm_100a.
# The host builder decides tree shape with float arithmetic: -ffp-contract=off.
# Device code: --fmad=false (plus explicit _rn intrinsics in exact_math.cuh).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="${NNBVH_OUT:-$HERE/../libnnbvh.so}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -std=c++17 -O3 -lineinfo --fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
    -gencode arch=compute_100a,code=sm_100a \
    -Xcompiler -fPIC,-ffp-contract=off,-O2,-pthread -shared \
    ${NNBVH_PTXAS_V:+-Xptxas -v} ${NNBVH_EXTRA_FLAGS} \
    -o "$OUT" "$HERE/engine.cu" "$HERE/build_host.cpp" -lcudart
echo "built $OUT"
