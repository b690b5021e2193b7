#!/bin/bash
# Build the C-ABI shared library for sm_100a (cross-compiles without a GPU).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="$HERE/pytorch-kaldi-cgs_b200/csrc"
OUT="$HERE/pytorch-kaldi-cgs_b200/libhcgs_b200.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo \
    -Xcompiler -fPIC --shared \
    -DHCGS_BUILD -o "$OUT" "$SRC"/*.cu -lcudart ${HCGS_NVCC_EXTRA}
echo "built $OUT"
