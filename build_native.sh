#!/bin/bash
# Build the C-ABI shared library for sm_100a (cross-compiles without a GPU).
# Every csrc/*.cu is compiled to build/*.o in parallel and only when it (or a header) changed; then one link.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="$HERE/pytorch-kaldi-cgs_b200/csrc"
OUT="$HERE/pytorch-kaldi-cgs_b200/libhcgs_b200.so"
OBJ="$HERE/build/obj"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -DHCGS_BUILD ${HCGS_NVCC_EXTRA}"
mkdir -p "$OBJ"
# a changed header or flag set rebuilds everything
STAMP="$OBJ/.flags"
NEWEST_HDR=$(ls -t "$SRC"/*.cuh "$HERE"/include/*.h 2>/dev/null | head -1)
if [ ! -f "$STAMP" ] || [ "$(cat "$STAMP")" != "$FLAGS" ] || [ "$NEWEST_HDR" -nt "$STAMP" ]; then
    rm -f "$OBJ"/*.o
    echo "$FLAGS" > "$STAMP"
fi
pids=()
for f in "$SRC"/*.cu; do
    o="$OBJ/$(basename "${f%.cu}").o"
    if [ ! -f "$o" ] || [ "$f" -nt "$o" ]; then
        ( "$NVCC" $FLAGS -c "$f" -o "$o.tmp" && mv "$o.tmp" "$o" ) &
        pids+=($!)
    fi
done
rc=0
for p in "${pids[@]}"; do wait "$p" || rc=1; done
[ $rc -eq 0 ] || { echo "nvcc failed" >&2; exit 1; }
"$NVCC" -gencode arch=compute_100a,code=sm_100a --shared -Xcompiler -fPIC -o "$OUT" "$OBJ"/*.o -lcudart
echo "built $OUT"
