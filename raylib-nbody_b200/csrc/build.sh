#!/bin/sh
# Builds libnbody_b200.so (sm_100a only) next to this script's parent package directory.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../libnbody_b200.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
    -Xcompiler -fPIC -Xcompiler -ffp-contract=off -shared \
    -o "$OUT" "$HERE/nbody_b200.cu" -ldl "$@"
echo "built $OUT"
