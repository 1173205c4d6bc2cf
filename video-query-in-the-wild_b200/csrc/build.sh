#!/bin/bash
# Builds libvqwild.so (sm_100a only) next to the Python package.  Usage: csrc/build.sh [extra nvcc flags]
set -e
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/../libvqwild.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
  -Xcompiler -fPIC,-fvisibility=hidden -shared "$@" \
  "$HERE/vq_api.cu" "$HERE/vq_gemm_topk.cu" "$HERE/vq_select.cu" "$HERE/vq_exact.cu" \
  "$HERE/vq_prepare.cu" "$HERE/vq_metrics.cu" "$HERE/vq_triplet.cu" "$HERE/vq_triplet_tc.cu" \
  -o "$OUT" -lcudart_static -ldl -lrt -lpthread
echo "built $OUT"
