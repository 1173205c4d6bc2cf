#!/bin/bash
# Builds libvqwild.so (sm_100a only) next to the Python package.  Usage: csrc/build.sh [extra nvcc flags]
# Translation units are compiled in parallel; an object is rebuilt when its source or any header is newer.
set -e
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/../libvqwild.so"
OBJ="$HERE/_obj"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden"
SRCS="vq_api vq_gemm_topk vq_select vq_sort vq_exact vq_prepare vq_metrics vq_triplet vq_triplet_tc vq_moment vq_windows vq_losses vq_noalloc"
mkdir -p "$OBJ"
echo "$FLAGS $*" > "$OBJ/.flags.new"
if ! cmp -s "$OBJ/.flags.new" "$OBJ/.flags"; then rm -f "$OBJ"/*.o; mv "$OBJ/.flags.new" "$OBJ/.flags"; fi
newest_hdr=$(ls -t "$HERE"/*.cuh "$HERE"/*.h "$HERE"/*.inl "$HERE/../../include/vqwild.h" | head -1)
pids=()
for s in $SRCS; do
  [ -f "$HERE/$s.cu" ] || continue
  o="$OBJ/$s.o"
  if [ ! -f "$o" ] || [ "$HERE/$s.cu" -nt "$o" ] || [ "$newest_hdr" -nt "$o" ]; then
    "$NVCC" $FLAGS "$@" -c "$HERE/$s.cu" -o "$o" &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait "$p"; done
objs=""
for s in $SRCS; do [ -f "$OBJ/$s.o" ] && objs="$objs $OBJ/$s.o"; done
"$NVCC" -gencode arch=compute_100a,code=sm_100a -shared $objs -o "$OUT" -lcudart_static -ldl -lrt -lpthread
echo "built $OUT"
