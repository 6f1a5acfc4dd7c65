#!/bin/bash
# Builds grace_b200/lib/libgrace_b200.so for sm_100a (cross-compiles without a GPU).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../lib"
# GB_DEBUG=1: the checking build (device-side bounds asserts -> GB_STATUS_ASSERT), kept next to the release one;
# select it at run time with GRACE_B200_LIB=grace_b200/lib/debug/libgrace_b200.so
if [ -n "$GB_DEBUG" ]; then OUT="$OUT/debug"; GB_NVCC_EXTRA="$GB_NVCC_EXTRA -DGB_DEBUG"; fi
mkdir -p "$OUT"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -Wall -Xcompiler -Wno-unknown-pragmas ${GB_NVCC_EXTRA}"
SRCS="engine.cu plan.cu ops.cu k1_sparse.cu k1_topk2.cu k1_adaq.cu k2_quant.cu k2_sketch.cu k2_inceptionn.cu"
[ -f "$HERE/k0_dense.cu" ] && SRCS="$SRCS k0_dense.cu"
[ -f "$HERE/k3_powersgd.cu" ] && SRCS="$SRCS k3_powersgd.cu"
[ -f "$HERE/k1_static.cu" ] && SRCS="$SRCS k1_static.cu"
pids=""
for s in $SRCS; do
  "$NVCC" $FLAGS -c "$HERE/$s" -o "$OUT/${s%.cu}.o" &
  pids="$pids $!"
done
for p in $pids; do wait $p; done
OBJS=""
for s in $SRCS; do OBJS="$OBJS $OUT/${s%.cu}.o"; done
"$NVCC" -shared -gencode arch=compute_100a,code=sm_100a $OBJS -o "$OUT/libgrace_b200.so"
echo "built $OUT/libgrace_b200.so"
