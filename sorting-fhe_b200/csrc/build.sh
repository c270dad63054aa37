#!/bin/bash
# Build libckks_b200.so for sm_100a (in-tree; the .so travels to the GPU box, it is git-ignored).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
OUT=../libckks_b200.so
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fno-strict-aliasing --fmad=false"
OBJS=""
PIDS=""
for f in ntt kernels keyswitch encode context evaluator cheby bootstrap capi; do
  if [ ! -f build/$f.o ] || [ $f.cu -nt build/$f.o ] || [ engine.h -nt build/$f.o ] || [ modarith.cuh -nt build/$f.o ] || [ evaluator.h -nt build/$f.o ] || [ ../../include/ckks_b200.h -nt build/$f.o ]; then
    mkdir -p build
    rm -f build/$f.o
    $NVCC $FLAGS $EXTRA -c $f.cu -o build/$f.o &
    PIDS="$PIDS $!"
  fi
  OBJS="$OBJS build/$f.o"
done
for p in $PIDS; do wait $p || { echo "compile failed" >&2; exit 1; }; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o $OUT $OBJS -lcudart
echo "built $OUT"
