#!/bin/bash
# A/B builds of the engine: recompiles the named sources with extra flags and links them with the stock objects into
# sorting-fhe_b200/libckks_b200_<name>.so (select with CKKS_B200_LIB=<path>).
#   tools/build_variant.sh tb11 "-DNTT_TB_MAX=11" ntt
set -e
cd "$(dirname "$0")/../sorting-fhe_b200/csrc"
NAME=$1; EXTRA=$2; shift 2
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fno-strict-aliasing --fmad=false"
OBJS=""
for f in ntt kernels keyswitch encode context evaluator cheby bootstrap capi; do
  o=build/$f.o
  for v in "$@"; do if [ "$v" = "$f" ]; then o=build/${f}_$NAME.o; $NVCC $FLAGS $EXTRA -c $f.cu -o $o & fi; done
  OBJS="$OBJS $o"
done
wait
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o ../libckks_b200_$NAME.so $OBJS -lcudart
echo "built libckks_b200_$NAME.so"
