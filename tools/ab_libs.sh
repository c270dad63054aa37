#!/bin/bash
# A/B of engine BUILDS inside whole sorts and the key-switch microbenchmark, each variant twice, interleaved.  The sorter layer
# (libsortfhe_host.so) binds libckks_b200.so by rpath, so CKKS_B200_LIB alone would leave the sorts on the stock build: the
# variant is copied over sorting-fhe_b200/libckks_b200.so for its runs (run this on the GPU box copy, not in the source tree).
#   tools/ab_libs.sh libckks_b200.so libckks_b200_<variant>.so ...
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
D=sorting-fhe_b200
mkdir -p /tmp/ab_libs
for v in "$@"; do cp $D/$v /tmp/ab_libs/$v; done
for rep in 1 2; do
  for v in "$@"; do
    cp /tmp/ab_libs/$v $D/libckks_b200.so
    python bench.py --no-cpu-baseline --steps 10 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('$v', round(d['ms_per_step'],2), round(d['e2e']['value']*1e3,2), d['gpu_launches'], {k:round(x,2) for k,x in d['breakdown_ms'].items() if x})"
    python tools/microbench.py 17 30 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin); print('   ks_us', [round(d[k]['ks_us'],1) for k in ('nl32','nl24','nl16','nl8')], 'ntt_us', [round(d[k]['ntt_us'],1) for k in ('nl32','nl8')])"
  done
done
cp /tmp/ab_libs/$1 $D/libckks_b200.so
