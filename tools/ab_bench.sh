#!/bin/bash
# A/B of the engine switches inside whole sorts: every variant twice, interleaved.  (profiles/r02_ab_in_sort.txt was taken when
# the row interleave was still the default and CKKS_B200_NTT_NO_INTERLEAVE switched it off: its `nomix` line is today's default.)
cd "$GRAFT_REPO_ROOT"
run() { name=$1; shift; env "$@" python bench.py --no-cpu-baseline --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('$name', round(d['ms_per_step'],2), round(d['e2e']['value']*1e3,2), d['gpu_launches'], {k:round(v,1) for k,v in d['breakdown_ms'].items() if v})"; }
for rep in 1 2; do
  run default X=1
  run noarena CKKS_B200_ARENA_MB=0
  run oldmd CKKS_B200_OLD_MODDOWN=1
  run mix CKKS_B200_NTT_INTERLEAVE=1
  run oldmd_noarena CKKS_B200_OLD_MODDOWN=1 CKKS_B200_ARENA_MB=0
done
