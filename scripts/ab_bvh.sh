#!/bin/bash
# A/B of LBVH variants inside ONE box (boxes differ by +-5 %): ab/<name>/librtb200.so swapped in turn.
# usage: VARIANTS="base B" WL=c4 ROUNDS=2 scripts/ab_bvh.sh      (a variant "X+aniso" sets RTB200_BVH_ANISO=1)
for round in $(seq ${ROUNDS:-2}); do
for v in ${VARIANTS:-base B}; do
  lib=${v%%+*}
  cp ab/$lib/librtb200.so ray-tracing-gpu_b200/lib/librtb200.so
  unset RTB200_BVH_ANISO
  [[ $v == *+aniso ]] && export RTB200_BVH_ANISO=1
  python bench.py --workload ${WL:-c4} --steps ${STEPS:-10} --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
w=d['roofline'].get('whole_frame',{})
print('$v', round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['stage_ms_per_frame'].items()}, 'nodes/ray', round(w.get('nodes_per_ray',0),2), 'tests/ray', round(w.get('tri_tests_per_ray',0),2), flush=True)"
done
done
