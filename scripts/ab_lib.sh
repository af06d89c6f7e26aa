#!/bin/bash
# A/B of library variants inside ONE box: ab/<name>/librtb200.so swapped in turn (scripts/build_variant.sh).
# usage: VARIANTS="base t128c7" WL=c2 ROUNDS=2 scripts/ab_lib.sh
cp ray-tracing-gpu_b200/lib/librtb200.so /tmp/librtb200.keep
for round in $(seq ${ROUNDS:-2}); do
for v in ${VARIANTS:-base}; do
  cp ab/$v/librtb200.so ray-tracing-gpu_b200/lib/librtb200.so
  python bench.py --workload ${WL:-c2} --steps ${STEPS:-50} --warmup 5 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v', 'ms/frame', round(d['ms_per_step'],5), 'kernel', round(d['roofline']['kernel_ms'],5), 'sync-calls', round(d.get('ms_per_frame_synchronous_calls',0),5), 'e2e', round(d['e2e']['ms_per_step'],4), flush=True)"
done
done
cp /tmp/librtb200.keep ray-tracing-gpu_b200/lib/librtb200.so
