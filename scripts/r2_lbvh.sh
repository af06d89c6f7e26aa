#!/bin/bash
OUT=gpurun_out; TAG=${1:-r2v}; mkdir -p $OUT
for w in c4 c5; do
  timeout 120 python bench.py --workload $w --no-cpu --no-nested --steps 5 > $OUT/${TAG}_$w.json 2> $OUT/${TAG}_$w.err
  python - <<PY
import json
d=json.loads(open("$OUT/${TAG}_$w.json").read().strip().splitlines()[-1])
print("$w", round(d['ms_per_frame'],3), d['stage_ms_per_frame'], 'e2e', round(d['e2e']['ms_per_frame'],3), 'nodes/ray', round(d['roofline']['whole_frame']['nodes_per_ray'],2))
PY
done
