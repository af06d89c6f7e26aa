#!/bin/bash
# FIRST GPU call of round 2 (one GPU, ~4 min): what round 1 could not measure any more.
#   1. the whole GPU suite on the final tree of round 1 (queue order of the tiled path on by default)
#   2. the A/B of the queue order and of the prepared, OFF-by-default variants
#        RTB200_NEAREST_CONES=4   (k_trace_nearest MULTI: up to 4 cones per warp of secondary rays)
#        RTB200_BVH_TILE_RAYS=1   (16x16 patches on the LBVH path)
#      frames must be bit-identical in every column
#   3. fresh bench lines of the four configs (the c3 line in profiles/ predates the queue order)
#   4. the shadow kernel of config 3 under ncu with the new order
# usage: gpurun --timeout 600 -- 'bash scripts/round2_first.sh r2a'
OUT=gpurun_out; TAG=${1:-r2a}
mkdir -p $OUT
set -x
timeout 120 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > $OUT/${TAG}_tests.log
timeout 90 python scripts/ab_grouping.py > $OUT/${TAG}_ab_grouping.log 2>&1
for w in c4 c5; do
  for v in 0 1; do
    if [ $v = 1 ]; then export RTB200_BVH_TILE_RAYS=1; else unset RTB200_BVH_TILE_RAYS; fi
    timeout 60 python bench.py --workload $w --no-cpu --steps 5 --warmup 3 > $OUT/${TAG}_${w}_tile${v}.json 2> $OUT/${TAG}_${w}_tile${v}.err
  done
done
unset RTB200_BVH_TILE_RAYS
timeout 100 python bench.py > $OUT/${TAG}_c2.json 2> $OUT/${TAG}_c2.err
timeout 100 python bench.py --workload c3 > $OUT/${TAG}_c3.json 2> $OUT/${TAG}_c3.err
RTB200_NEAREST_CONES=4 timeout 60 python bench.py --workload c3 --no-cpu > $OUT/${TAG}_c3_cones4.json 2> $OUT/${TAG}_c3_cones4.err
timeout 60 python bench.py --workload c4 > $OUT/${TAG}_c4.json 2> $OUT/${TAG}_c4.err
timeout 90 python bench.py --workload c5 > $OUT/${TAG}_c5.json 2> $OUT/${TAG}_c5.err
timeout 120 ncu --set full --clock-control none --import-source on -k regex:k_trace_shadow -c 1 -f -o $OUT/${TAG}_c3_shadow \
    python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu > $OUT/${TAG}_c3_ncu_shadow.log 2>&1
timeout 120 ncu --set full --clock-control none --import-source on -k regex:k_trace_nearest -s 1 -c 1 -f -o $OUT/${TAG}_c3_nearest_l1 \
    python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu > $OUT/${TAG}_c3_ncu_nearest.log 2>&1
ls -la $OUT
