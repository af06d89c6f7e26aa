#!/bin/bash
# round 2: the default bench line (all blocks, CPU legs, parity) + ncu captures of the dominant kernels of every config
OUT=gpurun_out; TAG=${1:-r2p}
mkdir -p $OUT
( time timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench_default.json 2> $OUT/${TAG}_bench_default.err ) 2> $OUT/${TAG}_bench_default.time
( time timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err ) 2> $OUT/${TAG}_bench_reference.time
cap() {  # tag workload kernel-regex skip
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$3 -s $4 -c 1 -f -o $OUT/${TAG}_$1 \
      python bench.py --workload $2 --steps 1 --warmup 1 --no-cpu --no-nested > $OUT/${TAG}_$1.log 2>&1
}
cap c3_plain_nearest c3 k_trace_nearest 0
cap c3_cull_shadow c3 k_cull_shadow 0
cap c3_cull_nearest_l1 c3 k_cull_nearest 1        # launches per frame: <primary>, <level 1>, <level 2>, ...
cap c3_lane_nearest_l1 c3 k_lane_nearest 0
cap c3_lane_shadow_l0 c3 k_lane_shadow 0
cap c4_bvh_shadow c4 k_bvh_shadow 0
cap c4_bvh_nearest c4 k_bvh_nearest 0
cap c5_bvh_shadow c5 k_bvh_shadow 0
cap c2_resident c2 k_resident_level 0
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $OUT/${TAG}_c3_launches.csv \
    python bench.py --workload c3 --steps 2 --warmup 1 --no-cpu --no-nested > $OUT/${TAG}_c3_launches.log 2>&1
# ten .ncu-rep files are more than a call may bring back (64 MiB): summarise them here, keep two for source-level budgets
python scripts/ncu_extract.py $OUT $TAG $OUT/extracted > $OUT/${TAG}_extract.log 2>&1
for n in c3_plain_nearest c3_cull_nearest_l1 c3_lane_nearest_l1 c3_lane_shadow_l0 c4_bvh_shadow c4_bvh_nearest c2_resident; do rm -f $OUT/${TAG}_$n.ncu-rep; done
ls -la $OUT | tail -30
