#!/bin/bash
# evidence refresh after a change to the culled scan only: the default bench line, the reference arm, ncu captures of the
# four culled-scan kernels and the launch list of config 3 (the other captures of scripts/r2_profile.sh stay valid)
OUT=gpurun_out; TAG=${1:-r2f}
mkdir -p $OUT
( time timeout 600 python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench_default.json 2> $OUT/${TAG}_bench_default.err ) 2> $OUT/${TAG}_bench_default.time
( time timeout 200 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err ) 2> $OUT/${TAG}_bench_reference.time
cap() {  # tag workload kernel-regex skip
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:$3 -s $4 -c 1 -f -o $OUT/${TAG}_$1 \
      python bench.py --workload $2 --steps 1 --warmup 1 --no-cpu --no-nested > $OUT/${TAG}_$1.log 2>&1
}
cap c3_cull_shadow c3 k_cull_shadow 0
cap c3_cull_nearest_l1 c3 k_cull_nearest 1
cap c3_lane_nearest_l1 c3 k_lane_nearest 0
cap c3_lane_shadow_l0 c3 k_lane_shadow 0
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $OUT/${TAG}_c3_launches.csv \
    python bench.py --workload c3 --steps 2 --warmup 1 --no-cpu --no-nested > $OUT/${TAG}_c3_launches.log 2>&1
mkdir -p $OUT/extracted_$TAG
for n in c3_cull_shadow c3_cull_nearest_l1 c3_lane_nearest_l1 c3_lane_shadow_l0; do
  python scripts/ncu_summary.py $OUT/${TAG}_$n.ncu-rep $OUT/extracted_$TAG/r2_${n}_ncu_full.txt > /dev/null 2>&1
done
for n in c3_cull_nearest_l1 c3_lane_nearest_l1; do rm -f $OUT/${TAG}_$n.ncu-rep; done
ls -la $OUT | tail -16
