#!/bin/bash
# Profile capture for profiles/ (run under gpurun on ONE GPU, after the plain commands have exited 0).
# usage: scripts/gpu_profile.sh <tag>
TAG=${1:-r1}
OUT=gpurun_out
set -x
python bench.py --workload c2 --steps 2 --warmup 1 --no-cpu > $OUT/${TAG}_c2_plain.json 2> $OUT/${TAG}_c2_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_c2_launches.csv \
    python bench.py --workload c2 --steps 2 --warmup 1 --no-cpu > $OUT/${TAG}_c2_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_resident_level -s 3 -c 1 -f -o $OUT/${TAG}_c2_resident \
    python bench.py --workload c2 --steps 2 --warmup 1 --no-cpu > $OUT/${TAG}_c2_full.log 2>&1
python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu > $OUT/${TAG}_c3_plain.json 2> $OUT/${TAG}_c3_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/${TAG}_c3_launches.csv \
    python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu > $OUT/${TAG}_c3_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_trace_nearest -c 1 -f -o $OUT/${TAG}_c3_nearest \
    python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu > $OUT/${TAG}_c3_full1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_trace_shadow -c 1 -f -o $OUT/${TAG}_c3_shadow \
    python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu > $OUT/${TAG}_c3_full2.log 2>&1
python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu > $OUT/${TAG}_c4_plain.json 2> $OUT/${TAG}_c4_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/${TAG}_c4_launches.csv \
    python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu > $OUT/${TAG}_c4_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_bvh_shadow -c 1 -f -o $OUT/${TAG}_c4_bvh_shadow \
    python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu > $OUT/${TAG}_c4_full1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_bvh_nearest -c 1 -f -o $OUT/${TAG}_c4_bvh_nearest \
    python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu > $OUT/${TAG}_c4_full2.log 2>&1
ls -la $OUT
