#!/bin/bash
# Last measurement call of round 1 (ONE GPU, under gpurun, short budget): steps ordered by priority, every
# step under its own timeout and writing its own file, so a call that is cut off keeps what finished.
#   1. the whole GPU test-suite + smoke() on the final build
#   2. bench lines of the LBVH workloads (their kernels changed after profiles/r1_c4|c5_bench.json)
#   3. ncu launch list + --set full captures of the LBVH traversal kernels (c4), shadow kernel of c5
#   4. headline c2 line + reference arm
# usage: scripts/final_round2.sh <tag>
TAG=${1:-r1g}
OUT=gpurun_out
mkdir -p $OUT
set -x
date +%s > $OUT/${TAG}_t0
timeout 420 python -m pytest tests -m gpu -x -q --durations=8 > $OUT/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> $OUT/${TAG}_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> $OUT/${TAG}_smoke.log
timeout 200 python bench.py --workload c4 > $OUT/${TAG}_c4.json 2> $OUT/${TAG}_c4.err
timeout 200 python bench.py --workload c5 > $OUT/${TAG}_c5.json 2> $OUT/${TAG}_c5.err
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/${TAG}_c4_launches.csv \
    python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu > $OUT/${TAG}_c4_ncu.log 2>&1
timeout 150 ncu --set full --clock-control none --import-source on -k regex:k_bvh_shadow -c 1 -f -o $OUT/${TAG}_c4_bvh_shadow \
    python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu > $OUT/${TAG}_c4_full1.log 2>&1
timeout 150 ncu --set full --clock-control none --import-source on -k regex:k_bvh_nearest -c 1 -f -o $OUT/${TAG}_c4_bvh_nearest \
    python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu > $OUT/${TAG}_c4_full2.log 2>&1
timeout 200 python bench.py > $OUT/${TAG}_c2.json 2> $OUT/${TAG}_c2.err
timeout 120 python bench.py --impl reference > $OUT/${TAG}_c2_reference.json 2> $OUT/${TAG}_c2_reference.err
timeout 200 ncu --set full --clock-control none --import-source on -k regex:k_bvh_shadow -c 1 -f -o $OUT/${TAG}_c5_bvh_shadow \
    python bench.py --workload c5 --steps 1 --warmup 1 --no-cpu > $OUT/${TAG}_c5_full1.log 2>&1
timeout 200 python bench.py --workload c3 > $OUT/${TAG}_c3.json 2> $OUT/${TAG}_c3.err
date +%s > $OUT/${TAG}_t1
ls -la $OUT | tail -30
