#!/bin/bash
# Measurement refresh after the warp-bundle culling / resident-kernel changes (ONE GPU, under gpurun).
TAG=${1:-r1n}
OUT=gpurun_out
mkdir -p $OUT
set -x
timeout 300 python -m pytest tests -m gpu -x -q --durations=5 > $OUT/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> $OUT/${TAG}_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> $OUT/${TAG}_smoke.log
timeout 200 python bench.py > $OUT/${TAG}_c2.json 2> $OUT/${TAG}_c2.err
timeout 120 python bench.py --impl reference > $OUT/${TAG}_c2_reference.json 2> $OUT/${TAG}_c2_reference.err
timeout 200 python bench.py --workload c3 > $OUT/${TAG}_c3.json 2> $OUT/${TAG}_c3.err
timeout 200 python bench.py --workload c4 > $OUT/${TAG}_c4.json 2> $OUT/${TAG}_c4.err
timeout 200 python bench.py --workload c5 > $OUT/${TAG}_c5.json 2> $OUT/${TAG}_c5.err
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_c2_launches.csv \
    python bench.py --steps 20 --warmup 3 --no-cpu > $OUT/${TAG}_c2_ncu.log 2>&1
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/${TAG}_c3_launches.csv \
    python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu > $OUT/${TAG}_c3_ncu.log 2>&1
timeout 200 ncu --set full --clock-control none --import-source on -k regex:k_trace_shadow -c 1 -f -o $OUT/${TAG}_c3_shadow_culled \
    python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu > $OUT/${TAG}_c3_full.log 2>&1
timeout 100 python scripts/bench_scenes.py > $OUT/${TAG}_scenes.jsonl 2> $OUT/${TAG}_scenes.err
ls -la $OUT | tail -20
