#!/bin/bash
# End-of-round measurement set (ONE GPU, under gpurun): the four bench workloads with their CPU arms, the
# reference arm, then the ncu launch list and one --set full capture of the headline kernel.
# usage: scripts/final_round.sh <tag>
TAG=${1:-r1f}
OUT=gpurun_out
set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $OUT/${TAG}_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1
timeout 400 python bench.py > $OUT/${TAG}_c2.json 2> $OUT/${TAG}_c2.err || exit 1
timeout 300 python bench.py --impl reference > $OUT/${TAG}_c2_reference.json 2> $OUT/${TAG}_c2_reference.err
timeout 400 python bench.py --workload c3 > $OUT/${TAG}_c3.json 2> $OUT/${TAG}_c3.err
timeout 400 python bench.py --workload c4 > $OUT/${TAG}_c4.json 2> $OUT/${TAG}_c4.err
timeout 500 python bench.py --workload c5 > $OUT/${TAG}_c5.json 2> $OUT/${TAG}_c5.err
timeout 300 python scripts/bench_scenes.py > $OUT/${TAG}_scenes.jsonl 2> $OUT/${TAG}_scenes.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_c2_launches.csv \
    python bench.py --steps 20 --warmup 3 --no-cpu > $OUT/${TAG}_c2_ncu.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_resident_level -s 5 -c 1 -f -o $OUT/${TAG}_c2_resident \
    python bench.py --steps 20 --warmup 3 --no-cpu > $OUT/${TAG}_c2_full.log 2>&1
ls -la $OUT | tail -20
