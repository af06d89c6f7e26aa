#!/bin/bash
# ncu --set full of one launch of a kernel of config 3 (args: tag, kernel regex, launches to skip)
OUT=gpurun_out; TAG=$1; K=$2; SKIP=${3:-0}
mkdir -p $OUT
timeout 300 ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -f -o $OUT/${TAG} \
    python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu --no-nested > $OUT/${TAG}.log 2>&1
tail -3 $OUT/${TAG}.log
