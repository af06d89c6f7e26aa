#!/bin/bash
# A/B of library variants on config 3 inside ONE box (scripts/build_variant.sh makes ab/<name>/librtb200.so).
# usage: VARIANTS="base small pair" ROUNDS=2 scripts/ab_c3.sh [tag]
OUT=gpurun_out; TAG=${1:-ab}
mkdir -p $OUT
for round in $(seq ${ROUNDS:-2}); do
for v in ${VARIANTS:-base}; do
  RTB200_LIB=$PWD/ab/$v/librtb200.so timeout ${LIMIT:-40} python scripts/ab_c3.py ${FRAMES:-7} 2>&1 | tail -3
done
done | tee $OUT/${TAG}_ab_c3.log
