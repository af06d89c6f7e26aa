#!/bin/bash
# A/B of library variants on configs 5 and 4 inside ONE box.  usage: VARIANTS="chunk8 chunk32" ROUNDS=2 scripts/ab_c5.sh [tag]
OUT=gpurun_out; TAG=${1:-ab}
mkdir -p $OUT
for round in $(seq ${ROUNDS:-2}); do
for v in ${VARIANTS:-base}; do
  RTB200_LIB=$PWD/ab/$v/librtb200.so timeout ${LIMIT:-60} python scripts/ab_c5.py ${FRAMES:-3} 2>&1 | tail -3
done
done | tee $OUT/${TAG}_ab_c5.log
