#!/bin/bash
# thresholds of the culled scan on config 3 inside ONE box: which tasks go to the lane kernels (RTB200_LANE_SIN,
# RTB200_LANE_SIN_SHADOW) and below which level size tasks shrink to 32 rays (RTB200_SMALL_LEVEL)
OUT=gpurun_out; TAG=${1:-sweep}
mkdir -p $OUT
run() { echo "== $*"; env "$@" timeout ${LIMIT:-30} python scripts/ab_c3.py ${FRAMES:-5} 2>&1 | tail -2; }
{
run X=default
for a in 0.02 0.05 0.2; do run RTB200_LANE_SIN=$a; done
for b in 0.05 0.1 0.15 0.4; do run RTB200_LANE_SIN_SHADOW=$b; done
for c in 0 100000 3000000; do run RTB200_SMALL_LEVEL=$c; done
run RTB200_LANE_SIN=0.05 RTB200_LANE_SIN_SHADOW=0.1
run RTB200_LANE_SIN=0.02 RTB200_LANE_SIN_SHADOW=0.05
} | tee $OUT/${TAG}_sweep_c3.log
