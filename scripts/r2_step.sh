#!/bin/bash
# one GPU call of round 2: GPU test subset + config-3 bench lines (argument = tag)
OUT=gpurun_out; TAG=${1:-r2b}
mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_stream.py tests/test_gpu_synth_golden.py tests/test_gpu_sharding.py tests/test_gpu_fullsize.py -m gpu -q -x 2>&1 | tail -15 > $OUT/${TAG}_tests.log
timeout 120 python bench.py --workload c3 --no-cpu --no-nested --steps 5 > $OUT/${TAG}_c3.json 2> $OUT/${TAG}_c3.err
RTB200_CULLSCAN=0 timeout 120 python bench.py --workload c3 --no-cpu --no-nested --steps 5 > $OUT/${TAG}_c3_tiled.json 2> $OUT/${TAG}_c3_tiled.err
tail -3 $OUT/${TAG}_tests.log; tail -c 600 $OUT/${TAG}_c3.err
