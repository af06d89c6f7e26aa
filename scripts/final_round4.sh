#!/bin/bash
# Last call of round 1 (about 100 s of GPU budget left): steps ordered by priority, each writing its own file.
OUT=gpurun_out; TAG=${1:-r1r}
set -x
timeout 60 python -m pytest tests/test_gpu_stream.py tests/test_gpu_fullsize.py tests/test_gpu_bvh.py -m gpu -q -x 2>&1 | tail -4 > $OUT/${TAG}_tests.log
timeout 80 python bench.py > $OUT/${TAG}_c2.json 2> $OUT/${TAG}_c2.err
timeout 60 python bench.py --workload c3 > $OUT/${TAG}_c3.json 2> $OUT/${TAG}_c3.err
timeout 40 python bench.py --workload c4 --no-cpu > $OUT/${TAG}_c4.json 2> $OUT/${TAG}_c4.err
timeout 40 python bench.py --workload c5 --no-cpu > $OUT/${TAG}_c5.json 2> $OUT/${TAG}_c5.err
