#!/bin/bash
# compute-sanitizer is CLOSED on this GPU pool (profiles/r2_sanitizer_closed_on_this_pool.log).  Instead: the CHECKED
# build of the library (make checked: device-side bounds checks, RT_CHECK in rt_math.cuh) under the small frames of
# scripts/sanitize_driver.py and under the whole GPU test-suite.  Argument: number of GPUs of this call (1 or 2).
N=${1:-1}; OUT=gpurun_out; mkdir -p $OUT
export RTB200_LIB=ab/checked/librtb200.so
if [ "$N" = "1" ]; then
  timeout 600 python scripts/sanitize_driver.py > $OUT/r2_checked_build_driver.log 2>&1; echo "exit code $?" >> $OUT/r2_checked_build_driver.log
  tail -3 $OUT/r2_checked_build_driver.log
  timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -5 > $OUT/r2_checked_build_gpu_tests.log; cat $OUT/r2_checked_build_gpu_tests.log
else
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 \
      scripts/sanitize_driver.py > $OUT/r2_checked_build_driver_${N}gpu.log 2>&1; echo "exit code $?" >> $OUT/r2_checked_build_driver_${N}gpu.log
  grep -v Warn $OUT/r2_checked_build_driver_${N}gpu.log | tail -6
fi
