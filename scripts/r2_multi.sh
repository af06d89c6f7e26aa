#!/bin/bash
# multi-GPU step (argument 1 = number of GPUs, 2 = tag): frame-assembly parity, then the default bench line (config 5 strong)
N=${1:-2}; TAG=${2:-r2m}; OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 300 $TR scripts/check_multigpu.py > $OUT/${TAG}_parity_${N}gpu.log 2>&1; tail -4 $OUT/${TAG}_parity_${N}gpu.log
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 > $OUT/${TAG}_c5_${N}gpu.json 2> $OUT/${TAG}_c5_${N}gpu.err; tail -c 400 $OUT/${TAG}_c5_${N}gpu.err
timeout 300 $TR bench.py --gpus $N --steps 20 --warmup 3 --workload c2 --no-cpu > $OUT/${TAG}_c2_${N}gpu.json 2> $OUT/${TAG}_c2_${N}gpu.err; tail -c 400 $OUT/${TAG}_c2_${N}gpu.err
