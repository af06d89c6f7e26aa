#!/bin/bash
# Builds ab/<name>/librtb200.so = the current objects with ONE translation unit recompiled with extra -D flags
# (for A/B runs inside one gpurun call; ab/ is git-ignored but travels to the GPU box).
# usage: scripts/build_variant.sh <name> <file.cu> <flags...>      e.g. scripts/build_variant.sh t128 rt_resident.cu -DRES_THREADS=128 -DRES_MIN_CTAS=7
set -e
NAME=$1; SRC=$2; shift 2
ROOT=$(cd "$(dirname "$0")/.." && pwd)
PKG=$ROOT/ray-tracing-gpu_b200
mkdir -p $ROOT/ab/$NAME
OBJ=$ROOT/ab/$NAME/${SRC%.cu}.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -ccbin /usr/bin/g++ \
    -Xcompiler -fPIC,-Wall,-Wno-unused-function -Xptxas -warn-spills "$@" -c $PKG/csrc/$SRC -o $OBJ 2>&1 | grep -v "^$" | grep -i "spill\|error" || true
OTHERS=$(ls $PKG/build/*.o | grep -v "/${SRC%.cu}.o")
/usr/local/cuda/bin/nvcc -shared -ccbin /usr/bin/g++ -o $ROOT/ab/$NAME/librtb200.so $OBJ $OTHERS -ldl 2>/dev/null
cuobjdump -res-usage $OBJ 2>/dev/null | grep -A1 "k_resident_levelILb1ELb1" | grep -o "REG:[0-9]*" || true
ls -la $ROOT/ab/$NAME/librtb200.so
