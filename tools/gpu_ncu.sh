#!/bin/bash
# ncu capture of the render kernel (after a plain run of the same command).  bash tools/gpu_ncu.sh [tag] [extra bench args]
set -u
TAG=${1:-p}; shift
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline $*"
timeout 600 $CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_ -s 3 -c 1 -f -o $OUT/${TAG}_render $CMD > $OUT/${TAG}_ncu_full.log 2>&1
echo "exit $?"; tail -3 $OUT/${TAG}_ncu_full.log
