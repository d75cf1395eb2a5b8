#!/bin/bash
# ncu --set full capture of the render kernel of one config.  bash tools/gpu_ncu2.sh tag cfg scale
set -u
TAG=$1; CFG=$2; SCALE=$3
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python tools/ncu_cfg.py $CFG $SCALE > $OUT/${TAG}_plain.log 2>&1 && cat $OUT/${TAG}_plain.log &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_ -s 2 -c 1 -f -o $OUT/${TAG}_render python tools/ncu_cfg.py $CFG $SCALE > $OUT/${TAG}_ncu.log 2>&1
echo "exit $?"; tail -3 $OUT/${TAG}_ncu.log
