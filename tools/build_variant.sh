#!/bin/bash
# Developer aid: builds build_var/<name>.so from the current sources with extra nvcc flags (A/B with tools/gpu_ab.sh).
#   bash tools/build_variant.sh t288 -DRTE_FAST_THREADS=288
set -eu
NAME=$1; shift
ROOT=$(cd "$(dirname "$0")/.." && pwd)
D=$ROOT/build_var/obj_$NAME
mkdir -p $D
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xptxas -v -Xcompiler -fPIC,-ffp-contract=off,-O2,-Wall -I$ROOT/raytracingengine_b200/csrc"
for f in rte_api rte_render rte_tonemap rte_peak; do
  $NV "$@" -c $ROOT/raytracingengine_b200/csrc/$f.cu -o $D/$f.o > $D/$f.log 2>&1 &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $ROOT/build_var/$NAME.so $D/*.o -cudart shared
grep -A2 "render_fast_kernelILi2ELb0ELb0" $D/rte_render.log | grep -E "spill|Used" || true
