#!/bin/bash
# A/B of library builds under build_var/: device time of the render kernel on the given configs.
OUT=gpurun_out; mkdir -p $OUT
TAG=${1:-ab}; shift
CFGS=${*:-c2}
echo "default:"; python tools/quick_bench.py $CFGS 2>&1 | tee $OUT/${TAG}_default.txt
for lib in build_var/*.so; do echo "$lib:"; RTE_B200_LIB=$PWD/$lib python tools/quick_bench.py $CFGS 2>&1 | tee $OUT/${TAG}_$(basename $lib .so).txt; done
