#!/bin/bash
# Short GPU visit: parity tests + smoke + bench + per-config device times.  bash tools/gpu_quick.sh [tag]
set -u
TAG=${1:-q}
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest -m gpu"; timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee $OUT/${TAG}_pytest_gpu.txt
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5 | tee $OUT/${TAG}_smoke.txt
echo "== bench"; timeout 900 python bench.py --steps 20 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "exit $?"; cat $OUT/${TAG}_bench.json; tail -5 $OUT/${TAG}_bench.err
echo "== bench reference"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; cat $OUT/${TAG}_bench_ref.json; tail -3 $OUT/${TAG}_bench_ref.err
echo "== quick bench all configs"; timeout 600 python tools/quick_bench.py 2>&1 | tee $OUT/${TAG}_quick.txt
