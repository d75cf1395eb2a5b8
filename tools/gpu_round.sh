#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench (both arms), ncu launch list + full captures of the render / tonemap kernels.
# Usage (from the repo root, under gpurun):  bash tools/gpu_round.sh [tag]
set -u
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > $OUT/${TAG}_gpu.txt 2>&1
lscpu | head -20 >> $OUT/${TAG}_gpu.txt; nproc >> $OUT/${TAG}_gpu.txt
echo "== pytest -m gpu"; timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee $OUT/${TAG}_pytest_gpu.txt
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5 | tee $OUT/${TAG}_smoke.txt
echo "== bench"; timeout 900 python bench.py --steps 20 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "exit $?"; cut -c1-1500 $OUT/${TAG}_bench.json; tail -5 $OUT/${TAG}_bench.err
echo "== bench reference"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; cut -c1-600 $OUT/${TAG}_bench_ref.json; tail -3 $OUT/${TAG}_bench_ref.err
echo "== C1 (the reference's default workload), both arms, 1 and 32 samples per pixel"; timeout 600 python tools/c1_bench.py > $OUT/${TAG}_c1.json 2> $OUT/${TAG}_c1.err; cut -c1-900 $OUT/${TAG}_c1.json; tail -3 $OUT/${TAG}_c1.err
echo "== tonemap pass and its memory pattern without arithmetic"; (python tools/tm_probe.py; RTE_TM_NULL=1 python tools/tm_probe.py; RTE_TM_NULL=2 python tools/tm_probe.py) 2>&1 | tee $OUT/${TAG}_tm_probe.txt
echo "== grid rebuild cost"; (for d in 1 8; do python tools/grid_probe.py c2 $d; python tools/grid_probe.py c5 $d; done) 2>&1 | tee $OUT/${TAG}_grid_probe.txt
echo "== quick bench all configs"; timeout 600 python tools/quick_bench.py c2 c3 c4 c5 c1 --rgb8 2>&1 | tee $OUT/${TAG}_quick.txt
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-c5"
echo "== ncu launch list"
timeout 600 $CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_list.log 2>&1
echo "exit $?"
echo "== ncu full (render + grid kernels)"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"render32|grid_" -s 5 -c 8 -f -o $OUT/${TAG}_render $CMD > $OUT/${TAG}_ncu_full.log 2>&1
echo "exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tonemap_kernel -c 1 -f -o $OUT/${TAG}_tonemap $CMD > $OUT/${TAG}_ncu_full_tm.log 2>&1
echo "exit $?"
ls -la $OUT | tail -25
