#!/bin/bash
# Strong-scaling visit on an N-GPU box: bash tools/gpu_scale.sh tag N  (bench.py at N ranks: C2 peer / nccl gather, C5 auto-white)
TAG=$1; N=$2; OUT=gpurun_out; mkdir -p $OUT
run() { name=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline "$@" > $OUT/${TAG}_${name}_n$N.json 2> $OUT/${TAG}_${name}_n$N.err; echo "$name n=$N exit $?"; python -c "
import json,sys
d=json.load(open('$OUT/${TAG}_${name}_n$N.json'))
print(' ', d['value'], d['unit'], d['ms_per_step'], 'ms/step; e2e', d['e2e']['value'], '; kernel_ms', d['roofline'].get('kernel_ms'))" 2>&1 | tail -1; }
run c2peer
run c2nccl --gather nccl
run c5 --workload c5
