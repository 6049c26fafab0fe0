#!/bin/bash
# usage: call22.sh N : torchrun default bench line at N GPUs (both arms)
set -x
N=$1
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29531 bench.py --gpus $N --steps 5 --warmup 3 > $O/c22_bench_${N}gpu.json 2> $O/c22_bench_${N}gpu.err
timeout 600 $TR --master-port 29532 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $O/c22_ref_${N}gpu.json 2> $O/c22_ref_${N}gpu.err
echo done
