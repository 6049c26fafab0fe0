#!/bin/bash
# GPU call 12 (8 GPUs): torchrun bench at N=8 (default line + config-5 sweep), multi-device host entry over 1/2/4/8 devices
set -x
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 > $O/c12_bench_8gpu.json 2> $O/c12_bench_8gpu.err
timeout 900 $TR --master-port 29522 bench.py --gpus 8 --sweep > $O/c12_sweep_N8.jsonl 2> $O/c12_sweep_N8.err
for d in 0 0,1 0,1,2,3 0,1,2,3,4,5,6,7; do
  timeout 300 python tools/e2e_multi.py --devices $d >> $O/c12_e2e_multi.jsonl 2>> $O/c12_e2e_multi.err
done
nvidia-smi topo -m > $O/c12_topo.txt 2>&1
echo done
