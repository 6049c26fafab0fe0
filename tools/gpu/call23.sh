#!/bin/bash
# GPU call 23 (4 GPUs): single-process multi-device e2e with plain vs NUMA-interleaved pinned buffers
set -x
mkdir -p gpurun_out
O=gpurun_out
cat /sys/devices/system/node/online > $O/c23_numa.txt 2>&1; nproc >> $O/c23_numa.txt; lscpu | grep -i "numa\|socket\|model name" >> $O/c23_numa.txt 2>&1
for il in 0 1 0 1; do
  timeout 300 python tools/e2e_multi.py --devices 0,1,2,3 --interleave $il >> $O/c23_e2e_multi.jsonl 2>> $O/c23_e2e_multi.err
done
echo done
