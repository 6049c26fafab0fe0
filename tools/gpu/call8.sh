#!/bin/bash
# GPU call 8: ncu of the lock-step IK kernel (W = 16) on a short run
set -x
mkdir -p gpurun_out
O=gpurun_out
CMD="python bench.py --workload ik --links 6 --batch 30000 --steps 1 --warmup 3 --no-cpu --no-e2e"
$CMD > $O/c8_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:nls_chain_ik -s 3 -c 1 -o $O/c8_ik16 $CMD > $O/c8_ncu.log 2>&1
python tools/ncu_summary.py $O/c8_ik16.ncu-rep > $O/c8_ik16_summary.txt 2>&1
python tools/ncu_windows.py $O/c8_ik16.ncu-rep 500 >> $O/c8_ik16_summary.txt 2>&1
python tools/ncu_stalls.py $O/c8_ik16.ncu-rep > $O/c8_ik16_stalls.txt 2>&1
rm -f $O/*.ncu-rep
echo done
