#!/bin/bash
# GPU call 25: ncu of the lock-step IK kernel at HEAD (W = 24) on a short run
set -x
mkdir -p gpurun_out
O=gpurun_out
CMD="python bench.py --workload ik --links 6 --batch 30000 --steps 1 --warmup 3 --no-cpu --no-e2e"
$CMD > $O/c25_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:nls_solve_kernel -s 3 -c 1 -o $O/c25_ik24 $CMD > $O/c25_ncu.log 2>&1
python tools/ncu_summary.py $O/c25_ik24.ncu-rep > $O/c25_ik24_summary.txt 2>&1
python tools/ncu_windows.py $O/c25_ik24.ncu-rep 500 >> $O/c25_ik24_summary.txt 2>&1
python tools/ncu_stalls.py $O/c25_ik24.ncu-rep > $O/c25_ik24_stalls.txt 2>&1
rm -f $O/*.ncu-rep
echo done
