#!/bin/bash
# GPU call 11: ncu --set full of the tpp iterate kernel with the cp.async refill
set -x
mkdir -p gpurun_out
O=gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-also --no-e2e"
$CMD > $O/c11_plain_n8.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:qp_tpp_iterate -s 3 -c 1 -o $O/c11_tpp_n8 $CMD > $O/c11_ncu_tpp.log 2>&1
python tools/ncu_summary.py $O/c11_tpp_n8.ncu-rep > $O/c11_tpp_n8_summary.txt 2>&1
python tools/ncu_windows.py $O/c11_tpp_n8.ncu-rep 200 >> $O/c11_tpp_n8_summary.txt 2>&1
python tools/ncu_stalls.py $O/c11_tpp_n8.ncu-rep > $O/c11_tpp_n8_stalls.txt 2>&1
rm -f $O/*.ncu-rep
echo done
