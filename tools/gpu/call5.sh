#!/bin/bash
# GPU call 5: ncu evidence (launch list of the default bench command, full captures of the three main kernels, warp16 TMA
# vs lane staging), linearization bench, config-5 sweep at N=1
set -x
mkdir -p gpurun_out
O=gpurun_out
python tools/lin_bench.py --n 64 --r 128 --batch 65536 > $O/c5_lin_n64.json 2> $O/c5_lin.err
python tools/lin_bench.py --n 32 --r 64 --k 8 --batch 262144 >> $O/c5_lin_n64.json 2>> $O/c5_lin.err
python tools/lin_bench.py --n 12 --r 24 --k 3 --batch 1048576 >> $O/c5_lin_n64.json 2>> $O/c5_lin.err
python bench.py --sweep > $O/c5_sweep_N1.jsonl 2> $O/c5_sweep.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-also"
$CMD > $O/c5_plain_n8.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/c5_launches_n8.csv $CMD > $O/c5_ncu_list.log 2>&1
$CMD > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:qp_tpp_iterate -s 3 -c 1 -o $O/c5_tpp_n8 $CMD > $O/c5_ncu_tpp.log 2>&1
$CMD > /dev/null 2>&1 && \
ncu --set full --clock-control none -k regex:qp_tpp_guess -s 3 -c 1 -o $O/c5_tppguess_n8 $CMD > $O/c5_ncu_tppg.log 2>&1
CMD64="python bench.py --n 64 --k 16 --batch 8288 --no-cpu --no-e2e --no-also --steps 1 --warmup 3"
$CMD64 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:qp_wd_kernel -s 3 -c 1 -o $O/c5_wd64 $CMD64 > $O/c5_ncu_wd64.log 2>&1
CMDW="python tools/w16_ab.py --n 6 --k 2 --batch 262144"
$CMDW > $O/c5_w16_plain.json 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:qp_warp16 -s 3 -c 1 -o $O/c5_w16_tma $CMDW > $O/c5_ncu_w16a.log 2>&1
$CMDW > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:qp_warp16 -s 16 -c 1 -o $O/c5_w16_lanes $CMDW > $O/c5_ncu_w16b.log 2>&1
# summarise on the box, return text only (the .ncu-rep files exceed the 64 MiB return limit)
for r in tpp_n8 tppguess_n8 wd64 w16_tma w16_lanes; do
  if [ -f $O/c5_$r.ncu-rep ]; then
    python tools/ncu_summary.py $O/c5_$r.ncu-rep > $O/c5_${r}_summary.txt 2>&1
    python tools/ncu_windows.py $O/c5_$r.ncu-rep 400 >> $O/c5_${r}_summary.txt 2>&1
  fi
done
rm -f $O/*.ncu-rep
du -sh $O
echo done
