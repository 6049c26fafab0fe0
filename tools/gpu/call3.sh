#!/bin/bash
# GPU call 3: n=64 kernel variants (v0 baseline+diet, v1 +pipelined backsub, v2 +look-ahead columns), ncu source-level
# profiles of v1/v2, lane-activity microbenchmark, measured test fractions, parity for the remaining tpp shapes
set -x
mkdir -p gpurun_out
O=gpurun_out
export MO_RECORD_FRACTIONS=$PWD/$O/c3_gpu_test_fractions.jsonl
rm -f $MO_RECORD_FRACTIONS
timeout 900 python -m pytest tests -x -q -m gpu > $O/c3_pytest.log 2>&1; echo "rc=$?" >> $O/c3_pytest.log
unset MO_RECORD_FRACTIONS
./tools/ubench/lanes_bench > $O/c3_lanes_bench.txt 2>&1
B="python bench.py --no-cpu --no-e2e --no-also --steps 5 --warmup 3"
for v in v0 v1 v2; do
  for cfg in "64 16 65536" "32 8 262144" "16 8 524288"; do set -- $cfg
    MINI_OPT_B200_LIB=$PWD/mini_opt_b200/libminiopt_b200_$v.so $B --n $1 --k $2 --batch $3 > $O/c3_bench_n$1_$v.json 2>> $O/c3_bench.err
  done
done
P="python tools/parity.py --out $O/c3_parity.jsonl"
$P --n 8 --k 0 --batch 262144 > $O/c3_par.log 2>&1
$P --n 4 --k 1 --batch 262144 >> $O/c3_par.log 2>&1
$P --n 10 --k 2 --batch 131072 >> $O/c3_par.log 2>&1
$P --n 64 --k 16 --batch 32768 >> $O/c3_par.log 2>&1
# ncu: v1 and v2 on a short n=64 run (plain run first, same command)
for v in v1 v2; do
  CMD="python bench.py --n 64 --k 16 --batch 8288 --no-cpu --no-e2e --no-also --steps 1 --warmup 3"
  MINI_OPT_B200_LIB=$PWD/mini_opt_b200/libminiopt_b200_$v.so $CMD > $O/c3_ncu_plain_$v.log 2>&1 && \
  MINI_OPT_B200_LIB=$PWD/mini_opt_b200/libminiopt_b200_$v.so ncu --set full --clock-control none --import-source on -k regex:qp_wd_kernel -s 3 -c 1 -o $O/c3_wd64_$v $CMD > $O/c3_ncu_$v.log 2>&1
done
echo done
