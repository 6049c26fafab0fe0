#!/bin/bash
# GPU call 2: look-ahead schedule of the n=64 kernel (A/B + phase trace), warp16 kernel (parity + TMA A/B)
set -x
mkdir -p gpurun_out
O=gpurun_out
ALT=$PWD/mini_opt_b200/libminiopt_b200_alt.so
TR=$PWD/mini_opt_b200/libminiopt_b200_trace.so
timeout 600 python -m pytest tests/test_gpu_qp.py -x -q -m gpu > $O/c2_pytest.log 2>&1; echo "rc=$?" >> $O/c2_pytest.log
MINI_OPT_B200_LIB=$ALT timeout 600 python -m pytest tests/test_gpu_qp.py -x -q -m gpu -k "32 or 16 or dmma" > $O/c2_pytest_alt.log 2>&1; echo "rc=$?" >> $O/c2_pytest_alt.log
B="python bench.py --no-cpu --no-e2e --no-also --steps 5 --warmup 3"
for cfg in "64 16 65536" "32 8 262144" "16 8 524288"; do set -- $cfg
  $B --n $1 --k $2 --batch $3 > $O/c2_bench_n$1_default.json 2>> $O/c2_bench.err
  MINI_OPT_B200_LIB=$ALT $B --n $1 --k $2 --batch $3 > $O/c2_bench_n$1_alt.json 2>> $O/c2_bench.err
done
MINI_OPT_B200_LIB=$TR python bench.py --n 64 --k 16 --batch 65536 --no-cpu --no-e2e --no-also --steps 2 --warmup 3 2>&1 | grep "wd trace" | tail -3 > $O/c2_trace_n64_la.log
python tools/parity.py --n 64 --k 16 --batch 65536 --out $O/c2_parity.jsonl > $O/c2_par.log 2>&1
python tools/parity.py --n 12 --k 3 --batch 131072 --out $O/c2_parity.jsonl >> $O/c2_par.log 2>&1
python tools/parity.py --n 6 --k 2 --batch 131072 --out $O/c2_parity.jsonl >> $O/c2_par.log 2>&1
python tools/parity.py --n 8 --k 2 --batch 131072 --triplets --out $O/c2_parity.jsonl >> $O/c2_par.log 2>&1
python tools/w16_ab.py --n 12 --k 3 --batch 262144 > $O/c2_w16_ab.json 2> $O/c2_w16_ab.err
python tools/w16_ab.py --n 6 --k 2 --batch 524288 >> $O/c2_w16_ab.json 2>> $O/c2_w16_ab.err
python bench.py --steps 5 --warmup 3 > $O/c2_bench_full.json 2> $O/c2_bench_full.err
echo done
