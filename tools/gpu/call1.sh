#!/bin/bash
# GPU call 1 (round 2): tests, parity on identical bytes, baselines, n=64 phase trace.
set -x
mkdir -p gpurun_out
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/c1_pytest.log 2>&1; echo "pytest rc=$?" >> $O/c1_pytest.log
P="python tools/parity.py --out $O/c1_parity.jsonl"
$P --n 8 --k 2 --batch 1048576 > $O/c1_par_n8.log 2>&1
$P --n 8 --k 2 --batch 1048576 --torch-gen >> $O/c1_par_n8.log 2>&1
$P --n 8 --k 2 --batch 262144 --guess naive >> $O/c1_par_n8.log 2>&1
$P --n 8 --k 2 --batch 131072 --triplets >> $O/c1_par_n8.log 2>&1
$P --n 64 --k 16 --batch 65536 > $O/c1_par_n64.log 2>&1
$P --n 64 --k 16 --batch 4096 --triplets >> $O/c1_par_n64.log 2>&1
$P --n 32 --k 8 --batch 131072 >> $O/c1_par_n64.log 2>&1
$P --n 16 --k 8 --batch 131072 >> $O/c1_par_n64.log 2>&1
$P --ik --links 6 --batch 100000 > $O/c1_par_ik.log 2>&1
$P --ik --links 6 --batch 20000 --generic >> $O/c1_par_ik.log 2>&1
$P --ik --links 12 --batch 20000 >> $O/c1_par_ik.log 2>&1
$P --ik --links 12 --batch 20000 --generic >> $O/c1_par_ik.log 2>&1
$P --ik --links 6 --batch 20000 --n-eq 1 >> $O/c1_par_ik.log 2>&1
$P --ik --links 6 --batch 20000 --n-eq 1 --generic >> $O/c1_par_ik.log 2>&1
$P --ik --links 6 --batch 20000 --max-qp-iterations 5 >> $O/c1_par_ik.log 2>&1
python bench.py --steps 10 --warmup 3 > $O/c1_bench_n8.json 2> $O/c1_bench_n8.err
python bench.py --n 64 --k 16 --batch 65536 --cpu-sample 16384 --steps 5 --warmup 3 > $O/c1_bench_n64.json 2> $O/c1_bench_n64.err
python bench.py --n 32 --k 8 --batch 262144 --no-cpu --steps 5 --warmup 3 > $O/c1_bench_n32.json 2> $O/c1_bench_n32.err
MINI_OPT_B200_LIB=$PWD/mini_opt_b200/libminiopt_b200_trace.so python bench.py --n 64 --k 16 --batch 65536 --no-cpu --no-e2e --steps 2 --warmup 3 > $O/c1_trace_n64.log 2>&1
nvidia-smi > $O/c1_smi.log 2>&1
lscpu | head -30 > $O/c1_lscpu.log 2>&1
echo done
