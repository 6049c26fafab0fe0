#!/bin/bash
set -x
mkdir -p gpurun_out
O=gpurun_out
MINI_OPT_B200_LIB=mini_opt_b200/libminiopt_b200_trace.so python bench.py --workload ik --links 6 --batch 100000 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/c21_ik_trace6.log 2>&1
MINI_OPT_B200_LIB=mini_opt_b200/libminiopt_b200_trace.so python bench.py --workload ik --links 12 --batch 50000 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/c21_ik_trace12.log 2>&1
echo done
