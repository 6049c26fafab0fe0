#!/bin/bash
# GPU call 15: phase trace of the n=64 / n=32 warp-per-problem DMMA kernel (debug build with -DMO_WD_TRACE)
set -x
mkdir -p gpurun_out
O=gpurun_out
MINI_OPT_B200_LIB=mini_opt_b200/libminiopt_b200_trace.so python bench.py --n 64 --k 16 --batch 8288 --no-cpu --no-e2e --no-also --steps 1 --warmup 1 > $O/c15_trace64.log 2>&1
MINI_OPT_B200_LIB=mini_opt_b200/libminiopt_b200_trace.so python bench.py --n 32 --k 8 --batch 14208 --no-cpu --no-e2e --no-also --steps 1 --warmup 1 > $O/c15_trace32.log 2>&1
echo done
