#!/bin/bash
set -x
mkdir -p gpurun_out
O=gpurun_out
python bench.py --n 64 --k 16 --batch 65536 --no-cpu --no-e2e --no-also --steps 5 --warmup 3 > $O/c27_base.json 2>&1
MINI_OPT_B200_LIB=mini_opt_b200/libminiopt_b200_trace.so python bench.py --n 64 --k 16 --batch 65536 --no-cpu --no-e2e --no-also --steps 5 --warmup 3 > $O/c27_inline.json 2>&1
python bench.py --n 32 --k 8 --batch 262144 --no-cpu --no-e2e --no-also --steps 5 --warmup 3 > $O/c27_base32.json 2>&1
MINI_OPT_B200_LIB=mini_opt_b200/libminiopt_b200_trace.so python bench.py --n 32 --k 8 --batch 262144 --no-cpu --no-e2e --no-also --steps 5 --warmup 3 > $O/c27_inline32.json 2>&1
echo done
