#!/bin/bash
# GPU call 16: lock-step CTAs of the DMMA warp kernel: parity tests, A/B, phase trace
set -x
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_qp.py -x -q -m gpu -k "dmma or n64 or n32 or n16 or wd or medium or shapes" > $O/c16_pytest.log 2>&1; echo "rc=$?" >> $O/c16_pytest.log
python tools/wd_ab.py > $O/c16_wd_ab.jsonl 2> $O/c16_wd_ab.err
python tools/wd_ab.py --n 32 --k 8 --batch 262144 >> $O/c16_wd_ab.jsonl 2>> $O/c16_wd_ab.err
python tools/wd_ab.py --n 16 --k 8 --batch 524288 >> $O/c16_wd_ab.jsonl 2>> $O/c16_wd_ab.err
MINI_OPT_B200_LIB=mini_opt_b200/libminiopt_b200_trace.so python bench.py --n 64 --k 16 --batch 8288 --no-cpu --no-e2e --no-also --steps 1 --warmup 1 > $O/c16_trace64.log 2>&1
echo done
