#!/bin/bash
set -x
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_nls.py tests/test_gpu_nls_families.py -x -q -m gpu > $O/c24_pytest.log 2>&1; echo "rc=$?" >> $O/c24_pytest.log
python tools/ik_ab.py --links 6 --batch 100000 > $O/c24_ik_ab.jsonl 2> $O/c24_ik_ab.err
python tools/ik_ab.py --links 12 --batch 100000 --reps 2 >> $O/c24_ik_ab.jsonl 2>> $O/c24_ik_ab.err
echo done
