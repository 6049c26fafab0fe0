#!/bin/bash
# GPU call 7: lock-step IK kernel A/B over warps per CTA; NLS tests
set -x
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_nls.py -x -q -m gpu > $O/c7_pytest_nls.log 2>&1; echo "rc=$?" >> $O/c7_pytest_nls.log
python tools/ik_ab.py --links 6 --batch 100000 > $O/c7_ik_ab.jsonl 2> $O/c7_ik_ab.err
python tools/ik_ab.py --links 12 --batch 100000 --reps 2 >> $O/c7_ik_ab.jsonl 2>> $O/c7_ik_ab.err
python tools/ik_ab.py --links 6 --batch 50000 --n-eq 1 >> $O/c7_ik_ab.jsonl 2>> $O/c7_ik_ab.err
echo done
