#!/bin/bash
# GPU call 17: chain-IK kernel at 32 resident warps per SM (64 registers): one CTA of 32 warps, or two CTAs of 16
set -x
mkdir -p gpurun_out
O=gpurun_out
python tools/ik_ab.py --links 6 --batch 100000 > $O/c17_ik_ab.jsonl 2> $O/c17_ik_ab.err
python tools/ik_ab.py --links 12 --batch 100000 --reps 2 >> $O/c17_ik_ab.jsonl 2>> $O/c17_ik_ab.err
python tools/ik_ab.py --links 6 --batch 50000 --n-eq 1 >> $O/c17_ik_ab.jsonl 2>> $O/c17_ik_ab.err
echo done
