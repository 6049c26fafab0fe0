#!/bin/bash
# GPU call 26: tpp refill A/B on awkward batch sizes (bit-identity of every output array)
set -x
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/c26_tpp_ab.jsonl
for b in 1 2 7 31 32 33 129 1025 37889 75777; do
  python tools/tpp_ab.py --batch $b --reps 2 >> $O/c26_tpp_ab.jsonl 2>> $O/c26_tpp_ab.err
  python tools/tpp_ab.py --n 4 --k 1 --batch $b --reps 2 >> $O/c26_tpp_ab.jsonl 2>> $O/c26_tpp_ab.err
done
echo done
