#!/bin/bash
# GPU call 10: tpp async refill: parity tests, A/B, default bench
set -x
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_qp.py -x -q -m gpu > $O/c10_pytest_qp.log 2>&1; echo "rc=$?" >> $O/c10_pytest_qp.log
python tools/tpp_ab.py > $O/c10_tpp_ab.jsonl 2> $O/c10_tpp_ab.err
python tools/tpp_ab.py --k 0 --batch 524288 >> $O/c10_tpp_ab.jsonl 2>> $O/c10_tpp_ab.err
python tools/tpp_ab.py --n 4 --k 1 --batch 524288 >> $O/c10_tpp_ab.jsonl 2>> $O/c10_tpp_ab.err
timeout 600 python bench.py --no-also > $O/c10_bench.json 2> $O/c10_bench.err; echo "rc=$?" >> $O/c10_bench.err
echo done
