#!/bin/bash
# GPU call 9: IK barrier-mask A/B, full GPU suite on HEAD, smoke, default bench (both arms)
set -x
mkdir -p gpurun_out
O=gpurun_out
python tools/ik_ab.py --links 6 --batch 100000 > $O/c9_ik_ab.jsonl 2> $O/c9_ik_ab.err
python tools/ik_ab.py --links 12 --batch 100000 --reps 2 >> $O/c9_ik_ab.jsonl 2>> $O/c9_ik_ab.err
python tools/ik_ab.py --links 6 --batch 50000 --n-eq 1 >> $O/c9_ik_ab.jsonl 2>> $O/c9_ik_ab.err
MO_RECORD_FRACTIONS=$O/c9_fractions.jsonl timeout 1500 python -m pytest tests -x -q -m gpu > $O/c9_pytest.log 2>&1; echo "rc=$?" >> $O/c9_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/c9_smoke.log 2>&1; echo "rc=$?" >> $O/c9_smoke.log
timeout 600 python bench.py > $O/c9_bench.json 2> $O/c9_bench.err; echo "rc=$?" >> $O/c9_bench.err
timeout 600 python bench.py --impl reference > $O/c9_bench_ref.json 2> $O/c9_bench_ref.err; echo "rc=$?" >> $O/c9_bench_ref.err
echo done
