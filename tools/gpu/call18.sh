#!/bin/bash
# GPU call 18: full GPU suite on HEAD, smoke, default bench (both arms)
set -x
mkdir -p gpurun_out
O=gpurun_out
MO_RECORD_FRACTIONS=$O/c18_fractions.jsonl timeout 1500 python -m pytest tests -x -q -m gpu > $O/c18_pytest.log 2>&1; echo "rc=$?" >> $O/c18_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/c18_smoke.log 2>&1; echo "rc=$?" >> $O/c18_smoke.log
timeout 600 python bench.py > $O/c18_bench.json 2> $O/c18_bench.err; echo "rc=$?" >> $O/c18_bench.err
timeout 600 python bench.py --impl reference > $O/c18_bench_ref.json 2> $O/c18_bench_ref.err; echo "rc=$?" >> $O/c18_bench_ref.err
echo done
