#!/bin/bash
# GPU call 4: full GPU test suite (with measured fractions), the no-alloc test alone in a fresh process, full bench line
set -x
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $O/c4_pytest_zz_alone.log 2>&1; echo "rc=$?" >> $O/c4_pytest_zz_alone.log
export MO_RECORD_FRACTIONS=$PWD/$O/c4_gpu_test_fractions.jsonl
rm -f $MO_RECORD_FRACTIONS
timeout 900 python -m pytest tests -q -m gpu > $O/c4_pytest.log 2>&1; echo "rc=$?" >> $O/c4_pytest.log
unset MO_RECORD_FRACTIONS
python bench.py --steps 5 --warmup 3 > $O/c4_bench_full.json 2> $O/c4_bench_full.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/c4_bench_ref.json 2> $O/c4_bench_ref.err
python -c "import __graft_entry__ as g; g.smoke()" > $O/c4_smoke.log 2>&1
echo done
