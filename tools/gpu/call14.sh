#!/bin/bash
# GPU call 14: full GPU suite on HEAD, smoke, default bench (both arms), N=1 sweep, launch list of the default bench command
set -x
mkdir -p gpurun_out
O=gpurun_out
MO_RECORD_FRACTIONS=$O/c14_fractions.jsonl timeout 1500 python -m pytest tests -x -q -m gpu > $O/c14_pytest.log 2>&1; echo "rc=$?" >> $O/c14_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/c14_smoke.log 2>&1; echo "rc=$?" >> $O/c14_smoke.log
timeout 600 python bench.py > $O/c14_bench.json 2> $O/c14_bench.err; echo "rc=$?" >> $O/c14_bench.err
timeout 600 python bench.py --impl reference > $O/c14_bench_ref.json 2> $O/c14_bench_ref.err; echo "rc=$?" >> $O/c14_bench_ref.err
timeout 900 python bench.py --sweep > $O/c14_sweep_N1.jsonl 2> $O/c14_sweep.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-also"
$CMD > $O/c14_plain_n8.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/c14_launches_n8.csv $CMD > $O/c14_ncu_list.log 2>&1
echo done
