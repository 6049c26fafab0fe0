#!/bin/bash
set -x
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > $O/c29_pytest.log 2>&1; echo "rc=$?" >> $O/c29_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/c29_smoke.log 2>&1; echo "rc=$?" >> $O/c29_smoke.log
echo done
