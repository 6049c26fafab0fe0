#!/bin/bash
# GPU call 13: device residual families + host logging tests; NLS regression; IK A/B sanity after the family refactor
set -x
mkdir -p gpurun_out
O=gpurun_out
MO_RECORD_FRACTIONS=$O/c13_fractions.jsonl timeout 900 python -m pytest tests/test_gpu_nls_families.py tests/test_cpp_host_api.py tests/test_gpu_nls.py -q -m gpu -s > $O/c13_pytest.log 2>&1; echo "rc=$?" >> $O/c13_pytest.log
python tools/ik_ab.py --links 6 --batch 100000 > $O/c13_ik_ab.jsonl 2> $O/c13_ik_ab.err
echo done
