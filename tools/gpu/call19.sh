#!/bin/bash
set -x
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_qp.py -x -q -m gpu -k "refill" > $O/c19_pytest.log 2>&1; echo "rc=$?" >> $O/c19_pytest.log
echo done
