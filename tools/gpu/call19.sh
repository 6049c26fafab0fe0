#!/bin/bash
set -x
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_cpp_host_api.py -x -q -m gpu > $O/c19_pytest.log 2>&1; echo "rc=$?" >> $O/c19_pytest.log
echo done
