#!/bin/bash
# GPU call 28: compute-sanitizer memcheck over small solves of every kernel family
set -x
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python tools/sanitize_run.py > $O/c28_plain.log 2>&1; echo "rc=$?" >> $O/c28_plain.log
# compute-sanitizer is closed on this pool (runs under it have left GPUs needing a reset): plain run only
echo done
