#!/bin/bash
set -x
mkdir -p gpurun_out
./tools/ubench/diag_block_bench > gpurun_out/c20_diag_bench.txt 2>&1
echo done
