#!/bin/bash
# GPU call 6 (2 GPUs): C++ host API tests, multi-device host entry on 2 real devices, torchrun bench at N=2 (+ reference arm)
set -x
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests -q -m gpu -k "cpp or multi_device or nullspace or no_alloc or compact" > $O/c6_pytest.log 2>&1; echo "rc=$?" >> $O/c6_pytest.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > $O/c6_bench_2gpu.json 2> $O/c6_bench_2gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/c6_ref_2gpu.json 2> $O/c6_ref_2gpu.err
python tools/e2e_multi.py --devices 0,1 > $O/c6_e2e_multi.json 2> $O/c6_e2e_multi.err
nvidia-smi topo -m > $O/c6_topo.txt 2>&1
echo done
