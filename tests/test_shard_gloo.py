"""Multi-process host logic on CPU (gloo, world_size 2): contiguous sharding of a batch across ranks, independent
solves per rank (the oracle stands in for the GPU here: tests may use it), one gather at the end, identical to the
unsharded solve. No collective is needed on the solve path."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


def test_shard_bounds_partition():
    from mini_opt_b200 import shard
    for batch in (0, 1, 7, 1000, 1 << 20):
        for world in (1, 2, 3, 4, 8):
            b = [shard.shard_bounds(batch, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == batch
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.shard_bounds(10, 2, 2)


def _worker(rank, world, port, batch, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, str(ROOT))
    import torch
    import torch.distributed as dist
    from mini_opt_b200 import abi, shard, synth
    from oracle import oracle_py as O
    dist.init_process_group("gloo", rank=rank, world_size=world)
    arrs = synth.box_qp_batch_numpy(8, 2, batch, seed=99, explicit_constraints=False)
    mine = shard.shard_qp_batch(arrs, rank, world)
    mine = abi.QPBatchArrays(**{k: (np.ascontiguousarray(v) if isinstance(v, np.ndarray) else v) for k, v in mine.__dict__.items()})
    out, steps = O.qp_solve("oracle", mine, synth.solve_params())
    t = abi.QPOutArrays(vars=torch.from_numpy(out.vars), status=torch.from_numpy(out.status), iters=torch.from_numpy(out.iters),
                        kkt_final=torch.from_numpy(out.kkt_final))
    full = shard.gather_qp_out(t, batch)
    total = torch.tensor([float(steps)], dtype=torch.float64)
    dist.all_reduce(total)  # the three timing scalars of bench.py travel the same way
    if rank == 0:
        np.savez(tmp, vars=full.vars.numpy(), status=full.status.numpy(), iters=full.iters.numpy(), total=total.numpy())
    dist.destroy_process_group()


@pytest.mark.parametrize("batch", [1001, 64])
def test_two_rank_sharded_solve_matches_unsharded(tmp_path, oracle, batch):
    import torch.multiprocessing as mp
    from mini_opt_b200 import synth
    port = 29500 + (os.getpid() % 2000)
    tmp = str(tmp_path / "gathered.npz")
    mp.spawn(_worker, args=(2, port, batch, tmp), nprocs=2, join=True)
    got = np.load(tmp)
    arrs = synth.box_qp_batch_numpy(8, 2, batch, seed=99, explicit_constraints=False)
    ref, steps = oracle.qp_solve("oracle", arrs, synth.solve_params())
    assert np.array_equal(got["vars"], ref.vars) and np.array_equal(got["status"], ref.status)
    assert np.array_equal(got["iters"], ref.iters) and int(got["total"][0]) == steps
