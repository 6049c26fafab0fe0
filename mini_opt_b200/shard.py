"""Multi-GPU sharding of a batch: problems are independent, so rank g of G owns the contiguous slice
[g*B/G, (g+1)*B/G) of every batch array, solves it on its own GPU with no collective on the solve path, and the
results are gathered once at the end. ``torch.distributed`` (nccl on GPUs, gloo in the CPU tests) is plumbing only."""
from __future__ import annotations

import dataclasses

from . import abi


def shard_bounds(batch: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced partition: sizes differ by at most one, order preserved."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return (batch * rank) // world, (batch * (rank + 1)) // world


def shard_qp_batch(arrs: abi.QPBatchArrays, rank: int, world: int) -> abi.QPBatchArrays:
    """This rank's slice of a QP batch (views, no copies)."""
    lo, hi = shard_bounds(arrs.batch, rank, world)
    kw = {f.name: getattr(arrs, f.name) for f in dataclasses.fields(arrs)}
    for name in ("G", "c", "A_eq", "b_eq", "ineq_var", "ineq_a", "ineq_b", "ineq_count"):
        if kw[name] is not None:
            kw[name] = kw[name][lo:hi]
    kw["batch"] = hi - lo
    return abi.QPBatchArrays(**kw)


def gather_rows(local, batch: int, group=None):
    """All-gather the per-problem rows of one output tensor (torch, any device) back into batch order."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    sizes = [shard_bounds(batch, r, world)[1] - shard_bounds(batch, r, world)[0] for r in range(world)]
    cap = max(sizes)
    pad = torch.zeros((cap,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[:s] for p, s in zip(parts, sizes)], dim=0)


def gather_qp_out(out: abi.QPOutArrays, batch: int, group=None) -> abi.QPOutArrays:
    """The "gather at the end": every field of the per-rank outputs, in the original problem order."""
    kw = {}
    for f in dataclasses.fields(out):
        v = getattr(out, f.name)
        kw[f.name] = None if v is None else gather_rows(v, batch, group)
    return abi.QPOutArrays(**kw)
