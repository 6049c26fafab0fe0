"""Rebuild QPBatchArrays from the golden .npz files (tests/golden/, made by tools/make_golden.py)."""
import numpy as np

from mini_opt_b200 import abi


def generated_batch(g) -> abi.QPBatchArrays:
    B, n = g["c"].shape
    return abi.QPBatchArrays(n=n, k=0, m=n, batch=B, G=np.ascontiguousarray(g["G"]), c=np.ascontiguousarray(g["c"]),
                             ineq_var=np.ascontiguousarray(g["ineq_var"]), ineq_a=np.ascontiguousarray(g["ineq_a"]),
                             ineq_b=np.ascontiguousarray(g["ineq_b"]), ineq_count=np.ascontiguousarray(g["ineq_count"]))


def box_batch(g, n: int, explicit: bool = False) -> abi.QPBatchArrays:
    t = f"n{n}"
    B = g[f"{t}_c"].shape[0]
    k = g[f"{t}_b_eq"].shape[1]
    arrs = abi.QPBatchArrays(n=n, k=k, m=2 * n, batch=B, G=np.ascontiguousarray(g[f"{t}_G"]),
                             c=np.ascontiguousarray(g[f"{t}_c"]), A_eq=np.ascontiguousarray(g[f"{t}_A_eq"]),
                             b_eq=np.ascontiguousarray(g[f"{t}_b_eq"]), ineq_b=np.ascontiguousarray(g[f"{t}_ineq_b"]),
                             flags=abi.MO_QP_FLAG_BOX_PAIRS)
    if explicit:
        arrs.flags = 0
        arrs.ineq_var = np.ascontiguousarray(np.broadcast_to(np.repeat(np.arange(n, dtype=np.int32), 2), (B, 2 * n)))
        arrs.ineq_a = np.ascontiguousarray(np.broadcast_to(np.tile([1.0, -1.0], n), (B, 2 * n)))
    return arrs


def rel_x_err(vars_a, vars_b, n):
    xa, xb = np.asarray(vars_a)[:, :n], np.asarray(vars_b)[:, :n]
    return np.abs(xa - xb).max(axis=1) / np.maximum(1.0, np.abs(xb).max(axis=1))
