"""Harness-side callers of the C ABI (tests, bench). The product host API is the C++ header
``include/mini_opt_b200/qp.hpp``; this module only packs numpy / torch buffers into the C structs.

torch is used for device memory and streams only.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import abi, lib


def to_device(arrs: abi.QPBatchArrays, device) -> abi.QPBatchArrays:
    """Copy a host (numpy) batch to ``device`` as torch tensors."""
    import torch

    def conv(a):
        return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to(device)

    return abi.QPBatchArrays(n=arrs.n, k=arrs.k, m=arrs.m, batch=arrs.batch, G=conv(arrs.G), c=conv(arrs.c),
                             A_eq=conv(arrs.A_eq), b_eq=conv(arrs.b_eq), ineq_var=conv(arrs.ineq_var),
                             ineq_a=conv(arrs.ineq_a), ineq_b=conv(arrs.ineq_b), ineq_count=conv(arrs.ineq_count),
                             flags=arrs.flags)


def alloc_out_torch(batch, n, k, m, device, max_iterations=0, log=False, kkt=True, lagrange=True) -> abi.QPOutArrays:
    import torch

    f64 = dict(dtype=torch.float64, device=device)
    return abi.QPOutArrays(
        vars=torch.zeros((batch, abi.vars_stride(n, k, m)), **f64),
        status=torch.full((batch,), -1, dtype=torch.int32, device=device),
        iters=torch.zeros((batch,), dtype=torch.int32, device=device),
        kkt_final=torch.zeros((batch, 4), **f64) if kkt else None,
        lagrange=torch.zeros((batch, 2), **f64) if lagrange else None,
        iter_log=torch.full((batch, max_iterations, abi.MO_QP_ITER_LOG_STRIDE), float("nan"), **f64) if log else None,
    )


def out_to_numpy(out: abi.QPOutArrays) -> abi.QPOutArrays:
    def conv(t):
        return None if t is None else t.cpu().numpy()

    return abi.QPOutArrays(vars=conv(out.vars), status=conv(out.status), iters=conv(out.iters),
                           kkt_final=conv(out.kkt_final), lagrange=conv(out.lagrange), iter_log=conv(out.iter_log))


def _device_index(device) -> int:
    import torch

    d = torch.device(device)
    return d.index if d.index is not None else torch.cuda.current_device()


def solve_batch_device(arrs: abi.QPBatchArrays, params: abi.QPParams, out: abi.QPOutArrays, device, stream=None):
    """``mo_qp_ip_solve_batch`` on device-resident torch tensors; asynchronous on ``stream`` (torch current stream)."""
    import torch

    L = lib.load()
    st = stream if stream is not None else torch.cuda.current_stream(device)
    b, p, o = arrs.to_c(), params.to_c(), out.to_c()
    lib.check(L.mo_qp_ip_solve_batch(C.byref(b), C.byref(p), C.byref(o), _device_index(device), st.cuda_stream))


def solve_batch_host(arrs: abi.QPBatchArrays, params: abi.QPParams, out: abi.QPOutArrays, device_index: int = 0):
    """``mo_qp_ip_solve_batch_host`` on host (numpy) buffers; returns when the results are in ``out``."""
    L = lib.load()
    b, p, o = arrs.to_c(), params.to_c(), out.to_c()
    lib.check(L.mo_qp_ip_solve_batch_host(C.byref(b), C.byref(p), C.byref(o), int(device_index)))


def nullspace_solve_device(arrs: abi.QPBatchArrays, device):
    """``mo_qp_nullspace_solve_batch`` on device tensors: (x [B, n], status [B]) torch tensors."""
    import torch

    L = lib.load()
    x = torch.zeros((arrs.batch, arrs.n), dtype=torch.float64, device=device)
    status = torch.full((arrs.batch,), -1, dtype=torch.int32, device=device)
    b = arrs.to_c()
    st = torch.cuda.current_stream(device)
    lib.check(L.mo_qp_nullspace_solve_batch(C.byref(b), x.data_ptr(), status.data_ptr(), _device_index(device), st.cuda_stream))
    return x, status


def solve_batch_host_multi(arrs: abi.QPBatchArrays, params: abi.QPParams, out: abi.QPOutArrays, devices):
    """``mo_qp_ip_solve_batch_host_multi``: one host batch sharded over several devices of this node from one process."""
    L = lib.load()
    b, p, o = arrs.to_c(), params.to_c(), out.to_c()
    devs = (C.c_int * len(devices))(*[int(d) for d in devices])
    lib.check(L.mo_qp_ip_solve_batch_host_multi(C.byref(b), C.byref(p), C.byref(o), devs, len(devices)))


def pack_lower(G: np.ndarray) -> np.ndarray:
    """[B, n, n] column-major problems -> [B, n(n+1)/2] packed lower triangles (MO_QP_FLAG_PACKED_G: column j holds rows
    j..n-1). ``G[p][c][r]`` is element (r, c)."""
    B, n, _ = G.shape
    cols = [G[:, j, j:] for j in range(n)]
    return np.ascontiguousarray(np.concatenate(cols, axis=1))


def newton_step_device(arrs: abi.QPBatchArrays, mu: float, strategy: int, vars_t, device, want=("delta", "ip", "kkt")):
    """``mo_qp_ip_newton_step_batch`` on device tensors; ``vars_t`` [B, n+2m+k] is updated in place."""
    import torch

    L = lib.load()
    B = arrs.batch
    f64 = dict(dtype=torch.float64, device=device)
    delta = torch.zeros_like(vars_t) if "delta" in want else None
    ip = torch.zeros((B, 6), **f64) if "ip" in want else None
    kkt = torch.zeros((B, 4), **f64) if "kkt" in want else None
    status = torch.full((B,), -1, dtype=torch.int32, device=device)
    b = arrs.to_c()
    ptr = lambda t: None if t is None else t.data_ptr()
    st = torch.cuda.current_stream(device)
    lib.check(L.mo_qp_ip_newton_step_batch(C.byref(b), float(mu), int(strategy), ptr(vars_t), ptr(delta), ptr(ip),
                                           ptr(kkt), ptr(status), _device_index(device), st.cuda_stream))
    return dict(vars=vars_t, delta=delta, ip=ip, kkt=kkt, status=status)
