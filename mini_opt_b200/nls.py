"""Harness-side callers of the nonlinear least squares entry points of the C ABI (tests, bench): packs torch
device tensors into ``mo_chain_ik_batch`` / ``mo_nls_linearize_args``. torch = device memory and streams only."""
from __future__ import annotations

import ctypes as C
import dataclasses

import numpy as np

from . import abi, lib
from .qp import _device_index


def _to_dev(a, device):
    import torch
    return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to(device)


def chain_to_device(ch: abi.ChainIKArrays, device) -> abi.ChainIKArrays:
    kw = {f.name: getattr(ch, f.name) for f in dataclasses.fields(ch)}
    for name in ("link_pose", "link_mask", "cost_w", "cost_target", "eq_w", "eq_target", "ineq_var", "ineq_a",
                 "ineq_b", "x0"):
        kw[name] = _to_dev(kw[name], device)
    return abi.ChainIKArrays(**kw)


def alloc_nls_out_torch(batch: int, n: int, device) -> abi.NLSOutArrays:
    import torch
    i32 = dict(dtype=torch.int32, device=device)
    return abi.NLSOutArrays(x=torch.zeros((batch, n), dtype=torch.float64, device=device),
                            termination=torch.full((batch,), -1, **i32), iterations=torch.zeros((batch,), **i32),
                            qp_iterations=torch.zeros((batch,), **i32), ls_steps=torch.zeros((batch,), **i32),
                            errors=torch.zeros((batch, 2), dtype=torch.float64, device=device))


def nls_out_to_numpy(o: abi.NLSOutArrays) -> abi.NLSOutArrays:
    return abi.NLSOutArrays(**{f.name: getattr(o, f.name).cpu().numpy() for f in dataclasses.fields(o)})


def solve_chain_ik_device(ch: abi.ChainIKArrays, params: abi.NLSParams, out: abi.NLSOutArrays, device, stream=None):
    """``mo_nls_solve_chain_ik_batch`` on device tensors; asynchronous on ``stream``."""
    import torch
    st = stream if stream is not None else torch.cuda.current_stream(device)
    c, p, o = ch.to_c(), params.to_c(), out.to_c()
    lib.check(lib.load().mo_nls_solve_chain_ik_batch(C.byref(c), C.byref(p), C.byref(o), _device_index(device),
                                                     st.cuda_stream))


def family_dims(family: int):
    """(n, cost rows, equality rows) of a functor residual family (``mo_nls_family_dims``)."""
    n, rc, re = C.c_int32(), C.c_int32(), C.c_int32()
    lib.check(lib.load().mo_nls_family_dims(family, C.byref(n), C.byref(rc), C.byref(re)))
    return n.value, rc.value, re.value


def solve_family_device(family: int, x0, params: abi.NLSParams, out: abi.NLSOutArrays, device, ineq=None, data=None,
                        stream=None):
    """``mo_nls_solve_family_batch`` on device tensors: ``x0`` [batch, n]; ``ineq`` = (var int32 [m], a [m], b [m]) device
    tensors shared by the batch; ``data`` optional [batch, stride] per-problem constants of the functor."""
    import torch
    st = stream if stream is not None else torch.cuda.current_stream(device)
    var, a, b = ineq if ineq is not None else (None, None, None)
    c = abi.MoNlsFamilyBatch(family, 0 if var is None else int(var.numel()), int(x0.shape[0]), abi._addr(var), abi._addr(a),
                             abi._addr(b), abi._addr(x0), abi._addr(data), 0 if data is None else int(data.shape[1]), 0)
    p, o = params.to_c(), out.to_c()
    lib.check(lib.load().mo_nls_solve_family_batch(C.byref(c), C.byref(p), C.byref(o), _device_index(device), st.cuda_stream))


_CHAIN_ARRAYS = ("link_pose", "link_mask", "cost_w", "cost_target", "eq_w", "eq_target", "ineq_var", "ineq_a", "ineq_b", "x0")


class ChainIKHostPipeline:
    """Host-buffer path of the chain-IK solve for the harness (bench ``e2e``): the problem arrays live in pinned host
    memory, every call copies them to preallocated device buffers, runs ``mo_nls_solve_chain_ik_batch`` and copies the
    results back to pinned host memory, all on one stream. ``h2d_bytes`` / ``d2h_bytes`` are counted from the tensors."""

    def __init__(self, ch: abi.ChainIKArrays, device):
        import torch
        self.device = device
        kw = {f.name: getattr(ch, f.name) for f in dataclasses.fields(ch)}
        self.host, self.dev = {}, {}
        for name in _CHAIN_ARRAYS:
            if kw[name] is not None:
                self.host[name] = torch.from_numpy(np.ascontiguousarray(kw[name])).pin_memory()
                self.dev[name] = torch.empty_like(self.host[name], device=device)
        self.ch = abi.ChainIKArrays(**{**kw, **{k: self.dev.get(k) for k in _CHAIN_ARRAYS}})
        self.out = alloc_nls_out_torch(ch.batch, ch.n, device)
        self.host_out = {f.name: torch.empty_like(getattr(self.out, f.name), device="cpu").pin_memory()
                         for f in dataclasses.fields(self.out)}
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in self.host.values())
        self.d2h_bytes = sum(t.numel() * t.element_size() for t in self.host_out.values())

    def solve(self, params: abi.NLSParams):
        import torch
        st = torch.cuda.current_stream(self.device)
        for name, h in self.host.items():
            self.dev[name].copy_(h, non_blocking=True)
        solve_chain_ik_device(self.ch, params, self.out, self.device, st)
        for name, h in self.host_out.items():
            h.copy_(getattr(self.out, name), non_blocking=True)
        st.synchronize()
        return self.host_out


def chain_fk_device(ch: abi.ChainIKArrays, params_t, device):
    """``mo_chain_fk_batch``: dict(translation, rotation_q, t_D_params, r_D_params) of device tensors."""
    import torch
    B, n = params_t.shape
    f64 = dict(dtype=torch.float64, device=device)
    t, q = torch.zeros((B, 3), **f64), torch.zeros((B, 4), **f64)
    tD, rD = torch.zeros((B, n, 3), **f64), torch.zeros((B, n, 3), **f64)
    c = ch.to_c()
    c.batch = B
    st = torch.cuda.current_stream(device)
    lib.check(lib.load().mo_chain_fk_batch(C.byref(c), params_t.data_ptr(), t.data_ptr(), q.data_ptr(), tD.data_ptr(),
                                           rD.data_ptr(), _device_index(device), st.cuda_stream))
    return dict(translation=t, rotation_q=q, t_D_params=tD, r_D_params=rD)


def linearize_device(a: abi.LinearizeArrays, device) -> abi.LinearizeOut:
    """``mo_nls_linearize_batch`` on host arrays copied to ``device``; returns device tensors."""
    import torch
    kw = {f.name: getattr(a, f.name) for f in dataclasses.fields(a)}
    for name in ("x", "J_cost", "r_cost_v", "J_eq", "r_eq_v", "lam", "ineq_var", "ineq_a", "ineq_b"):
        kw[name] = _to_dev(kw[name], device)
    d = abi.LinearizeArrays(**kw)
    B, n, k, m = a.batch, a.n, a.r_eq, a.m
    f64 = dict(dtype=torch.float64, device=device)
    o = abi.LinearizeOut(G=torch.full((B, n, n), float("nan"), **f64), c=torch.zeros((B, n), **f64),
                         A_eq=torch.zeros((B, n, k), **f64), b_eq=torch.zeros((B, k), **f64),
                         ineq_b_shifted=torch.zeros((B, m), **f64), errors=torch.zeros((B, 2), **f64))
    args = abi.linearize_to_c(d, o)
    st = torch.cuda.current_stream(device)
    lib.check(lib.load().mo_nls_linearize_batch(C.byref(args), _device_index(device), st.cuda_stream))
    torch.cuda.synchronize()
    return o
