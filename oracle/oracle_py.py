"""TEST INFRASTRUCTURE — loader for the CPU checkers. Never imported by the product package.

* ``liboracle.so``            plain-C restatement of the reference algorithm (oracle/*.c)
* ``_ref/libminiopt_ref.so``  the unmodified reference compiled from /root/reference + extern "C" harness

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs
may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

from mini_opt_b200 import abi

HERE = Path(__file__).resolve().parent
ORACLE_SO = HERE / "liboracle.so"
REF_SO = HERE / "_ref" / "libminiopt_ref.so"
REFERENCE_ROOT = Path(os.environ.get("MINI_OPT_REFERENCE", "/root/reference"))


def build(verbose: bool = False) -> None:
    """Compile the C restatement and, when the reference sources are present, oracle/_ref."""
    cmd = ["make", "-C", str(HERE), f"REF={REFERENCE_ROOT}", "-j8", "all"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:], res.stderr[-4000:])
    if res.returncode != 0:
        raise RuntimeError("oracle build failed")


_oracle = None
_ref = None


def _bind_qp(lib, prefix: str, with_threads: bool):
    solve = getattr(lib, f"{prefix}_qp_solve_batch")
    solve.restype = C.c_int64
    solve.argtypes = [C.POINTER(abi.MoQpBatch), C.POINTER(abi.MoQpParams), C.POINTER(abi.MoQpOut)] + (
        [C.c_int] if with_threads else [])
    step = getattr(lib, f"{prefix}_qp_newton_step_batch")
    step.restype = C.c_int
    step.argtypes = [C.POINTER(abi.MoQpBatch), C.c_double, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                     C.c_void_p, C.c_void_p]


def _bind_nls(lib, prefix: str, with_threads: bool):
    lin = getattr(lib, f"{prefix}_nls_linearize_batch")
    lin.restype = C.c_int
    lin.argtypes = [C.POINTER(abi.MoNlsLinearizeArgs)]
    fk = getattr(lib, f"{prefix}_chain_fk_batch")
    fk.restype = C.c_int
    fk.argtypes = [C.POINTER(abi.MoChainIkBatch)] + [C.c_void_p] * 5
    solve = getattr(lib, f"{prefix}_nls_solve_chain_ik_batch")
    solve.restype = C.c_int64
    solve.argtypes = [C.POINTER(abi.MoChainIkBatch), C.POINTER(abi.MoNlsParams), C.POINTER(abi.MoNlsOut)] + (
        [C.c_int] if with_threads else [])


def oracle_lib():
    global _oracle
    if _oracle is None:
        if not ORACLE_SO.exists():
            build()
        _oracle = C.CDLL(str(ORACLE_SO))
        _bind_qp(_oracle, "oracle", with_threads=False)
        _bind_nls(_oracle, "oracle", with_threads=False)
    return _oracle


def ref_available() -> bool:
    return REF_SO.exists() or (REFERENCE_ROOT / "source" / "qp.cc").exists()


def ref_lib():
    global _ref
    if _ref is None:
        if not REF_SO.exists():
            build()
        _ref = C.CDLL(str(REF_SO))
        _bind_qp(_ref, "ref", with_threads=True)
        _bind_nls(_ref, "ref", with_threads=True)
        _ref.ref_qp_solve_batch_timed.restype = C.c_int64
        _ref.ref_qp_solve_batch_timed.argtypes = [C.POINTER(abi.MoQpBatch), C.POINTER(abi.MoQpParams),
                                                  C.POINTER(abi.MoQpOut), C.c_int, C.POINTER(C.c_double)]
        _ref.ref_generate_random_qp.restype = C.c_int
        _ref.ref_generate_random_qp.argtypes = [C.c_int, C.c_int, C.c_double] + [C.c_void_p] * 6
        _ref.ref_hardware_concurrency.restype = C.c_int
    return _ref


def _check_host(arrs: abi.QPBatchArrays):
    for name in ("G", "c", "A_eq", "b_eq", "ineq_var", "ineq_a", "ineq_b", "ineq_count"):
        a = getattr(arrs, name)
        if a is not None:
            assert isinstance(a, np.ndarray) and a.flags["C_CONTIGUOUS"], name


def qp_solve(which: str, arrs: abi.QPBatchArrays, params: abi.QPParams, log: bool = False, nthreads: int = 1,
             user_vars: np.ndarray | None = None):
    """Solve a HOST batch with the C restatement (``which='oracle'``) or the reference (``'ref'``).

    Returns (QPOutArrays of numpy arrays, total Newton steps).
    """
    _check_host(arrs)
    out = abi.alloc_out_numpy(arrs.batch, arrs.n, arrs.k, arrs.m, params.max_iterations, log)
    if user_vars is not None:
        out.vars[...] = user_vars
    b, p, o = arrs.to_c(), params.to_c(), out.to_c()
    if which == "oracle":
        total = oracle_lib().oracle_qp_solve_batch(C.byref(b), C.byref(p), C.byref(o))
    elif which == "ref":
        total = ref_lib().ref_qp_solve_batch(C.byref(b), C.byref(p), C.byref(o), int(nthreads))
    else:
        raise ValueError(which)
    if total < 0:
        raise RuntimeError(f"{which}_qp_solve_batch failed: {total}")
    return out, int(total)


def qp_solve_timed(which: str, arrs: abi.QPBatchArrays, params: abi.QPParams, nthreads: int = 1):
    """Like ``qp_solve`` but returns (outputs, total Newton steps, seconds) where seconds covers the solve loop only:
    for the reference the problems are unpacked into ``mini_opt::QP`` objects and the outputs allocated before the
    clock starts (``ref_qp_solve_batch_timed``); the scalar C port is timed around its call."""
    import time
    _check_host(arrs)
    out = abi.alloc_out_numpy(arrs.batch, arrs.n, arrs.k, arrs.m, params.max_iterations, False)
    b, p, o = arrs.to_c(), params.to_c(), out.to_c()
    if which == "ref":
        sec = C.c_double(0.0)
        total = ref_lib().ref_qp_solve_batch_timed(C.byref(b), C.byref(p), C.byref(o), int(nthreads), C.byref(sec))
        seconds = float(sec.value)
    else:
        t0 = time.perf_counter()
        total = oracle_lib().oracle_qp_solve_batch(C.byref(b), C.byref(p), C.byref(o))
        seconds = time.perf_counter() - t0
    if total < 0:
        raise RuntimeError(f"{which} timed solve failed: {total}")
    return out, int(total), seconds


def qp_newton_step(which: str, arrs: abi.QPBatchArrays, mu: float, strategy: int, vars_in: np.ndarray):
    """One ``Iterate(mu, strategy)`` per problem. Returns dict(vars, delta, ip, kkt, status)."""
    _check_host(arrs)
    B, stride = arrs.batch, abi.vars_stride(arrs.n, arrs.k, arrs.m)
    v = np.ascontiguousarray(vars_in, dtype=np.float64).reshape(B, stride).copy()
    delta = np.zeros((B, stride))
    ip = np.zeros((B, 6))
    kkt = np.zeros((B, 4))
    status = np.full(B, -1, dtype=np.int32)
    b = arrs.to_c()
    fn = oracle_lib().oracle_qp_newton_step_batch if which == "oracle" else ref_lib().ref_qp_newton_step_batch
    rc = fn(C.byref(b), float(mu), int(strategy), v.ctypes.data, delta.ctypes.data, ip.ctypes.data,
            kkt.ctypes.data, status.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"{which} newton step failed: {rc}")
    return dict(vars=v, delta=delta, ip=ip, kkt=kkt, status=status)


def qp_nullspace(which: str, arrs: abi.QPBatchArrays):
    """``QPNullSpaceSolver::Solve`` (qp.cc:679-729) per problem of a host batch: (x [B, n], status [B])."""
    _check_host(arrs)
    x = np.zeros((arrs.batch, arrs.n))
    status = np.full(arrs.batch, -1, dtype=np.int32)
    lib = oracle_lib() if which == "oracle" else ref_lib()
    fn = getattr(lib, f"{which}_qp_nullspace_batch")
    fn.restype = C.c_int
    fn.argtypes = [C.POINTER(abi.MoQpBatch), C.c_void_p, C.c_void_p]
    b = arrs.to_c()
    rc = fn(C.byref(b), x.ctypes.data, status.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"{which}_qp_nullspace_batch failed: {rc}")
    return x, status


def generate_random_qp(seed: int, dim: int = 8, p_inequality: float = 0.5):
    """The reference's own ``GenerateRandomQP`` (test/qp_test.cc:483-524). Needs oracle/_ref."""
    lib = ref_lib()
    G = np.zeros((dim, dim))
    c = np.zeros(dim)
    var = np.zeros(dim, dtype=np.int32)
    a = np.zeros(dim)
    b = np.zeros(dim)
    sol = np.zeros(dim)
    cnt = lib.ref_generate_random_qp(int(seed), int(dim), float(p_inequality), G.ctypes.data, c.ctypes.data,
                                     var.ctypes.data, a.ctypes.data, b.ctypes.data, sol.ctypes.data)
    return dict(G=G, c=c, var=var[:cnt].copy(), a=a[:cnt].copy(), b=b[:cnt].copy(), solution=sol, count=cnt)


def _lib(which: str):
    return oracle_lib() if which == "oracle" else ref_lib()


def nls_linearize(which: str, a: abi.LinearizeArrays) -> abi.LinearizeOut:
    """``LinearizeAndFillQP`` (nonlinear.cc:170-214) on dense per-problem Jacobians (host arrays)."""
    out = abi.alloc_linearize_out_numpy(a)
    args = abi.linearize_to_c(a, out)
    rc = getattr(_lib(which), f"{which}_nls_linearize_batch")(C.byref(args))
    if rc != 0:
        raise RuntimeError(f"{which}_nls_linearize_batch failed: {rc}")
    return out


def chain_fk(which: str, chain: abi.ChainIKArrays, params: np.ndarray):
    """``ActuatorChain::Update`` per row of ``params``: dict(translation, rotation_q, t_D_params, r_D_params)."""
    B, n = params.shape
    params = np.ascontiguousarray(params, dtype=np.float64)
    t = np.zeros((B, 3))
    q = np.zeros((B, 4))
    tD = np.zeros((B, n, 3))
    rD = np.zeros((B, n, 3))
    c = chain.to_c()
    c.batch = B
    rc = getattr(_lib(which), f"{which}_chain_fk_batch")(C.byref(c), params.ctypes.data, t.ctypes.data, q.ctypes.data,
                                                         tD.ctypes.data, rD.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"{which}_chain_fk_batch failed: {rc}")
    return dict(translation=t, rotation_q=q, t_D_params=tD, r_D_params=rD)


def nls_solve_chain(which: str, chain: abi.ChainIKArrays, params: abi.NLSParams, nthreads: int = 1):
    """``ConstrainedNonlinearLeastSquares::Solve`` per problem. Returns (NLSOutArrays, total QP iterations)."""
    out = abi.alloc_nls_out_numpy(chain.batch, chain.n)
    c, p, o = chain.to_c(), params.to_c(), out.to_c()
    if which == "oracle":
        total = oracle_lib().oracle_nls_solve_chain_ik_batch(C.byref(c), C.byref(p), C.byref(o))
    else:
        total = ref_lib().ref_nls_solve_chain_ik_batch(C.byref(c), C.byref(p), C.byref(o), int(nthreads))
    if total < 0:
        raise RuntimeError(f"{which}_nls_solve_chain_ik_batch failed: {total}")
    return out, int(total)
