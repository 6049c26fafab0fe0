"""ctypes mirror of ``include/mini_opt_b200.h`` (struct layouts, enums, argument packing).

Plain data only: nothing here touches CUDA. ``mini_opt_b200.lib`` binds the product library with these
types; the test-only oracle loader (``oracle/oracle_py.py``) reuses them so that the reference, the C
restatement and the CUDA path all consume identical byte buffers.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

MO_ABI_VERSION = 1

# mo_error
MO_OK = 0
MO_ERR_INVALID_ARGUMENT = -1
MO_ERR_INVALID_PARAMS = -2
MO_ERR_UNSUPPORTED_SHAPE = -3
MO_ERR_CUDA = -4
MO_ERR_NO_DEVICE = -5

# per-problem status (QPInteriorPointTerminationState, structs.hpp:97-102, + batched failure codes)
MO_QP_SATISFIED_KKT_TOL = 0
MO_QP_MAX_ITERATIONS = 1
MO_QP_FAILED_FACTORIZATION = 2
MO_QP_NONPOSITIVE_SLACK = 3

# BarrierStrategy (structs.hpp:24-31)
MO_BARRIER_COMPLEMENTARITY = 0
MO_BARRIER_FIXED_DECREASE = 1
MO_BARRIER_PREDICTOR_CORRECTOR = 2

# InitialGuessMethod (structs.hpp:34-41)
MO_GUESS_NAIVE = 0
MO_GUESS_SOLVE_EQUALITY_CONSTRAINED = 1
MO_GUESS_USER_PROVIDED = 2

MO_QP_FLAG_BOX_PAIRS = 1
MO_QP_ITER_LOG_STRIDE = 14

_dp = C.c_void_p  # data pointers are passed as raw addresses (host or device)


class MoQpBatch(C.Structure):
    _fields_ = [
        ("n", C.c_int32),
        ("k", C.c_int32),
        ("m", C.c_int32),
        ("flags", C.c_int32),
        ("batch", C.c_int64),
        ("G", _dp),
        ("c", _dp),
        ("A_eq", _dp),
        ("b_eq", _dp),
        ("ineq_var", _dp),
        ("ineq_a", _dp),
        ("ineq_b", _dp),
        ("ineq_count", _dp),
    ]


class MoQpParams(C.Structure):
    _fields_ = [
        ("initial_mu", C.c_double),
        ("sigma", C.c_double),
        ("termination_kkt_tol", C.c_double),
        ("termination_complementarity_tol", C.c_double),
        ("max_iterations", C.c_int32),
        ("barrier_strategy", C.c_int32),
        ("decrease_mu_only_on_small_error", C.c_int32),
        ("initial_guess_method", C.c_int32),
        ("initialize_mu_with_complementarity", C.c_int32),
        ("reserved", C.c_int32),
    ]


class MoQpOut(C.Structure):
    _fields_ = [
        ("vars", _dp),
        ("status", _dp),
        ("iters", _dp),
        ("kkt_final", _dp),
        ("lagrange", _dp),
        ("iter_log", _dp),
    ]


@dataclass
class QPParams:
    """``QPInteriorPointSolver::Params`` (include/mini_opt/qp.hpp:134-164), same names and defaults."""

    initial_mu: float = 1.0
    sigma: float = 0.5
    termination_kkt_tol: float = 1.0e-9
    termination_complementarity_tol: float = 1.0e-6
    max_iterations: int = 10
    barrier_strategy: int = MO_BARRIER_COMPLEMENTARITY
    decrease_mu_only_on_small_error: bool = False
    initial_guess_method: int = MO_GUESS_NAIVE
    initialize_mu_with_complementarity: bool = False

    def to_c(self) -> MoQpParams:
        return MoQpParams(
            self.initial_mu,
            self.sigma,
            self.termination_kkt_tol,
            self.termination_complementarity_tol,
            int(self.max_iterations),
            int(self.barrier_strategy),
            int(bool(self.decrease_mu_only_on_small_error)),
            int(self.initial_guess_method),
            int(bool(self.initialize_mu_with_complementarity)),
            0,
        )


def vars_stride(n: int, k: int, m: int) -> int:
    return n + 2 * m + k


def _addr(a) -> int | None:
    """Address of a numpy array (host) or of anything exposing ``data_ptr()`` (torch tensors)."""
    if a is None:
        return None
    if hasattr(a, "data_ptr"):
        return int(a.data_ptr())
    return int(a.ctypes.data)


@dataclass
class QPBatchArrays:
    """A uniform-shape batch of QPs as flat arrays (numpy on the host or torch tensors on a device).

    Shapes: G [B,n,n] holding each problem COLUMN-major (i.e. ``G[p].T`` is the row-major matrix; for the
    symmetric G this only matters for which triangle is "lower"), c [B,n], A_eq [B,n,k] (column-major
    k x n == C-order [n][k]), b_eq [B,k], ineq_var [B,m] int32, ineq_a/ineq_b [B,m], ineq_count [B] int32.
    """

    n: int
    k: int
    m: int
    batch: int
    G: object
    c: object
    A_eq: object = None
    b_eq: object = None
    ineq_var: object = None
    ineq_a: object = None
    ineq_b: object = None
    ineq_count: object = None
    flags: int = 0

    def to_c(self) -> MoQpBatch:
        return MoQpBatch(
            self.n, self.k, self.m, self.flags, self.batch,
            _addr(self.G), _addr(self.c), _addr(self.A_eq), _addr(self.b_eq),
            _addr(self.ineq_var), _addr(self.ineq_a), _addr(self.ineq_b), _addr(self.ineq_count),
        )


@dataclass
class QPOutArrays:
    vars: object
    status: object
    iters: object
    kkt_final: object = None
    lagrange: object = None
    iter_log: object = None

    def to_c(self) -> MoQpOut:
        return MoQpOut(_addr(self.vars), _addr(self.status), _addr(self.iters), _addr(self.kkt_final),
                       _addr(self.lagrange), _addr(self.iter_log))


def alloc_out_numpy(batch: int, n: int, k: int, m: int, max_iterations: int = 0, log: bool = False) -> QPOutArrays:
    return QPOutArrays(
        vars=np.zeros((batch, vars_stride(n, k, m))),
        status=np.full(batch, -1, dtype=np.int32),
        iters=np.zeros(batch, dtype=np.int32),
        kkt_final=np.zeros((batch, 4)),
        lagrange=np.zeros((batch, 2)),
        iter_log=np.full((batch, max_iterations, MO_QP_ITER_LOG_STRIDE), np.nan) if log else None,
    )
