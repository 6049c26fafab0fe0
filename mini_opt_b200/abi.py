"""ctypes mirror of ``include/mini_opt_b200.h`` (struct layouts, enums, argument packing).

Plain data only: nothing here touches CUDA. ``mini_opt_b200.lib`` binds the product library with these
types; the test-only oracle loader (``oracle/oracle_py.py``) reuses them so that the reference, the C
restatement and the CUDA path all consume identical byte buffers.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

MO_ABI_VERSION = 2

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
MO_QP_FLAG_PACKED_G = 2
MO_QP_FLAG_UNIFORM_BOX = 4
MO_QP_ITER_LOG_STRIDE = 14
MO_NOALLOC_MAX_N = 128

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
        ("x", _dp),
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
    x: object = None

    def to_c(self) -> MoQpOut:
        return MoQpOut(_addr(self.vars), _addr(self.status), _addr(self.iters), _addr(self.kkt_final),
                       _addr(self.lagrange), _addr(self.iter_log), _addr(self.x))


def alloc_out_numpy(batch: int, n: int, k: int, m: int, max_iterations: int = 0, log: bool = False) -> QPOutArrays:
    return QPOutArrays(
        vars=np.zeros((batch, vars_stride(n, k, m))),
        status=np.full(batch, -1, dtype=np.int32),
        iters=np.zeros(batch, dtype=np.int32),
        kkt_final=np.zeros((batch, 4)),
        lagrange=np.zeros((batch, 2)),
        iter_log=np.full((batch, max_iterations, MO_QP_ITER_LOG_STRIDE), np.nan) if log else None,
    )


# ---------------------------------------------------------------------------------------------------------
# Nonlinear least squares (ConstrainedNonlinearLeastSquares, nonlinear.hpp / nonlinear.cc)
# ---------------------------------------------------------------------------------------------------------
MO_LS_ARMIJO_BACKTRACK = 0
MO_LS_POLYNOMIAL_APPROXIMATION = 1

MO_NLS_MAX_ITERATIONS = 0
MO_NLS_SATISFIED_ABSOLUTE_TOL = 1
MO_NLS_SATISFIED_RELATIVE_TOL = 2
MO_NLS_SATISFIED_FIRST_ORDER_TOL = 3
MO_NLS_MAX_LAMBDA = 4
MO_NLS_QP_INDEFINITE = 5
MO_NLS_USER_CALLBACK = 6
MO_NLS_QP_FAILED = 7

MO_RETRACT_ADD = 0
MO_RETRACT_MOD_PI = 1


class MoNlsParams(C.Structure):
    _fields_ = [
        ("max_iterations", C.c_int32),
        ("max_qp_iterations", C.c_int32),
        ("termination_kkt_tolerance", C.c_double),
        ("absolute_exit_tol", C.c_double),
        ("relative_exit_tol", C.c_double),
        ("absolute_first_derivative_tol", C.c_double),
        ("max_line_search_iterations", C.c_int32),
        ("line_search_strategy", C.c_int32),
        ("armijo_search_tau", C.c_double),
        ("equality_penalty_initial", C.c_double),
        ("equality_penalty_scale_factor", C.c_double),
        ("equality_penalty_rho", C.c_double),
        ("lambda_initial", C.c_double),
        ("lambda_failure_init", C.c_double),
        ("lambda_decrease_on_success", C.c_double),
        ("lambda_decrease_on_restore", C.c_double),
        ("max_lambda", C.c_double),
        ("min_lambda", C.c_double),
        ("retraction", C.c_int32),
        ("inner_qp_solver", C.c_int32),
    ]


@dataclass
class NLSParams:
    """``ConstrainedNonlinearLeastSquares::Params`` (include/mini_opt/nonlinear.hpp:37-124), same defaults."""

    max_iterations: int = 10
    max_qp_iterations: int = 10
    termination_kkt_tolerance: float = 1.0e-6
    absolute_exit_tol: float = 1.0e-12
    relative_exit_tol: float = 1.0e-5
    absolute_first_derivative_tol: float = 1.0e-6
    max_line_search_iterations: int = 2
    line_search_strategy: int = MO_LS_POLYNOMIAL_APPROXIMATION
    armijo_search_tau: float = 0.8
    equality_penalty_initial: float = 1.0
    equality_penalty_scale_factor: float = 1.01
    equality_penalty_rho: float = 0.1
    lambda_initial: float = 0.0
    lambda_failure_init: float = 1.0e-2
    lambda_decrease_on_success: float = 0.1
    lambda_decrease_on_restore: float = 0.8
    max_lambda: float = 1.0
    min_lambda: float = 0.0
    retraction: int = MO_RETRACT_ADD
    inner_qp_solver: int = 0  # MO_NLS_QP_AUTO; 1 = MO_NLS_QP_GENERIC

    def to_c(self) -> MoNlsParams:
        return MoNlsParams(
            int(self.max_iterations), int(self.max_qp_iterations), self.termination_kkt_tolerance,
            self.absolute_exit_tol, self.relative_exit_tol, self.absolute_first_derivative_tol,
            int(self.max_line_search_iterations), int(self.line_search_strategy), self.armijo_search_tau,
            self.equality_penalty_initial, self.equality_penalty_scale_factor, self.equality_penalty_rho,
            self.lambda_initial, self.lambda_failure_init, self.lambda_decrease_on_success,
            self.lambda_decrease_on_restore, self.max_lambda, self.min_lambda, int(self.retraction), int(self.inner_qp_solver))


class MoNlsLinearizeArgs(C.Structure):
    _fields_ = [
        ("n", C.c_int32), ("r_cost", C.c_int32), ("cost_block", C.c_int32), ("r_eq", C.c_int32),
        ("m", C.c_int32), ("reserved", C.c_int32), ("batch", C.c_int64),
        ("x", _dp), ("J_cost", _dp), ("r_cost_v", _dp), ("J_eq", _dp), ("r_eq_v", _dp), ("lam", _dp),
        ("ineq_var", _dp), ("ineq_a", _dp), ("ineq_b", _dp),
        ("G", _dp), ("c", _dp), ("A_eq", _dp), ("b_eq", _dp), ("ineq_b_shifted", _dp), ("errors", _dp),
    ]


class MoChainIkBatch(C.Structure):
    _fields_ = [
        ("links", C.c_int32), ("n", C.c_int32), ("n_cost", C.c_int32), ("cost_block", C.c_int32),
        ("n_eq", C.c_int32), ("m", C.c_int32), ("batch", C.c_int64),
        ("link_pose", _dp), ("link_mask", _dp), ("cost_w", _dp), ("cost_target", _dp), ("eq_w", _dp),
        ("eq_target", _dp), ("ineq_var", _dp), ("ineq_a", _dp), ("ineq_b", _dp), ("x0", _dp),
    ]


MO_NLS_FAMILY_ROSENBROCK, MO_NLS_FAMILY_HIMMELBLAU, MO_NLS_FAMILY_SPHERE_PRODUCT = 0, 1, 2


class MoNlsFamilyBatch(C.Structure):
    _fields_ = [
        ("family", C.c_int32), ("m", C.c_int32), ("batch", C.c_int64),
        ("ineq_var", _dp), ("ineq_a", _dp), ("ineq_b", _dp), ("x0", _dp), ("data", _dp),
        ("data_stride", C.c_int32), ("reserved", C.c_int32),
    ]


class MoNlsOut(C.Structure):
    _fields_ = [("x", _dp), ("termination", _dp), ("iterations", _dp), ("qp_iterations", _dp), ("ls_steps", _dp),
                ("errors", _dp)]


@dataclass
class ChainIKArrays:
    """A batch of transform-chain IK problems (``mo_chain_ik_batch``): numpy arrays or torch tensors."""

    links: int
    n: int
    n_cost: int
    cost_block: int
    n_eq: int
    m: int
    batch: int
    link_pose: object      # [links, 7] (qw,qx,qy,qz,tx,ty,tz)
    link_mask: object      # [links, 6] uint8
    cost_w: object         # [n_cost, 3]
    cost_target: object    # [batch, n_cost]
    eq_w: object = None
    eq_target: object = None
    ineq_var: object = None
    ineq_a: object = None
    ineq_b: object = None
    x0: object = None

    def to_c(self) -> MoChainIkBatch:
        return MoChainIkBatch(self.links, self.n, self.n_cost, self.cost_block, self.n_eq, self.m, self.batch,
                              _addr(self.link_pose), _addr(self.link_mask), _addr(self.cost_w),
                              _addr(self.cost_target), _addr(self.eq_w), _addr(self.eq_target),
                              _addr(self.ineq_var), _addr(self.ineq_a), _addr(self.ineq_b), _addr(self.x0))


@dataclass
class NLSOutArrays:
    x: object
    termination: object
    iterations: object
    qp_iterations: object
    ls_steps: object
    errors: object = None

    def to_c(self) -> MoNlsOut:
        return MoNlsOut(_addr(self.x), _addr(self.termination), _addr(self.iterations), _addr(self.qp_iterations),
                        _addr(self.ls_steps), _addr(self.errors))


def alloc_nls_out_numpy(batch: int, n: int) -> NLSOutArrays:
    return NLSOutArrays(x=np.zeros((batch, n)), termination=np.full(batch, -1, dtype=np.int32),
                        iterations=np.zeros(batch, dtype=np.int32), qp_iterations=np.zeros(batch, dtype=np.int32),
                        ls_steps=np.zeros(batch, dtype=np.int32), errors=np.zeros((batch, 2)))


@dataclass
class LinearizeArrays:
    """Inputs of ``mo_nls_linearize_batch`` (dense per-problem Jacobians); numpy or torch."""

    n: int
    r_cost: int
    cost_block: int
    r_eq: int
    m: int
    batch: int
    x: object
    J_cost: object      # [B, n, r_cost]  (column-major r_cost x n)
    r_cost_v: object    # [B, r_cost]
    J_eq: object = None
    r_eq_v: object = None
    lam: object = None  # [B]
    ineq_var: object = None
    ineq_a: object = None
    ineq_b: object = None


@dataclass
class LinearizeOut:
    G: object
    c: object
    A_eq: object
    b_eq: object
    ineq_b_shifted: object
    errors: object


def linearize_to_c(a: LinearizeArrays, o: LinearizeOut) -> MoNlsLinearizeArgs:
    return MoNlsLinearizeArgs(a.n, a.r_cost, a.cost_block, a.r_eq, a.m, 0, a.batch, _addr(a.x), _addr(a.J_cost),
                              _addr(a.r_cost_v), _addr(a.J_eq), _addr(a.r_eq_v), _addr(a.lam), _addr(a.ineq_var),
                              _addr(a.ineq_a), _addr(a.ineq_b), _addr(o.G), _addr(o.c), _addr(o.A_eq),
                              _addr(o.b_eq), _addr(o.ineq_b_shifted), _addr(o.errors))


def alloc_linearize_out_numpy(a: LinearizeArrays) -> LinearizeOut:
    B, n, k, m = a.batch, a.n, a.r_eq, a.m
    return LinearizeOut(G=np.zeros((B, n, n)), c=np.zeros((B, n)), A_eq=np.zeros((B, n, k)), b_eq=np.zeros((B, k)),
                        ineq_b_shifted=np.zeros((B, m)), errors=np.zeros((B, 2)))
