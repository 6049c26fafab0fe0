"""Seeded synthetic batches for BASELINE configs 2/4/5.

Recipe per problem: ``A ~ U(-1,1)^{n x n}``, ``G = A'A + 0.1 n I``, ``c ~ 4 U(-1,1)``, ``A_eq ~ U(-1,1)^{k x n}``,
``b_eq ~ 0.1 U(-1,1)``, box ``-1 <= x_i <= 1`` stored as the interleaved pairs ``(i,+1,1), (i,-1,1)``
(``MO_QP_FLAG_BOX_PAIRS``: constraint 2i is ``Var(i) >= -1``, 2i+1 is ``Var(i) <= 1``).

The same bytes are fed to the reference / oracle and to the CUDA path. Two generators with the same recipe:
numpy (tests, golden fixtures) and torch (bench: generates directly in HBM). They do NOT produce the same
stream; parity never compares across generators.
"""
from __future__ import annotations

import numpy as np

from . import abi


def box_qp_batch_numpy(n: int, k: int, batch: int, seed: int = 1234, explicit_constraints: bool = True
                       ) -> abi.QPBatchArrays:
    rng = np.random.default_rng(seed)
    A = rng.uniform(-1.0, 1.0, size=(batch, n, n))
    G = np.einsum("bji,bjk->bik", A, A) + 0.1 * n * np.eye(n)[None]
    G = 0.5 * (G + G.transpose(0, 2, 1))
    c = 4.0 * rng.uniform(-1.0, 1.0, size=(batch, n))
    A_eq = rng.uniform(-1.0, 1.0, size=(batch, k, n)) if k else None
    b_eq = 0.1 * rng.uniform(-1.0, 1.0, size=(batch, k)) if k else None
    m = 2 * n
    ineq_b = np.ones((batch, m))
    arrs = abi.QPBatchArrays(
        n=n, k=k, m=m, batch=batch,
        G=np.ascontiguousarray(G), c=np.ascontiguousarray(c),
        A_eq=np.ascontiguousarray(A_eq.transpose(0, 2, 1)) if k else None,  # column-major k x n
        b_eq=np.ascontiguousarray(b_eq) if k else None,
        ineq_b=ineq_b, flags=abi.MO_QP_FLAG_BOX_PAIRS)
    if explicit_constraints:
        var = np.repeat(np.arange(n, dtype=np.int32), 2)
        a = np.tile(np.array([1.0, -1.0]), n)
        arrs.ineq_var = np.ascontiguousarray(np.broadcast_to(var, (batch, m)))
        arrs.ineq_a = np.ascontiguousarray(np.broadcast_to(a, (batch, m)))
    return arrs


def box_qp_batch_torch(n: int, k: int, batch: int, seed: int, device, chunk: int = 1 << 18) -> abi.QPBatchArrays:
    """Same recipe, generated on ``device`` in chunks (bounded scratch memory)."""
    import torch

    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    f64 = dict(dtype=torch.float64, device=device)
    m = 2 * n
    G = torch.empty((batch, n, n), **f64)
    c = torch.empty((batch, n), **f64)
    A_eq = torch.empty((batch, n, k), **f64) if k else None
    b_eq = torch.empty((batch, k), **f64) if k else None
    eye = 0.1 * n * torch.eye(n, **f64)
    for lo in range(0, batch, chunk):
        hi = min(batch, lo + chunk)
        A = torch.rand((hi - lo, n, n), generator=gen, **f64) * 2 - 1
        Gc = torch.bmm(A.transpose(1, 2), A) + eye
        G[lo:hi] = 0.5 * (Gc + Gc.transpose(1, 2))
        c[lo:hi] = 4.0 * (torch.rand((hi - lo, n), generator=gen, **f64) * 2 - 1)
        if k:
            A_eq[lo:hi] = torch.rand((hi - lo, n, k), generator=gen, **f64) * 2 - 1
            b_eq[lo:hi] = 0.1 * (torch.rand((hi - lo, k), generator=gen, **f64) * 2 - 1)
    ineq_b = torch.ones((batch, m), **f64)
    return abi.QPBatchArrays(n=n, k=k, m=m, batch=batch, G=G, c=c, A_eq=A_eq, b_eq=b_eq, ineq_b=ineq_b,
                             flags=abi.MO_QP_FLAG_BOX_PAIRS)


def to_numpy(arrs: abi.QPBatchArrays, lo: int = 0, hi: int | None = None) -> abi.QPBatchArrays:
    """Host copy of problems [lo, hi) of a (possibly device-resident torch) batch."""
    hi = arrs.batch if hi is None else hi

    def conv(t):
        if t is None:
            return None
        s = t[lo:hi]
        return np.ascontiguousarray(s.cpu().numpy() if hasattr(s, "cpu") else s)

    return abi.QPBatchArrays(n=arrs.n, k=arrs.k, m=arrs.m, batch=hi - lo, G=conv(arrs.G), c=conv(arrs.c),
                             A_eq=conv(arrs.A_eq), b_eq=conv(arrs.b_eq), ineq_var=conv(arrs.ineq_var),
                             ineq_a=conv(arrs.ineq_a), ineq_b=conv(arrs.ineq_b),
                             ineq_count=conv(arrs.ineq_count), flags=arrs.flags)


def solve_params(guess: int = abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED) -> abi.QPParams:
    """The ``ComputeStepDirection`` configuration (source/nonlinear.cc:226-239) used by SURVEY §8(d)."""
    return abi.QPParams(initial_mu=1.0, sigma=0.1, termination_kkt_tol=1e-9, termination_complementarity_tol=1e-6,
                        max_iterations=30, barrier_strategy=abi.MO_BARRIER_COMPLEMENTARITY,
                        initial_guess_method=guess)


# ---------------------------------------------------------------------------------------------------------
# BASELINE config 3: transform-chain IK problems
# ---------------------------------------------------------------------------------------------------------
def chain_ik_batch_numpy(links: int, batch: int, seed: int = 4321, n_eq: int = 0) -> abi.ChainIKArrays:
    """Chains of ``links`` Z-revolute links, each ``Pose(Identity, (0.3, 0, 0))`` with mask {0,0,1,0,0,0} (the pattern
    of test/nonlinear_test.cc:833-837, 967-985). Cost: one 2-row Residual, effector (x, y) - target, target uniform
    inside 0.8 x reach. Joint limits Var(i) in [-pi/2, pi/2] (cf. :928-929), initial angles U(-0.5, 0.5).
    ``n_eq`` = 1 moves the x row into an equality constraint (the x_eq / y_res split of :840-859)."""
    rng = np.random.default_rng(seed)
    n = links
    pose = np.zeros((links, 7))
    pose[:, 0] = 1.0
    pose[:, 4] = 0.3
    mask = np.zeros((links, 6), dtype=np.uint8)
    mask[:, 2] = 1
    reach = 0.3 * (links - 1)  # the first link's translation comes before its own rotation? no: after; see FK
    reach = 0.3 * links
    radius = 0.8 * reach * np.sqrt(rng.uniform(0.0, 1.0, size=batch))
    phi = rng.uniform(-np.pi, np.pi, size=batch)
    target = np.stack([radius * np.cos(phi), radius * np.sin(phi)], axis=1)
    x0 = rng.uniform(-0.5, 0.5, size=(batch, n))
    var = np.repeat(np.arange(n, dtype=np.int32), 2)
    a = np.tile(np.array([1.0, -1.0]), n)
    b = np.full(2 * n, np.pi / 2)
    if n_eq == 0:
        cost_w = np.array([[1.0, 0, 0], [0, 1.0, 0]])
        return abi.ChainIKArrays(links=links, n=n, n_cost=2, cost_block=2, n_eq=0, m=2 * n, batch=batch,
                                 link_pose=pose, link_mask=mask, cost_w=cost_w,
                                 cost_target=np.ascontiguousarray(target), ineq_var=var, ineq_a=a, ineq_b=b, x0=x0)
    return abi.ChainIKArrays(links=links, n=n, n_cost=1, cost_block=1, n_eq=1, m=2 * n, batch=batch,
                             link_pose=pose, link_mask=mask, cost_w=np.array([[0.0, 1.0, 0.0]]),
                             cost_target=np.ascontiguousarray(target[:, 1:2]), eq_w=np.array([[1.0, 0.0, 0.0]]),
                             eq_target=np.ascontiguousarray(target[:, 0:1]), ineq_var=var, ineq_a=a, ineq_b=b, x0=x0)


def chain_slice(ch: abi.ChainIKArrays, lo: int, hi: int) -> abi.ChainIKArrays:
    """Problems [lo, hi) of a host chain-IK batch (the per-problem arrays are sliced, the shared ones kept)."""
    import dataclasses
    kw = {f.name: getattr(ch, f.name) for f in dataclasses.fields(ch)}
    for name in ("cost_target", "eq_target", "x0"):
        if kw[name] is not None:
            kw[name] = np.ascontiguousarray(kw[name][lo:hi])
    kw["batch"] = hi - lo
    return abi.ChainIKArrays(**kw)


def chain_ik_params(max_qp_iterations: int = 20) -> abi.NLSParams:
    """Params of TestDualActuatorBalancing (test/nonlinear_test.cc:1086-1102) + the ModPi retraction (:1076-1082).
    ``max_qp_iterations`` defaults to 20 instead of that test's 5: with 5 the REFERENCE itself brings only 9 % of the
    6-link problems below absolute_exit_tol (70 % with 20); pass 5 for the literal test configuration."""
    return abi.NLSParams(max_iterations=100, max_qp_iterations=max_qp_iterations, relative_exit_tol=1e-12,
                         absolute_first_derivative_tol=1e-10, absolute_exit_tol=1e-8,
                         termination_kkt_tolerance=1e-6, max_line_search_iterations=5,
                         line_search_strategy=abi.MO_LS_ARMIJO_BACKTRACK, lambda_failure_init=0.01,
                         armijo_search_tau=0.5, lambda_initial=0.001, min_lambda=1e-9,
                         retraction=abi.MO_RETRACT_MOD_PI)
