"""Nonlinear least squares test problems. Each case cites the reference test it mirrors
(``/root/reference/test/nonlinear_test.cc``, ``transform_chains_test.cc``, ``residual_test.cc``)."""
from __future__ import annotations

import numpy as np

from mini_opt_b200 import abi


def quat_from_euler_xyz(x, y, z):
    """R = Rx Ry Rz as (w, x, y, z) -- the XYZ branch of SO3FromEulerAngles_ (so3.hpp:318-326)."""
    chx, shx, chy, shy, chz, shz = np.cos(x / 2), np.sin(x / 2), np.cos(y / 2), np.sin(y / 2), np.cos(z / 2), np.sin(z / 2)
    return np.array([chx * chy * chz - shx * shy * shz, chx * shy * shz + chy * chz * shx,
                     chx * chz * shy - chy * shx * shz, chx * chy * shz + chz * shx * shy])


def two_angle_chain(inequalities: bool = True) -> tuple[abi.ChainIKArrays, abi.NLSParams]:
    """TestTwoAngleActuatorChain, second half (nonlinear_test.cc:827-960): 3 Z-revolute links (the last one switched
    off), cost effector.y - 0.6, equality effector.x - 0.45, Var(1) in [0, pi], grid of initial guesses."""
    pose = np.array([[1, 0, 0, 0, 0.0, 0, 0], [1, 0, 0, 0, 0.4, 0, 0], [1, 0, 0, 0, 0.4, 0, 0]], dtype=float)
    mask = np.array([[0, 0, 1, 0, 0, 0], [0, 0, 1, 0, 0, 0], [0, 0, 0, 0, 0, 0]], dtype=np.uint8)
    guesses = []
    theta0 = 0.1
    while theta0 <= np.pi / 2:
        theta1 = 0.001
        while theta1 <= np.pi / 2 - 0.001:
            guesses.append((theta0, theta1))
            theta1 += 0.1
        theta0 += 0.1
    x0 = np.array(guesses)
    B = x0.shape[0]
    ch = abi.ChainIKArrays(links=3, n=2, n_cost=1, cost_block=1, n_eq=1, m=2, batch=B, link_pose=pose, link_mask=mask,
                           cost_w=np.array([[0.0, 1.0, 0.0]]), cost_target=np.full((B, 1), 0.6),
                           eq_w=np.array([[1.0, 0.0, 0.0]]), eq_target=np.full((B, 1), 0.45),
                           ineq_var=np.array([1, 1], dtype=np.int32), ineq_a=np.array([1.0, -1.0]),
                           ineq_b=np.array([0.0, np.pi]), x0=np.ascontiguousarray(x0))
    prm = abi.NLSParams(max_iterations=50, max_qp_iterations=10, relative_exit_tol=1e-12,
                        absolute_first_derivative_tol=1e-10, absolute_exit_tol=1e-9, termination_kkt_tolerance=1e-6,
                        max_line_search_iterations=10, equality_penalty_initial=0.01,
                        line_search_strategy=abi.MO_LS_ARMIJO_BACKTRACK, lambda_failure_init=0.001,
                        armijo_search_tau=0.5, lambda_initial=0.001, min_lambda=1e-9, retraction=abi.MO_RETRACT_MOD_PI)
    return ch, prm


def random_general_chain(seed: int, links: int = 5, batch: int = 64) -> tuple[abi.ChainIKArrays, np.ndarray]:
    """Chains with arbitrary link poses and mixed activity masks (cf. transform_chains_test.cc:120-231): returns the
    chain description and a [batch, n] matrix of parameter vectors."""
    rng = np.random.default_rng(seed)
    pose = np.zeros((links, 7))
    mask = np.zeros((links, 6), dtype=np.uint8)
    for i in range(links):
        e = rng.uniform(-0.8, 0.8, 3)
        pose[i, :4] = quat_from_euler_xyz(*e)
        pose[i, 4:] = rng.uniform(-0.5, 0.5, 3)
        kind = rng.integers(0, 5)
        if kind == 0:
            mask[i] = [0, 0, 1, 0, 0, 0]
        elif kind == 1:
            mask[i] = [1, 1, 1, 0, 0, 0]
        elif kind == 2:
            mask[i] = [0, 1, 0, 1, 0, 1]
        elif kind == 3:
            mask[i] = [0, 0, 0, 1, 1, 0]
        else:
            mask[i] = [0, 0, 0, 0, 0, 0]
    if mask.sum() == 0:
        mask[0] = [1, 0, 1, 0, 1, 0]
    n = int(mask.sum())
    params = rng.uniform(-1.0, 1.0, (batch, n))
    ch = abi.ChainIKArrays(links=links, n=n, n_cost=0, cost_block=0, n_eq=0, m=0, batch=batch, link_pose=pose,
                           link_mask=mask, cost_w=None, cost_target=None)
    return ch, params


def random_linearize(seed: int, n: int, r_cost: int, cost_block: int, r_eq: int, batch: int = 32) -> abi.LinearizeArrays:
    """Dense-Jacobian linearization inputs (cf. residual_test.cc:51-182, nonlinear_test.cc:109-183)."""
    rng = np.random.default_rng(seed)
    m = 2 * n
    var = np.repeat(np.arange(n, dtype=np.int32), 2)
    return abi.LinearizeArrays(
        n=n, r_cost=r_cost, cost_block=cost_block, r_eq=r_eq, m=m, batch=batch,
        x=rng.uniform(-1, 1, (batch, n)), J_cost=rng.uniform(-2, 2, (batch, n, r_cost)),
        r_cost_v=rng.uniform(-1, 1, (batch, r_cost)),
        J_eq=rng.uniform(-2, 2, (batch, n, r_eq)) if r_eq else None,
        r_eq_v=rng.uniform(-1, 1, (batch, r_eq)) if r_eq else None,
        lam=np.where(rng.uniform(size=batch) < 0.3, 0.0, rng.uniform(0, 0.1, batch)),
        ineq_var=np.ascontiguousarray(np.broadcast_to(var, (batch, m))),
        ineq_a=np.ascontiguousarray(np.broadcast_to(np.tile([1.0, -1.0], n), (batch, m))),
        ineq_b=rng.uniform(1.0, 2.0, (batch, m)))
