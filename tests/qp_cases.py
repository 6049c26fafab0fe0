"""The literal QPs of the reference's own unit tests (BASELINE config 1), rebuilt from the numbers in
``/root/reference/test/qp_test.cc``. Each case cites the test it mirrors. Used by the golden-vector
script (tools/make_golden.py), by the oracle tests and by the GPU parity tests.
"""
from __future__ import annotations

import numpy as np

from mini_opt_b200 import abi


def build_quadratic(roots):
    """QPSolverTest::BuildQuadratic (qp_test.cc:58-73): sum_i (a_i x_i - b_i)^2 as x'Gx + c'x + const."""
    n = len(roots)
    G = np.zeros((n, n))
    c = np.zeros(n)
    for i, (a, b) in enumerate(roots):
        G[i, i] = a * a
        c[i] = -2 * a * b
    return G, c


def var_le(i, v):  # Var(i) <= v  (qp.hpp:81-83)
    return (i, -1.0, float(v))


def var_ge(i, v):  # Var(i) >= v  (qp.hpp:86-88)
    return (i, 1.0, -float(v))


def make_batch(G, c, A_eq=None, b_eq=None, constraints=()) -> abi.QPBatchArrays:
    """Single-problem batch from dense numpy pieces (G symmetric or lower-triangular, A_eq as k x n)."""
    n = G.shape[0]
    k = 0 if A_eq is None else A_eq.shape[0]
    m = len(constraints)
    arrs = abi.QPBatchArrays(
        n=n, k=k, m=m, batch=1,
        G=np.ascontiguousarray(G.T.reshape(1, n, n)),  # column-major storage of G
        c=np.ascontiguousarray(np.asarray(c, dtype=float).reshape(1, n)),
        A_eq=np.ascontiguousarray(A_eq.T.reshape(1, n, k)) if k else None,
        b_eq=np.ascontiguousarray(np.asarray(b_eq, dtype=float).reshape(1, k)) if k else None,
        ineq_var=np.array([[v for v, _, _ in constraints]], dtype=np.int32) if m else None,
        ineq_a=np.array([[a for _, a, _ in constraints]], dtype=float) if m else None,
        ineq_b=np.array([[b for _, _, b in constraints]], dtype=float) if m else None,
    )
    return arrs


def dummy_state(n, k, m, x_guess):
    """PutDummyValuesInSlacksAndMultipliers (qp_test.cc:84-97): s_i=2/(i+1), z_i=(i+1)/2, y_k=(k+1)^2."""
    v = np.zeros(n + 2 * m + k)
    v[:n] = x_guess
    for i in range(m):
        v[n + i] = 2.0 / (i + 1)
        v[n + m + k + i] = 0.5 * (i + 1)
    for j in range(k):
        v[n + m + j] = float((j + 1) * (j + 1))
    return v


def full_system_delta(arrs: abi.QPBatchArrays, state):
    """BuildFullSystem (qp.cc:595-655) + dense LU solve, as CheckAugmentedSolveAgainstPartialPivot does
    (qp_test.cc:101-138): the independent known answer for one Newton step with mu = 0.
    Returns delta in [dx|ds|dy|dz] order with dy, dz sign-flipped like the test."""
    n, k, m = arrs.n, arrs.k, arrs.m
    G = arrs.G[0].T
    G = np.tril(G) + np.tril(G, -1).T
    c = arrs.c[0]
    A = arrs.A_eq[0].T if k else np.zeros((0, n))
    b = arrs.b_eq[0] if k else np.zeros(0)
    Ai = np.zeros((m, n))
    bi = np.zeros(m)
    for i in range(m):
        Ai[i, arrs.ineq_var[0, i]] = arrs.ineq_a[0, i]
        bi[i] = arrs.ineq_b[0, i]
    x, s, y, z = state[:n], state[n:n + m], state[n + m:n + m + k], state[n + m + k:]
    T = n + 2 * m + k
    H = np.zeros((T, T))
    H[:n, :n] = G
    H[:n, n + m:n + m + k] = A.T
    H[n + m:n + m + k, :n] = A
    if m:
        H[:n, n + m + k:] = Ai.T
        H[n + m + k:, :n] = Ai
        H[n + m + k:, n:n + m] = -np.eye(m)
        H[n:n + m, n:n + m] = np.diag(z / s)
        H[n:n + m, n + m + k:] = -np.eye(m)
    r = np.zeros(T)
    r[:n] = G @ x + c - A.T @ y - Ai.T @ z
    r[n:n + m] = z
    r[n + m:n + m + k] = A @ x + b
    r[n + m + k:] = Ai @ x + bi - s
    d = np.linalg.solve(H, -r)
    d[n + m:n + m + k] *= -1
    d[n + m + k:] *= -1
    return d


def elimination_cases():
    """The four TestElimination* problems (qp_test.cc:168-241) -> list of (name, batch, x_guess)."""
    out = []
    G, c = build_quadratic([(0.5, 2.0), (5.0, 25.0), (3.0, 9.0)])
    out.append(("no_constraints", make_batch(G, c), np.array([0.0, -0.1, -0.3])))

    G, c = build_quadratic([(1.0, -0.5), (2.0, -2.0), (-4.0, 5.0)])
    A = np.array([[0.0, 1.0, -1.0]])
    out.append(("equality", make_batch(G, c, A, [-0.5]), np.array([0.3, -0.1, -0.3])))

    G, c = build_quadratic([(1.5, 3.0), (-1.0, 4.0)])
    cons = [var_ge(1, 0), var_le(0, 5), var_ge(0, -5)]
    out.append(("inequality", make_batch(G, c, constraints=cons), np.array([0.0, 2.0])))

    G, c = build_quadratic([(0.5, 2.0), (5.0, 25.0), (3.0, 9.0), (4.0, 1.0), (1.2, 2.4), (-1.0, 2.0), (-0.5, 2.0)])
    A = np.zeros((2, 7))
    A[0, 1], A[0, 4], A[1, 0] = 2.0, -1.0, 3.0
    cons = [(3, 4.0, -8.0), (5, 2.0, 1.0), (6, 1.0, 0.0)]
    out.append(("all", make_batch(G, c, A, [0.5, -2.0], cons), np.array([0.0, 0.1, 0.2, 0.55, 0.3, 0.7, 1.0])))
    return out


def solve_cases():
    """Fixed-answer Solve() tests (qp_test.cc:252-471) ->
    list of dict(name, batch, params, guesses, expect) where expect holds the asserted solution pieces."""
    P = abi.QPParams
    both = (abi.MO_GUESS_NAIVE, abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED)
    cases = []

    # TestWithSingleInequality (:252-287): residual x-5 linearised at 0 -> G=[1], c=[-5]; Var(0) <= 4
    cases.append(dict(
        name="single_inequality",
        batch=make_batch(np.array([[1.0]]), [-5.0], constraints=[var_le(0, 4)]),
        params=P(termination_kkt_tol=1e-9, sigma=0.1, initial_mu=0.1), guesses=(abi.MO_GUESS_NAIVE,),
        expect=dict(x=[4.0], s=[0.0], tol=1e-6, z_gt=[(0, 1.0 - 1e-6)])))

    # TestWithInequalitiesActive (:290-332): J=diag(1,-4), r(0)=(-2,-16) -> G=J'J, c=J'r
    J = np.diag([1.0, -4.0]); r = np.array([-2.0, -16.0])
    cases.append(dict(
        name="inequalities_active",
        batch=make_batch(J.T @ J, J.T @ r, constraints=[var_le(0, 1.0), var_ge(1, -3.0)]),
        params=P(termination_kkt_tol=1e-12, initial_mu=0.1, sigma=0.1), guesses=both,
        expect=dict(x=[1.0, -3.0], s=[0.0, 0.0], tol=1e-6)))

    # TestWithInequalitiesPartiallyActive (:335-375)
    J = np.diag([1.0, -1.0, 0.5]); r = np.array([-1.0, -3.0, -5.0])
    cases.append(dict(
        name="inequalities_partially_active",
        batch=make_batch(J.T @ J, J.T @ r, constraints=[var_ge(1, -2.0), var_ge(0, -3.5)]),
        params=P(termination_kkt_tol=1e-12, sigma=0.1, initial_mu=0.1), guesses=both,
        expect=dict(x=[1.0, -2.0, 10.0], s0=0.0, z1=0.0, tol=1e-6)))

    # TestWithEqualitiesOnly (:379-411): exactly one iteration
    G, c = build_quadratic([(1.0, 0.5), (3.0, 2.0), (-4.0, 5.0), (0.25, 4)])
    A = np.zeros((2, 4)); A[0, 0], A[0, 2], A[1, 1], A[1, 3] = 1, -0.5, 0.25, 1.0
    cases.append(dict(
        name="equalities_only",
        batch=make_batch(G, c, A, [3.0, -2.0]),
        params=P(termination_kkt_tol=1e-6, max_iterations=1), guesses=(abi.MO_GUESS_NAIVE,),
        expect=dict(eq_residual_tol=1e-9, iters=1)))

    # TestWithFullyConstrainedEqualities (:414-436)
    G, c = build_quadratic([(1.0, -0.5), (1.0, -0.25), (1.0, 1.0)])
    cases.append(dict(
        name="fully_constrained_equalities",
        batch=make_batch(G, c, np.eye(3), [-1.0, -2.0, -3.0]),
        params=P(termination_kkt_tol=1e-6, max_iterations=1), guesses=(abi.MO_GUESS_NAIVE,),
        expect=dict(x=[1.0, 2.0, 3.0], tol=1e-9, y_gt=0.01, iters=1)))

    # TestWithInequalitiesAndEqualities (:439-471)
    G, c = build_quadratic([(1.0, 1.0), (5.0, -10.0), (10.0, 2.0)])
    cases.append(dict(
        name="inequalities_and_equalities",
        batch=make_batch(G, c, np.array([[0.0, 0.0, 1.0]]), [-2.0], [var_le(0, 0.5), var_ge(1, -1.0)]),
        params=P(termination_kkt_tol=1e-12, initial_mu=0.1, sigma=0.1), guesses=both,
        expect=dict(x=[0.5, -1.0, 2.0], s=[0.0, 0.0], tol=1e-6)))
    return cases


def check_solve_expectations(case, out: abi.QPOutArrays, arrs: abi.QPBatchArrays):
    """The reference test's own assertions, applied to a solve result (numpy arrays, batch of 1)."""
    n, k, m = arrs.n, arrs.k, arrs.m
    e = case["expect"]
    v = np.asarray(out.vars)[0]
    x, s, y, z = v[:n], v[n:n + m], v[n + m:n + m + k], v[n + m + k:]
    assert int(np.asarray(out.status)[0]) == abi.MO_QP_SATISFIED_KKT_TOL, case["name"]
    tol = e.get("tol", 1e-6)
    if "x" in e:
        np.testing.assert_allclose(x, e["x"], atol=tol, rtol=0)
    if "s" in e:
        np.testing.assert_allclose(s, e["s"], atol=tol, rtol=0)
    if "s0" in e:
        assert abs(s[0] - e["s0"]) < tol
    if "z1" in e:
        assert abs(z[1] - e["z1"]) < tol
    for idx, lo in e.get("z_gt", []):
        assert z[idx] > lo
    if "y_gt" in e:
        assert (y > e["y_gt"]).all()
    if "iters" in e:
        assert int(np.asarray(out.iters)[0]) == e["iters"]
    if "eq_residual_tol" in e:
        A = arrs.A_eq[0].T
        np.testing.assert_allclose(A @ x + arrs.b_eq[0], 0.0, atol=e["eq_residual_tol"])
