"""CPU tests (no GPU): pin the plain-C NLS oracle (oracle/nls_oracle.c) against the golden vectors recorded from the
reference (tests/golden/nls_chain.npz, tools/make_golden.py), the reference's own test assertions, and the live
reference (oracle/_ref) when present."""
import sys
from pathlib import Path

import numpy as np
import pytest

import nls_cases as N
from mini_opt_b200 import abi, synth

sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tools"))
from make_golden import IK_SETS, LINEARIZE_SHAPES  # noqa: E402


@pytest.fixture(scope="module")
def golden_nls():
    return np.load(Path(__file__).parent / "golden" / "nls_chain.npz")


def counts_same(a, b_term, b_it, b_qp, b_ls):
    return (a.termination == b_term) & (a.iterations == b_it) & (a.qp_iterations == b_qp) & (a.ls_steps == b_ls)


@pytest.mark.parametrize("seed", range(6))
def test_oracle_chain_fk_golden(oracle, golden_nls, seed):
    """ActuatorChain::Update on general chains (mixed masks, arbitrary poses): 1e-12 like transform_chains_test.cc."""
    ch, P = N.random_general_chain(seed)
    fk = oracle.chain_fk("oracle", ch, P)
    for k, v in fk.items():
        np.testing.assert_allclose(v, golden_nls[f"fk{seed}_{k}"], atol=1e-12, rtol=0)


def test_oracle_chain_jacobian_numerical(oracle):
    """transform_chains_test.cc:120-231: analytic translation Jacobian vs central differences (1e-9 tolerance scale)."""
    ch, P = N.random_general_chain(3, batch=8)
    fk = oracle.chain_fk("oracle", ch, P)
    h = 1e-6
    for j in range(ch.n):
        d = np.zeros_like(P); d[:, j] = h
        num = (oracle.chain_fk("oracle", ch, P + d)["translation"] - oracle.chain_fk("oracle", ch, P - d)["translation"]) / (2 * h)
        np.testing.assert_allclose(fk["t_D_params"][:, j, :], num, atol=2e-9, rtol=0)


@pytest.mark.parametrize("idx", range(len(LINEARIZE_SHAPES)))
def test_oracle_linearize_golden(oracle, golden_nls, idx):
    """LinearizeAndFillQP: J'J, J'r to 1e-12, strict upper triangle exactly 0 (residual_test.cc:51-182)."""
    n, rc, cb, re = LINEARIZE_SHAPES[idx]
    a = N.random_linearize(100 + idx, n, rc, cb, re)
    out = oracle.nls_linearize("oracle", a)
    for f in ("G", "c", "A_eq", "b_eq", "ineq_b_shifted", "errors"):
        np.testing.assert_allclose(getattr(out, f), golden_nls[f"lin{idx}_{f}"], atol=1e-12, rtol=0)
    assert np.all(np.triu(out.G.transpose(0, 2, 1), 1) == 0.0)
    # independent numpy statement of the same quantities
    J = a.J_cost.transpose(0, 2, 1)
    G = np.einsum("bri,brj->bij", J, J) + a.lam[:, None, None] * np.eye(n)
    np.testing.assert_allclose(np.tril(out.G.transpose(0, 2, 1)), np.tril(G), atol=1e-12)
    np.testing.assert_allclose(out.c, np.einsum("bri,br->bi", J, a.r_cost_v), atol=1e-12)
    np.testing.assert_allclose(out.errors[:, 0], 0.5 * (a.r_cost_v ** 2).sum(1), atol=1e-12)


def test_oracle_two_angle_chain(oracle, golden_nls):
    """TestTwoAngleActuatorChain with inequalities (nonlinear_test.cc:925-960): the reference's assertions and its
    recorded run -- identical termination, iteration, QP-iteration and line-search counts on all 240 starts."""
    ch, prm = N.two_angle_chain()
    out, _ = oracle.nls_solve_chain("oracle", ch, prm)
    g = golden_nls
    assert counts_same(out, g["two_angle_termination"], g["two_angle_iterations"], g["two_angle_qp_iterations"],
                       g["two_angle_ls_steps"]).all()
    np.testing.assert_allclose(out.x, g["two_angle_x"], rtol=1e-9, atol=1e-12)
    fk = oracle.chain_fk("oracle", ch, out.x)
    assert np.abs(fk["translation"][:, :2] - np.array([0.45, 0.6])).max() < 1e-3
    assert out.ls_steps.max() < 100


@pytest.mark.parametrize("tag,links,neq,qpi,B", IK_SETS)
def test_oracle_chain_ik_golden(oracle, golden_nls, tag, links, neq, qpi, B):
    """BASELINE config 3 problems vs the reference's recorded solves. Redundant chains make J'J + 1e-9 I nearly singular
    and the 5..20-iteration QP is often unconverged or (with equalities) divergent, so trajectories are rounding
    sensitive: identical counts are required on a stated fraction, identical termination state on a larger one."""
    ch = synth.chain_ik_batch_numpy(links, B, seed=777 + links, n_eq=neq)
    out, _ = oracle.nls_solve_chain("oracle", ch, synth.chain_ik_params(qpi))
    g = golden_nls
    same = counts_same(out, g[f"{tag}_termination"], g[f"{tag}_iterations"], g[f"{tag}_qp_iterations"], g[f"{tag}_ls_steps"])
    need = 0.70 if neq else 0.95
    assert same.mean() >= need, same.mean()
    assert (out.termination == g[f"{tag}_termination"]).mean() >= need
    dx = np.abs(out.x - g[f"{tag}_x"]).max(axis=1)
    assert np.mean(dx[same] < 1e-6) >= 0.9, np.mean(dx[same] < 1e-6)


def test_oracle_vs_live_reference_nls(oracle):
    if not oracle.REF_SO.exists():
        pytest.skip("oracle/_ref not built (reference sources absent)")
    ch = synth.chain_ik_batch_numpy(10, 200, seed=5)
    for strategy in (abi.MO_LS_ARMIJO_BACKTRACK, abi.MO_LS_POLYNOMIAL_APPROXIMATION):
        prm = synth.chain_ik_params()
        prm.line_search_strategy = strategy
        r, _ = oracle.nls_solve_chain("ref", ch, prm, nthreads=4)
        o, _ = oracle.nls_solve_chain("oracle", ch, prm)
        same = counts_same(o, r.termination, r.iterations, r.qp_iterations, r.ls_steps)
        assert same.mean() >= 0.95, (strategy, same.mean())


def test_oracle_nls_rejects_bad_params(oracle):
    ch, prm = N.two_angle_chain()
    for bad in (dict(max_qp_iterations=0), dict(armijo_search_tau=1.0), dict(lambda_initial=2.0), dict(absolute_exit_tol=0.0)):
        p = abi.NLSParams(**{**prm.__dict__, **bad})
        with pytest.raises(RuntimeError):
            oracle.nls_solve_chain("oracle", ch, p)


def test_equality_only_chain_uses_nullspace_like_reference():
    """Equality rows and no inequalities: QPNullSpaceSolver + the (18.36) penalty (nonlinear.cc:83-89, 492-500). The C
    restatement against the reference on the same chains."""
    import dataclasses
    from oracle import oracle_py as O
    if not O.ref_available():
        pytest.skip("reference library not built")
    ch = synth.chain_ik_batch_numpy(6, 1500, seed=11, n_eq=1)
    ch = dataclasses.replace(ch, m=0, ineq_var=None, ineq_a=None, ineq_b=None)
    prm = synth.chain_ik_params()
    ref, _ = O.nls_solve_chain("ref", ch, prm, nthreads=4)
    orc, _ = O.nls_solve_chain("oracle", ch, prm)
    assert (ref.termination == orc.termination).mean() >= 0.995
    same = (ref.termination == orc.termination) & (ref.iterations == orc.iterations) & (ref.ls_steps == orc.ls_steps) & \
           (ref.qp_iterations == orc.qp_iterations)
    assert same.mean() >= 0.96, same.mean()
