"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI, against
 (a) the plain-C oracle (oracle/qp_ip_oracle.c) on the same seeded inputs,
 (b) the golden vectors recorded from the reference itself (tests/golden/),
 (c) the reference's own unit-test assertions (tests/qp_cases.py <- test/qp_test.cc).

Tolerances. north_star: relative 1e-9 on x and the KKT norms, identical status and iteration count.
Problems whose start point is clamped onto a bound begin with s = 1e-9, z = 1e9 (qp.cc:472-481); their KKT
systems have condition ~1e18 and the trajectory is rounding-chaotic: even the plain-C oracle, which follows
Eigen's elimination order operation by operation, matches the reference's iteration count on only ~98-99 %
of those (tests/test_oracle_qp.py). Such cases are bounded by FRACTION here, never silently excluded; every
problem must still reach the same optimum (x to 1e-6) with the same status.
"""
import numpy as np
import pytest

import golden_util as gu
import qp_cases
from mini_opt_b200 import abi, synth

pytestmark = pytest.mark.gpu

X_RTOL = 1e-9
# Final KKT norms of a converged solve sit between the rounding floor (1e-16) and the termination tolerance
# (1e-9) and are cancellation noise there: they are compared to rtol 1e-6 OR within the termination tolerance.
# Norm parity where the norms carry information (1e9 .. 1e-6) is checked on the per-iteration logs.
KKT_ATOL = 1e-9


@pytest.fixture(scope="module")
def dev():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mini_opt_b200 import lib
    lib.load()  # must exist: no fallback
    return torch.device("cuda:0")


def gpu_solve(arrs, params, dev, log=False, user_vars=None):
    import torch
    from mini_opt_b200 import qp
    d = qp.to_device(arrs, dev)
    out = qp.alloc_out_torch(arrs.batch, arrs.n, arrs.k, arrs.m, dev, params.max_iterations, log)
    if user_vars is not None:
        out.vars.copy_(torch.from_numpy(user_vars))
    qp.solve_batch_device(d, params, out, dev)
    torch.cuda.synchronize()
    return qp.out_to_numpy(out)


def assert_log_close(got, want, cols=11):
    """Per-iteration log rows. Residual norms after a step can be pure cancellation noise (1e-16 x the
    1e9-sized multipliers of a clamped start), so the absolute tolerance scales with the largest entry."""
    got, want = np.asarray(got)[..., :cols], np.asarray(want)[..., :cols]
    scale = max(1.0, float(np.nanmax(np.abs(want)))) if want.size else 1.0
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-12 * scale, equal_nan=True)


def _record(**kw):
    """MO_RECORD_FRACTIONS=<file>: append the measured agreement fractions (used to pin the thresholds to measured - 0.3 %)."""
    import json, os
    path = os.environ.get("MO_RECORD_FRACTIONS")
    if path:
        kw["test"] = os.environ.get("PYTEST_CURRENT_TEST", "")
        with open(path, "a") as f:
            f.write(json.dumps(kw) + "\n")


def compare(out, ref, n, min_same=1.0, x_rtol=X_RTOL, min_x=1.0, min_status=1.0):
    """status / iteration-count / x agreement; fractions below 1.0 only for the rounding-chaotic cases above."""
    st_same = out.status == ref.status
    _s = (out.iters == ref.iters) & st_same
    _r = gu.rel_x_err(out.vars, ref.vars, n)
    _record(problems=int(len(st_same)), status=float(st_same.mean()), same=float(_s.mean()),
            x=float(np.mean(_r[_s] < x_rtol)) if _s.any() else None)
    assert st_same.mean() >= min_status, (st_same.mean(), np.flatnonzero(~st_same)[:10])
    same = (out.iters == ref.iters) & st_same
    assert same.mean() >= min_same, (same.mean(), np.flatnonzero(~same)[:10])
    rel = gu.rel_x_err(out.vars, ref.vars, n)
    assert np.mean(rel[same] < x_rtol) >= min_x, (np.mean(rel[same] < x_rtol), rel[same].max())
    both_ok = (out.status == 0) & (ref.status == 0)
    assert (rel[both_ok] < 1e-4).all(), rel[both_ok].max()  # same optimum (to the complementarity tolerance)
    return same, rel


@pytest.mark.parametrize("name_idx", range(4))
def test_newton_step_elimination_cases(dev, golden_fixed, name_idx):
    """TestElimination* (qp_test.cc:168-241): one Newton step (mu=0) on the dummy state, 1e-12 vs the
    reference's recorded delta AND vs the independent full-system LU answer."""
    import torch
    from mini_opt_b200 import qp
    name, arrs, x_guess = qp_cases.elimination_cases()[name_idx]
    state = qp_cases.dummy_state(arrs.n, arrs.k, arrs.m, x_guess)
    v = torch.from_numpy(state[None].copy()).to(dev)
    res = qp.newton_step_device(qp.to_device(arrs, dev), 0.0, abi.MO_BARRIER_COMPLEMENTARITY, v, dev)
    torch.cuda.synchronize()
    assert int(res["status"][0]) == 0
    delta = res["delta"][0].cpu().numpy()
    np.testing.assert_allclose(delta, golden_fixed[f"elim_{name}_delta"], atol=1e-12, rtol=0)
    np.testing.assert_allclose(res["vars"][0].cpu().numpy(), golden_fixed[f"elim_{name}_vars"], atol=1e-12, rtol=0)
    np.testing.assert_allclose(delta, qp_cases.full_system_delta(arrs, state), atol=1e-12, rtol=0)


@pytest.mark.parametrize("case", qp_cases.solve_cases(), ids=lambda c: c["name"])
def test_fixed_answer_solves(dev, golden_fixed, case):
    """qp_test.cc:252-471 (BASELINE config 1): the reference's assertions + its recorded outputs."""
    for guess in case["guesses"]:
        prm = abi.QPParams(**{**case["params"].__dict__, "initial_guess_method": guess})
        out = gpu_solve(case["batch"], prm, dev, log=True)
        qp_cases.check_solve_expectations(case, out, case["batch"])
        key = f"solve_{case['name']}_g{guess}"
        assert out.iters[0] == golden_fixed[key + "_iters"][0]
        assert out.status[0] == golden_fixed[key + "_status"][0]
        np.testing.assert_allclose(out.vars[0], golden_fixed[key + "_vars"], rtol=1e-9, atol=1e-12)
        its = int(out.iters[0])
        assert_log_close(out.iter_log[0, :its], golden_fixed[key + "_log"][:its])


@pytest.mark.parametrize("gname,guess", [("naive", abi.MO_GUESS_NAIVE), ("eq", abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED)])
def test_generated_problems(dev, golden_generated, gname, guess):
    """TestGeneratedProblems (qp_test.cc:527-574): the reference's own 1000 problems, ragged constraint counts."""
    g = golden_generated
    arrs = gu.generated_batch(g)
    prm = abi.QPParams(termination_kkt_tol=1e-12, max_iterations=30, initial_guess_method=guess)
    out = gpu_solve(arrs, prm, dev)
    n = arrs.n
    np.testing.assert_allclose(out.vars[:, :n], g["solution"], atol=5e-5, rtol=0)
    assert np.nanmax(np.abs(out.vars[:, n:2 * n])) < 5e-5
    cnt = g["ineq_count"]
    pad = np.arange(n)[None, :] >= cnt[:, None]
    assert np.isnan(out.vars[:, n:2 * n][pad]).all()  # unused slots are NaN
    assert (out.status == g[f"{gname}_status"]).all()
    same = out.iters == g[f"{gname}_iters"]
    assert same.mean() >= 0.99, same.mean()
    rel = gu.rel_x_err(out.vars, g[f"{gname}_vars"], n)
    assert (rel[same] < 1e-7).all() and np.median(rel) < 1e-12


@pytest.mark.parametrize("n", [8, 32, 64])
@pytest.mark.parametrize("gname,guess", [("naive", abi.MO_GUESS_NAIVE), ("eq", abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED)])
@pytest.mark.parametrize("explicit", [False, True], ids=["boxflag", "triplets"])
def test_box_synth_vs_reference_golden(dev, golden_box, n, gname, guess, explicit):
    """SURVEY 8(d) box QPs (configs 2/4/5 shapes) vs the reference's recorded solves. `boxflag` takes the tuned
    kernel where one exists, `triplets` the generic kernel."""
    g = golden_box
    arrs = gu.box_batch(g, n, explicit=explicit)
    out = gpu_solve(arrs, synth.solve_params(guess), dev, log=(n == 8))
    tag = f"n{n}_{gname}"

    class R:
        status, iters, vars = g[f"{tag}_status"], g[f"{tag}_iters"], g[f"{tag}_vars"]
    if gname == "naive" or n > 8:
        compare(out, R, n)  # identical status, iteration count and x (1e-9) on every problem
    else:  # n = 8, clamped starts: measured same 0.9883 (3 of 256), x 0.996, status 1.0
        compare(out, R, n, min_same=0.98, min_x=0.99, min_status=1.0)
    ok = out.iters == R.iters
    np.testing.assert_allclose(out.kkt_final[ok], g[f"{tag}_kkt"][ok], rtol=1e-6, atol=KKT_ATOL)
    tight = ok & (gu.rel_x_err(out.vars, R.vars, n) < X_RTOL)  # multipliers of the non-chaotic trajectories
    np.testing.assert_allclose(out.lagrange[tight], g[f"{tag}_lagrange"][tight], rtol=1e-6, atol=1e-9)
    if n == 8 and gname == "naive":
        for p in range(0, arrs.batch, 17):
            its = int(out.iters[p])
            assert_log_close(out.iter_log[p, :its], g[f"{tag}_log"][p, :its])


@pytest.mark.parametrize("n,k,B", [(8, 2, 20000), (8, 0, 5000), (4, 1, 5000), (6, 2, 2000), (12, 3, 2000), (16, 8, 1500), (32, 8, 1500), (64, 16, 400)])
@pytest.mark.parametrize("guess", [abi.MO_GUESS_NAIVE, abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED])
def test_box_synth_vs_oracle(dev, oracle, n, k, B, guess):
    """Fresh seeded batches: CUDA path vs the oracle on identical bytes."""
    arrs = synth.box_qp_batch_numpy(n, k, B, seed=4242 + n * 10 + k, explicit_constraints=False)
    prm = synth.solve_params(guess)
    ref, _ = oracle.qp_solve("oracle", arrs, prm)
    out = gpu_solve(arrs, prm, dev)
    if guess == abi.MO_GUESS_NAIVE:
        same, rel = compare(out, ref, n)  # measured 1.0 on every shape
    else:
        # clamped starts (s = 1e-9, z = 1e9). Measured on B200 (profiles/r02_gpu_test_fractions.jsonl), floors = measured - 0.3 %:
        # (8,2) same 0.99575 status 0.9998 x 0.99905; (6,2) same 0.9795 status 0.999 x 0.99847; every other shape 1.0
        floors = {(8, 2): (0.992, 0.9968, 0.996), (6, 2): (0.976, 0.996, 0.995)}.get((n, k), (0.997, 0.999, 0.997))
        same, rel = compare(out, ref, n, min_same=floors[0], min_status=floors[1], min_x=floors[2])
    np.testing.assert_allclose(out.kkt_final[same], ref.kkt_final[same], rtol=1e-6, atol=KKT_ATOL)


@pytest.mark.parametrize("strategy", [abi.MO_BARRIER_FIXED_DECREASE, abi.MO_BARRIER_PREDICTOR_CORRECTOR])
def test_barrier_strategies_vs_oracle(dev, oracle, strategy):
    arrs = synth.box_qp_batch_numpy(8, 2, 3000, seed=99, explicit_constraints=False)
    prm = synth.solve_params(abi.MO_GUESS_NAIVE)
    prm.barrier_strategy = strategy
    prm.decrease_mu_only_on_small_error = strategy == abi.MO_BARRIER_FIXED_DECREASE
    ref, _ = oracle.qp_solve("oracle", arrs, prm, log=True)
    out = gpu_solve(arrs, prm, dev, log=True)
    same, _ = compare(out, ref, 8)  # NAIVE start: measured 1.0
    p = int(np.flatnonzero(same)[0])
    its = int(out.iters[p])
    assert_log_close(out.iter_log[p, :its], ref.iter_log[p, :its], cols=14)


def test_user_provided_guess_and_flags(dev, oracle):
    arrs = synth.box_qp_batch_numpy(8, 2, 1000, seed=5, explicit_constraints=False)
    rng = np.random.default_rng(1)
    stride = abi.vars_stride(8, 2, 16)
    uv = np.zeros((1000, stride))
    uv[:, :8] = rng.uniform(-0.5, 0.5, (1000, 8))
    uv[:, 8:24] = rng.uniform(0.1, 1.0, (1000, 16))
    uv[:, 24:26] = rng.uniform(-1, 1, (1000, 2))
    uv[:, 26:] = rng.uniform(0.1, 1.0, (1000, 16))
    prm = synth.solve_params(abi.MO_GUESS_USER_PROVIDED)
    prm.initialize_mu_with_complementarity = True
    ref, _ = oracle.qp_solve("oracle", arrs, prm, user_vars=uv)
    out = gpu_solve(arrs, prm, dev, user_vars=uv)
    compare(out, ref, 8)  # interior user start: measured 1.0


@pytest.mark.parametrize("n,k,explicit,variant", [(8, 2, False, "tpp"), (6, 2, False, "warp16"), (8, 2, True, "warp16"),
                                                 (16, 8, False, "dmma_warp"), (20, 4, False, "generic")])
def test_nonpositive_slack_status(dev, oracle, n, k, explicit, variant):
    """F_ASSERT(all s > 0) of ComputeLDLT (qp.cc:285) is a per-problem status here: a USER_PROVIDED start with one slack <= 0
    ends as MO_QP_NONPOSITIVE_SLACK (3) on every kernel family, its neighbours in the batch are untouched, and the oracle's
    batched stand-in reports the same."""
    from mini_opt_b200 import lib
    B = 96
    arrs = synth.box_qp_batch_numpy(n, k, B, seed=21, explicit_constraints=explicit)
    if explicit:
        arrs.flags = 0  # plain (variable, a, b) triplets: not the box-pair fast path
    m = 2 * n
    assert lib.kernel_variant(n, k, m, arrs.flags).startswith(variant), lib.kernel_variant(n, k, m, arrs.flags)
    rng = np.random.default_rng(3)
    uv = np.zeros((B, abi.vars_stride(n, k, m)))
    uv[:, :n] = rng.uniform(-0.5, 0.5, (B, n))
    uv[:, n:n + m] = rng.uniform(0.1, 1.0, (B, m))
    uv[:, n + m:n + m + k] = rng.uniform(-1, 1, (B, k))
    uv[:, n + m + k:] = rng.uniform(0.1, 1.0, (B, m))
    bad = np.arange(0, B, 7)
    uv[bad, n + (bad % m)] = -0.25       # a negative slack ...
    uv[bad[::2], n + (bad[::2] % m)] = 0.0  # ... or exactly zero
    prm = synth.solve_params(abi.MO_GUESS_USER_PROVIDED)
    ref, _ = oracle.qp_solve("oracle", arrs, prm, user_vars=uv)
    out = gpu_solve(arrs, prm, dev, user_vars=uv)
    good = np.setdiff1d(np.arange(B), bad)
    assert (out.status[bad] == abi.MO_QP_NONPOSITIVE_SLACK).all(), out.status[bad]
    assert (ref.status[bad] == abi.MO_QP_NONPOSITIVE_SLACK).all(), ref.status[bad]
    assert (out.iters[bad] == 0).all()
    assert np.isin(out.status[good], (abi.MO_QP_SATISFIED_KKT_TOL, abi.MO_QP_MAX_ITERATIONS)).all()
    assert (out.status[good] == ref.status[good]).mean() >= 0.97


@pytest.mark.parametrize("n,k", [(8, 2), (4, 1)])
@pytest.mark.parametrize("B", [1, 33, 1025, 37889])
def test_tpp_refill_schedules_agree(dev, n, k, B):
    """The cp.async refill of the thread-per-problem kernel (issued mid-round, lanes join one round later) and the round-1
    register refill (mo_debug_set_tpp_async) must give bit-identical outputs, also for batches smaller than a warp, one past
    a warp, and one past the resident thread slots."""
    from mini_opt_b200 import lib
    L = lib.load()
    arrs = synth.box_qp_batch_numpy(n, k, B, seed=31, explicit_constraints=False)
    assert lib.kernel_variant(n, k, 2 * n, arrs.flags) == "tpp"
    prm = synth.solve_params()
    outs = []
    try:
        for mode in (1, 0):
            L.mo_debug_set_tpp_async(mode)
            outs.append(gpu_solve(arrs, prm, dev))
    finally:
        L.mo_debug_set_tpp_async(1)
    a, b = outs
    assert np.array_equal(a.status, b.status) and np.array_equal(a.iters, b.iters)
    assert np.array_equal(a.vars.view(np.int64), b.vars.view(np.int64))
    assert np.array_equal(a.kkt_final.view(np.int64), b.kkt_final.view(np.int64))
    assert (a.status >= 0).all() and (a.iters >= 1).all()


def test_max_iterations_and_failures(dev, oracle):
    """MAX_ITERATIONS status, and per-problem failure codes instead of exceptions."""
    arrs = synth.box_qp_batch_numpy(8, 2, 500, seed=11, explicit_constraints=False)
    prm = synth.solve_params(abi.MO_GUESS_NAIVE)
    prm.max_iterations = 3
    ref, _ = oracle.qp_solve("oracle", arrs, prm)
    out = gpu_solve(arrs, prm, dev)
    assert (out.status == abi.MO_QP_MAX_ITERATIONS).all() and (out.iters == 3).all()
    compare(out, ref, 8)
    # singular KKT: G = 0 -> zero leading pivots with non-zero columns -> FailedFactorization (qp.cc:303-307)
    bad = synth.box_qp_batch_numpy(8, 2, 64, seed=12, explicit_constraints=False)
    bad.G[:] = 0.0
    ref, _ = oracle.qp_solve("oracle", bad, synth.solve_params(abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED))
    for explicit in (False, True):
        b2 = bad
        if explicit:
            b2 = synth.box_qp_batch_numpy(8, 2, 64, seed=12, explicit_constraints=True)
            b2.G[:] = 0.0
            b2.flags = 0
        out = gpu_solve(b2, synth.solve_params(abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED), dev)
        assert (out.status == abi.MO_QP_FAILED_FACTORIZATION).all(), out.status
    assert (ref.status == abi.MO_QP_FAILED_FACTORIZATION).all()


def test_host_entry_point_matches_device(dev, oracle):
    """mo_qp_ip_solve_batch_host (H2D + solve + D2H, chunked) == device entry point, incl. ragged tail chunk."""
    from mini_opt_b200 import qp
    arrs = synth.box_qp_batch_numpy(8, 2, 10007, seed=21, explicit_constraints=False)
    prm = synth.solve_params()
    out_d = gpu_solve(arrs, prm, dev)
    out_h = abi.alloc_out_numpy(arrs.batch, 8, 2, 16)
    qp.solve_batch_host(arrs, prm, out_h, 0)
    assert np.array_equal(out_h.vars, out_d.vars) and np.array_equal(out_h.iters, out_d.iters)
    assert np.array_equal(out_h.status, out_d.status) and np.array_equal(out_h.kkt_final, out_d.kkt_final)


def test_empty_and_invalid(dev):
    from mini_opt_b200 import lib, qp
    arrs = synth.box_qp_batch_numpy(8, 2, 0, seed=1, explicit_constraints=False)
    out = abi.alloc_out_numpy(0, 8, 2, 16)
    qp.solve_batch_host(arrs, synth.solve_params(), out, 0)  # no-op
    arrs = synth.box_qp_batch_numpy(8, 2, 4, seed=1, explicit_constraints=False)
    out = abi.alloc_out_numpy(4, 8, 2, 16)
    for bad in (dict(initial_mu=0.0), dict(sigma=1.5), dict(termination_kkt_tol=0.0), dict(max_iterations=0)):
        with pytest.raises(lib.MiniOptError) as e:
            qp.solve_batch_host(arrs, abi.QPParams(**bad), out, 0)
        assert e.value.code == abi.MO_ERR_INVALID_PARAMS
    arrs.m = 15
    with pytest.raises(lib.MiniOptError) as e:
        qp.solve_batch_host(arrs, synth.solve_params(), out, 0)
    assert e.value.code == abi.MO_ERR_INVALID_ARGUMENT


def test_full_size_properties(dev):
    """BASELINE config 2 at full size (1M QPs, n=8, k=2): size-independent properties of the solution --
    status/iteration bounds, primal feasibility, KKT residuals recomputed on the host in numpy, and
    batch-permutation invariance (lane refill must not leak state between problems)."""
    import torch
    from mini_opt_b200 import qp
    B = 1 << 20
    arrs = synth.box_qp_batch_torch(8, 2, B, seed=7, device=dev)
    prm = synth.solve_params()
    out = qp.alloc_out_torch(B, 8, 2, 16, dev)
    qp.solve_batch_device(arrs, prm, out, dev)
    torch.cuda.synchronize()
    status, iters = out.status.cpu().numpy(), out.iters.cpu().numpy()
    assert np.isin(status, [0, 1]).all() and (status == 0).mean() > 0.995
    assert iters.min() >= 1 and iters.max() <= 30 and (iters[status == 1] == 30).all()
    ok = torch.from_numpy(status == 0).to(dev)
    x = out.vars[:, :8]
    s, y, z = out.vars[:, 8:24], out.vars[:, 24:26], out.vars[:, 26:42]
    G = torch.tril(arrs.G.transpose(1, 2))  # row-major lower triangle
    Gs = G + torch.tril(G, -1).transpose(1, 2)
    A = arrs.A_eq.transpose(1, 2)  # [B, k, n]
    r_d = torch.einsum("bij,bj->bi", Gs, x) + arrs.c - torch.einsum("bki,bk->bi", A, y) - (z[:, 0::2] - z[:, 1::2])
    r_pe = torch.einsum("bki,bi->bk", A, x) + arrs.b_eq
    assert float(r_d[ok].norm(dim=1).max()) < 1e-8 and float(r_pe[ok].norm(dim=1).max()) < 1e-8
    assert float(x[ok].abs().max()) <= 1.0 + 1e-8 and float(s[ok].min()) > 0 and float(z[ok].min()) > 0
    assert float((s * z)[ok].mean(dim=1).max()) < 1e-6
    # permutation invariance on a slice
    perm = torch.randperm(50000, device=dev)
    sub = abi.QPBatchArrays(n=8, k=2, m=16, batch=50000, G=arrs.G[:50000][perm].contiguous(),
                            c=arrs.c[:50000][perm].contiguous(), A_eq=arrs.A_eq[:50000][perm].contiguous(),
                            b_eq=arrs.b_eq[:50000][perm].contiguous(), ineq_b=arrs.ineq_b[:50000][perm].contiguous(),
                            flags=arrs.flags)
    out2 = qp.alloc_out_torch(50000, 8, 2, 16, dev)
    qp.solve_batch_device(sub, prm, out2, dev)
    torch.cuda.synchronize()
    assert torch.equal(out2.vars, out.vars[:50000][perm]) and torch.equal(out2.iters, out.iters[:50000][perm])


# The two DMMA kernels behind the same shapes: warp per problem (default) and the CTA-per-problem kernel, selected
# through the explicit debug hook of the ABI (mo_debug_set_dm_kernel; the library reads no environment variables).
DMMA_VARIANTS = {"warp": 0, "cta": 1}


@pytest.fixture
def dmma_variant(request):
    from mini_opt_b200 import lib
    L = lib.load()
    L.mo_debug_set_dm_kernel(DMMA_VARIANTS[request.param])
    yield request.param
    L.mo_debug_set_dm_kernel(0)


@pytest.mark.gpu
@pytest.mark.parametrize("dmma_variant", list(DMMA_VARIANTS), indirect=True)
@pytest.mark.parametrize("n,k,B", [(64, 16, 8192), (32, 8, 16384), (16, 8, 16384)])
def test_dmma_shapes_properties_and_determinism(dev, n, k, B, dmma_variant):
    """DMMA kernels at a batch that fills every SM with its full residency: KKT residuals recomputed on the host side
    in torch, and bit-identical results across repeated runs and under a batch permutation (a shared-memory race or
    state leaking between the problems of a persistent CTA shows up as nondeterminism)."""
    import torch
    from mini_opt_b200 import qp, lib
    m = 2 * n
    arrs = synth.box_qp_batch_torch(n, k, B, seed=11, device=dev)
    prm = synth.solve_params()
    want = "dmma_cta" if dmma_variant == "cta" else "dmma_warp"
    assert lib.load().mo_qp_kernel_variant(n, k, m, arrs.flags).decode() == want
    outs = []
    for _ in range(3):
        out = qp.alloc_out_torch(B, n, k, m, dev)
        qp.solve_batch_device(arrs, prm, out, dev)
        torch.cuda.synchronize()
        outs.append(out)
    for o in outs[1:]:
        assert torch.equal(o.vars, outs[0].vars) and torch.equal(o.iters, outs[0].iters) and torch.equal(o.status, outs[0].status)
    out = outs[0]
    status = out.status.cpu().numpy()
    assert np.isin(status, [0, 1]).all() and (status == 0).mean() > 0.99
    ok = out.status == 0
    x, s = out.vars[:, :n], out.vars[:, n:n + m]
    y, z = out.vars[:, n + m:n + m + k], out.vars[:, n + m + k:]
    G = torch.tril(arrs.G.transpose(1, 2))
    Gs = G + torch.tril(G, -1).transpose(1, 2)
    A = arrs.A_eq.transpose(1, 2)
    r_d = torch.einsum("bij,bj->bi", Gs, x) + arrs.c - torch.einsum("bki,bk->bi", A, y) - (z[:, 0::2] - z[:, 1::2])
    r_pe = torch.einsum("bki,bi->bk", A, x) + arrs.b_eq
    scale = 1.0 + float(arrs.c.abs().max())
    assert float(r_d[ok].norm(dim=1).max()) < 1e-7 * scale and float(r_pe[ok].norm(dim=1).max()) < 1e-7 * scale
    assert float(s[ok].min()) > 0 and float(z[ok].min()) > 0
    nb = B // 4
    perm = torch.randperm(nb, device=dev)
    sub = abi.QPBatchArrays(n=n, k=k, m=m, batch=nb, G=arrs.G[:nb][perm].contiguous(), c=arrs.c[:nb][perm].contiguous(),
                            A_eq=arrs.A_eq[:nb][perm].contiguous(), b_eq=arrs.b_eq[:nb][perm].contiguous(),
                            ineq_b=arrs.ineq_b[:nb][perm].contiguous(), flags=arrs.flags)
    out2 = qp.alloc_out_torch(nb, n, k, m, dev)
    qp.solve_batch_device(sub, prm, out2, dev)
    torch.cuda.synchronize()
    assert torch.equal(out2.vars, out.vars[:nb][perm]) and torch.equal(out2.iters, out.iters[:nb][perm])


@pytest.mark.gpu
@pytest.mark.parametrize("n,k,B", [(64, 16, 2048), (32, 8, 4096), (16, 8, 4096)])
@pytest.mark.parametrize("guess", [abi.MO_GUESS_NAIVE, abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED])
def test_dmma_kernel_variants_agree(dev, n, k, B, guess):
    """The warp-per-problem kernel and the CTA-per-problem kernel factor the same KKT matrices in a
    different order of operations (left-looking from shared-memory tiles vs right-looking in accumulator fragments):
    same status and iteration count on >= 97 % of the problems (start points clamped onto a bound are
    rounding-chaotic; NAIVE starts must agree on all), and x to relative 1e-9 where they do."""
    import torch
    from mini_opt_b200 import qp
    m = 2 * n
    arrs = synth.box_qp_batch_torch(n, k, B, seed=5, device=dev)
    prm = synth.solve_params(guess=guess)
    res = {}
    from mini_opt_b200 import lib
    for name, which in DMMA_VARIANTS.items():
        lib.load().mo_debug_set_dm_kernel(which)
        out = qp.alloc_out_torch(B, n, k, m, dev)
        qp.solve_batch_device(arrs, prm, out, dev)
        torch.cuda.synchronize()
        res[name] = qp.out_to_numpy(out)
    lib.load().mo_debug_set_dm_kernel(0)
    for name, ref_name in (("warp", "cta"),):
        got, ref = res[name], res[ref_name]
        same = (got.iters == ref.iters) & (got.status == ref.status)
        assert same.mean() >= (1.0 if guess == abi.MO_GUESS_NAIVE else 0.97), (name, same.mean())
        rel = np.abs(got.vars[:, :n] - ref.vars[:, :n]).max(axis=1) / np.maximum(1.0, np.abs(ref.vars[:, :n]).max(axis=1))
        assert rel[same].max() < 1e-9, (name, rel[same].max())


@pytest.mark.gpu
@pytest.mark.parametrize("dmma_variant", ["warp", "cta"], indirect=True)
@pytest.mark.parametrize("n,k", [(64, 16), (32, 8), (16, 8)])
def test_dmma_kernels_log_and_user_guess_vs_oracle(dev, oracle, n, k, dmma_variant):
    """The corners of the DMMA kernels that the golden box batches do not reach: the per-iteration log
    (QPInteriorPointIteration rows: kkt_initial, kkt_final, mu, alpha) and USER_PROVIDED start points with
    initialize_mu_with_complementarity, both against the CPU oracle on the same inputs."""
    m, B = 2 * n, 256
    arrs = synth.box_qp_batch_numpy(n, k, B, seed=21, explicit_constraints=False)
    prm = synth.solve_params(abi.MO_GUESS_NAIVE)
    ref, _ = oracle.qp_solve("oracle", arrs, prm, log=True)
    out = gpu_solve(arrs, prm, dev, log=True)
    compare(out, ref, n)
    for p in range(0, B, 13):
        its = int(out.iters[p])
        assert_log_close(out.iter_log[p, :its], ref.iter_log[p, :its])
    rng = np.random.default_rng(3)
    uv = np.zeros((B, abi.vars_stride(n, k, m)))
    uv[:, :n] = rng.uniform(-0.5, 0.5, (B, n))
    uv[:, n:n + m] = rng.uniform(0.1, 1.0, (B, m))
    uv[:, n + m:n + m + k] = rng.uniform(-1, 1, (B, k))
    uv[:, n + m + k:] = rng.uniform(0.1, 1.0, (B, m))
    prm = synth.solve_params(abi.MO_GUESS_USER_PROVIDED)
    prm.initialize_mu_with_complementarity = True
    ref, _ = oracle.qp_solve("oracle", arrs, prm, user_vars=uv)
    out = gpu_solve(arrs, prm, dev, user_vars=uv)
    compare(out, ref, n)  # interior user start on the DMMA kernels: measured 1.0


def test_host_entry_compact_formats(dev, oracle):
    """mo_qp_ip_solve_batch_host with MO_QP_FLAG_PACKED_G | MO_QP_FLAG_UNIFORM_BOX in and x-only out gives exactly the
    device path's results (the formats are expanded on the device before the same kernels run)."""
    from mini_opt_b200 import qp
    for n, k, B in ((8, 2, 20000), (32, 8, 3000), (12, 3, 3000)):
        arrs = synth.box_qp_batch_numpy(n, k, B, seed=321 + n, explicit_constraints=False)
        prm = synth.solve_params()
        want = gpu_solve(arrs, prm, dev)
        compact = abi.QPBatchArrays(n=n, k=k, m=2 * n, batch=B, G=qp.pack_lower(arrs.G), c=arrs.c, A_eq=arrs.A_eq,
                                    b_eq=arrs.b_eq, ineq_b=np.ascontiguousarray(arrs.ineq_b[0]),
                                    flags=arrs.flags | abi.MO_QP_FLAG_PACKED_G | abi.MO_QP_FLAG_UNIFORM_BOX)
        out = abi.QPOutArrays(vars=None, status=np.zeros(B, dtype=np.int32), iters=np.zeros(B, dtype=np.int32),
                              kkt_final=np.zeros((B, 4)), x=np.zeros((B, n)))
        qp.solve_batch_host(compact, prm, out, 0)
        assert np.array_equal(out.status, want.status) and np.array_equal(out.iters, want.iters)
        assert np.array_equal(out.x, want.vars[:, :n])
        np.testing.assert_array_equal(out.kkt_final, want.kkt_final)
    # the device entry point refuses the host-only formats
    import torch
    d = qp.to_device(compact, dev)
    o = qp.alloc_out_torch(B, n, k, 2 * n, dev)
    from mini_opt_b200 import lib
    with pytest.raises(lib.MiniOptError) as e:
        qp.solve_batch_device(d, prm, o, dev)
    assert e.value.code == abi.MO_ERR_INVALID_ARGUMENT


def test_host_entry_multi_device(dev):
    """mo_qp_ip_solve_batch_host_multi: the batch sharded over the visible devices from one process equals the
    single-device result (no collective; each shard lands in its slice of the caller's arrays)."""
    import torch
    from mini_opt_b200 import qp
    ndev = torch.cuda.device_count()
    devices = list(range(min(ndev, 4))) if ndev > 1 else [0, 0]  # one GPU: two shards queue on the same device
    B, n, k = 30001, 8, 2
    arrs = synth.box_qp_batch_numpy(n, k, B, seed=909, explicit_constraints=False)
    prm = synth.solve_params()
    one = abi.alloc_out_numpy(B, n, k, 2 * n)
    qp.solve_batch_host(arrs, prm, one, 0)
    many = abi.alloc_out_numpy(B, n, k, 2 * n)
    qp.solve_batch_host_multi(arrs, prm, many, devices)
    assert np.array_equal(one.status, many.status) and np.array_equal(one.iters, many.iters)
    np.testing.assert_array_equal(one.vars, many.vars)


def test_nullspace_solver_vs_oracle(dev, oracle):
    """mo_qp_nullspace_solve_batch (QPNullSpaceSolver, qp.cc:679-729) vs the C restatement on identical bytes, the
    closed-form case of TestNullSpaceSolver1 and the NOT_POSITIVE_DEFINITE status."""
    import torch
    from mini_opt_b200 import qp
    for n, k, B in ((6, 2, 3000), (12, 3, 2000), (16, 8, 1000), (64, 16, 200), (8, 8, 100), (3, 1, 100)):
        arrs = synth.box_qp_batch_numpy(n, k, B, seed=50 + n, explicit_constraints=False)
        want, ws = oracle.qp_nullspace("oracle", arrs)
        x, st = qp.nullspace_solve_device(qp.to_device(arrs, dev), dev)
        torch.cuda.synchronize()
        assert np.array_equal(st.cpu().numpy(), ws)
        np.testing.assert_allclose(x.cpu().numpy(), want, rtol=1e-9, atol=1e-11)
    semi = synth.box_qp_batch_numpy(6, 2, 64, seed=1, explicit_constraints=False)
    semi.G[:] = 0.0
    x, st = qp.nullspace_solve_device(qp.to_device(semi, dev), dev)
    assert (st.cpu().numpy() == 1).all() and (x.cpu().numpy() == 0).all()
    Am = np.array([[-2.0, 1.4], [2.2, -3.5]])
    bv = np.array([-0.8, 1.3])
    one = abi.QPBatchArrays(n=2, k=1, m=0, batch=1, G=np.ascontiguousarray((Am.T @ Am).T[None]),
                            c=np.ascontiguousarray((-Am.T @ bv)[None]), A_eq=np.array([[[0.0], [1.0]]]), b_eq=np.array([[-0.3]]))
    x, st = qp.nullspace_solve_device(qp.to_device(one, dev), dev)
    assert int(st[0]) == 0
    np.testing.assert_allclose(x.cpu().numpy()[0], [0.860859728506787, 0.3], atol=1e-12)
