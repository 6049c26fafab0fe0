#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by RUNNING THE REFERENCE (oracle/_ref, built from
/root/reference by oracle/Makefile). Run in the CPU container where /root/reference exists:

    python tools/make_golden.py

Fixtures (all float64 / int32, numpy .npz):
  qp_generated.npz   seeds 0..999 of the reference's own GenerateRandomQP(seed, 8, 0.5)
                     (test/qp_test.cc:483-524) + the reference's Solve() outputs for NAIVE and
                     SOLVE_EQUALITY_CONSTRAINED with the params of TestGeneratedProblems (:541-548).
  qp_fixed.npz       the literal QPs of test/qp_test.cc:168-471: reference Newton-step deltas on the dummy
                     state of the Elimination tests, and reference Solve() outputs of the fixed-answer tests.
  qp_box_synth.npz   small seeded samples of the SURVEY §8(d) box-QP distribution (n=8/32/64) with the
                     reference's outputs and per-iteration logs for both initial-guess methods.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from mini_opt_b200 import abi, synth  # noqa: E402
from oracle import oracle_py as O  # noqa: E402
import qp_cases  # noqa: E402

GOLD = ROOT / "tests" / "golden"


def generated_problems(count=1000, dim=8):
    G = np.zeros((count, dim, dim))
    c = np.zeros((count, dim))
    var = np.zeros((count, dim), dtype=np.int32)
    a = np.ones((count, dim))
    b = np.zeros((count, dim))
    cnt = np.zeros(count, dtype=np.int32)
    sol = np.zeros((count, dim))
    for seed in range(count):
        q = O.generate_random_qp(seed, dim, 0.5)
        G[seed] = q["G"].T  # column-major storage
        c[seed] = q["c"]
        cnt[seed] = q["count"]
        var[seed, :q["count"]] = q["var"]
        a[seed, :q["count"]] = q["a"]
        b[seed, :q["count"]] = q["b"]
        sol[seed] = q["solution"]
    arrs = abi.QPBatchArrays(n=dim, k=0, m=dim, batch=count, G=G, c=c, ineq_var=var, ineq_a=a, ineq_b=b,
                             ineq_count=cnt)
    save = dict(G=G, c=c, ineq_var=var, ineq_a=a, ineq_b=b, ineq_count=cnt, solution=sol)
    for name, guess in (("naive", abi.MO_GUESS_NAIVE), ("eq", abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED)):
        prm = abi.QPParams(termination_kkt_tol=1e-12, max_iterations=30, initial_guess_method=guess)
        out, total = O.qp_solve("ref", arrs, prm)
        save[f"{name}_vars"] = out.vars
        save[f"{name}_status"] = out.status
        save[f"{name}_iters"] = out.iters
        save[f"{name}_kkt"] = out.kkt_final
        print(f"generated[{name}]: total iterations {total}")
    np.savez_compressed(GOLD / "qp_generated.npz", **save)


def fixed_cases():
    save = {}
    for name, arrs, x_guess in qp_cases.elimination_cases():
        state = qp_cases.dummy_state(arrs.n, arrs.k, arrs.m, x_guess)
        res = O.qp_newton_step("ref", arrs, 0.0, abi.MO_BARRIER_COMPLEMENTARITY, state[None])
        assert res["status"][0] == 0
        save[f"elim_{name}_delta"] = res["delta"][0]
        save[f"elim_{name}_vars"] = res["vars"][0]
        save[f"elim_{name}_ip"] = res["ip"][0]
        # the reference test's own independent answer (full-system LU) must agree to 1e-12
        full = qp_cases.full_system_delta(arrs, state)
        assert np.abs(full - res["delta"][0]).max() < 1e-12, name
    for case in qp_cases.solve_cases():
        for guess in case["guesses"]:
            prm = abi.QPParams(**{**case["params"].__dict__, "initial_guess_method": guess})
            out, _ = O.qp_solve("ref", case["batch"], prm, log=True)
            qp_cases.check_solve_expectations(case, out, case["batch"])
            key = f"solve_{case['name']}_g{guess}"
            save[key + "_vars"] = out.vars[0]
            save[key + "_status"] = out.status
            save[key + "_iters"] = out.iters
            save[key + "_kkt"] = out.kkt_final[0]
            save[key + "_log"] = out.iter_log[0]
    np.savez_compressed(GOLD / "qp_fixed.npz", **save)
    print("fixed cases:", len(save), "arrays")


def box_synth():
    save = {}
    for n, k, B in ((8, 2, 256), (32, 8, 16), (64, 16, 4)):
        arrs = synth.box_qp_batch_numpy(n, k, B, seed=20260 + n)
        tag = f"n{n}"
        save.update({f"{tag}_G": arrs.G, f"{tag}_c": arrs.c, f"{tag}_A_eq": arrs.A_eq, f"{tag}_b_eq": arrs.b_eq,
                     f"{tag}_ineq_b": arrs.ineq_b})
        for gname, guess in (("naive", abi.MO_GUESS_NAIVE), ("eq", abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED)):
            out, total = O.qp_solve("ref", arrs, synth.solve_params(guess), log=True)
            save[f"{tag}_{gname}_vars"] = out.vars
            save[f"{tag}_{gname}_status"] = out.status
            save[f"{tag}_{gname}_iters"] = out.iters
            save[f"{tag}_{gname}_kkt"] = out.kkt_final
            save[f"{tag}_{gname}_lagrange"] = out.lagrange
            if n == 8:
                save[f"{tag}_{gname}_log"] = out.iter_log
            print(f"box_synth n={n} {gname}: {total} steps, status {np.bincount(out.status)}")
    np.savez_compressed(GOLD / "qp_box_synth.npz", **save)


def nls_golden():
    """tests/golden/nls_chain.npz: reference outputs for the NLS path (LinearizeAndFillQP, ActuatorChain::Update,
    ConstrainedNonlinearLeastSquares::Solve). Inputs are regenerated from seeds by tests/nls_cases.py / synth.py."""
    import nls_cases as N
    save = {}
    for seed in range(6):
        ch, P = N.random_general_chain(seed)
        fk = O.chain_fk("ref", ch, P)
        for k, v in fk.items():
            save[f"fk{seed}_{k}"] = v
    for i, (n, rc, cb, re) in enumerate(LINEARIZE_SHAPES):
        out = O.nls_linearize("ref", N.random_linearize(100 + i, n, rc, cb, re))
        for f in ("G", "c", "A_eq", "b_eq", "ineq_b_shifted", "errors"):
            save[f"lin{i}_{f}"] = getattr(out, f)
        assert np.all(np.triu(out.G.transpose(0, 2, 1), 1) == 0.0)  # strict upper triangle exactly 0 (residual_test.cc)
    ch, prm = N.two_angle_chain()
    out, total = O.nls_solve_chain("ref", ch, prm)
    fk = O.chain_fk("ref", ch, out.x)
    assert np.abs(fk["translation"][:, :2] - np.array([0.45, 0.6])).max() < 1e-3  # nonlinear_test.cc:949
    assert out.ls_steps.max() < 100                                                # nonlinear_test.cc:956
    for f in ("x", "termination", "iterations", "qp_iterations", "ls_steps", "errors"):
        save[f"two_angle_{f}"] = getattr(out, f)
    print(f"two_angle: {total} QP iterations, termination {np.bincount(out.termination)}")
    for tag, links, neq, qpi, B in IK_SETS:
        ch = synth.chain_ik_batch_numpy(links, B, seed=777 + links, n_eq=neq)
        out, total = O.nls_solve_chain("ref", ch, synth.chain_ik_params(qpi))
        for f in ("x", "termination", "iterations", "qp_iterations", "ls_steps", "errors"):
            save[f"{tag}_{f}"] = getattr(out, f)
        print(f"{tag}: {total} QP iterations, termination {np.bincount(out.termination, minlength=5)}")
    np.savez_compressed(GOLD / "nls_chain.npz", **save)


LINEARIZE_SHAPES = ((6, 3, 3, 0), (8, 4, 2, 2), (12, 6, 1, 3), (5, 2, 0, 1))
IK_SETS = (("ik6", 6, 0, 20, 256), ("ik12", 12, 0, 20, 128), ("ik6q5", 6, 0, 5, 256), ("ik8eq", 8, 1, 20, 128))


if __name__ == "__main__":
    GOLD.mkdir(parents=True, exist_ok=True)
    O.build()
    generated_problems()
    fixed_cases()
    box_synth()
    nls_golden()
    for f in sorted(GOLD.glob("*.npz")):
        print(f.name, f.stat().st_size, "bytes")
