"""TEST / MEASUREMENT INFRASTRUCTURE: parity of the CUDA path against the reference on IDENTICAL bytes.

north_star: relative 1e-9 on x and on the KKT residual norms, identical termination status and iteration count.
`compare_qp` turns one (GPU outputs, reference outputs) pair into the fractions bench.py prints in its `parity`
object and that the GPU tests pin (tests/test_gpu_qp.py). Problems whose equality-constrained start point is clamped
onto a bound (s = 1e-9, z = 1e9; qp.cc:472-481) are reported as their own partition: that is where the trajectories
are rounding-chaotic in the reference itself.

CLI (GPU box):  python tools/parity.py --n 8 --k 2 --batch 1048576 [--triplets] [--guess naive]
                python tools/parity.py --ik --links 6 --batch 100000 [--generic]
Imports oracle/ (the checker) — never imported by the product package.
"""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

from mini_opt_b200 import abi, synth  # noqa: E402


def rel_x_err(got_vars, ref_vars, n):
    d = np.abs(got_vars[:, :n] - ref_vars[:, :n]).max(axis=1)
    return d / np.maximum(1.0, np.abs(ref_vars[:, :n]).max(axis=1))


def clamped_start_mask(arrs: abi.QPBatchArrays, guess: int) -> np.ndarray:
    """True where ComputeInitialGuess (qp.cc:439-482) clamps x onto a bound (so some s starts at 1e-9)."""
    n, k, B = arrs.n, arrs.k, arrs.batch
    if guess != abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED:
        x0 = np.zeros((B, n))
    else:
        G = np.asarray(arrs.G)
        Gs = np.tril(G.transpose(0, 2, 1))  # G[p] is column-major: lower triangle of the matrix = upper of the C array
        Gs = Gs + np.tril(Gs, -1).transpose(0, 2, 1)
        K = np.zeros((B, n + k, n + k))
        K[:, :n, :n] = Gs
        if k:
            A = np.asarray(arrs.A_eq).transpose(0, 2, 1)  # [B, k, n]
            K[:, n:, :n] = A
            K[:, :n, n:] = A.transpose(0, 2, 1)
        rhs = np.concatenate([-np.asarray(arrs.c)] + ([-np.asarray(arrs.b_eq)] if k else []), axis=1)
        x0 = np.empty((B, n))
        for lo in range(0, B, 1 << 16):
            hi = min(B, lo + (1 << 16))
            x0[lo:hi] = np.linalg.solve(K[lo:hi], rhs[lo:hi, :, None])[:, :n, 0]
    ib = np.asarray(arrs.ineq_b)
    if arrs.flags & abi.MO_QP_FLAG_BOX_PAIRS:
        lo_, hi_ = -ib[:, 0::2], ib[:, 1::2]
        return ((x0 <= lo_) | (x0 >= hi_)).any(axis=1)
    # generic triplets: a x_v + b <= 1e-9 at the start
    var, a = np.asarray(arrs.ineq_var), np.asarray(arrs.ineq_a)
    val = a * np.take_along_axis(x0, var.astype(np.int64), axis=1) + ib
    return (val <= 1e-9).any(axis=1)


def compare_qp(got: abi.QPOutArrays, ref: abi.QPOutArrays, n: int, clamped: np.ndarray | None = None) -> dict:
    st = got.status == ref.status
    it = got.iters == ref.iters
    both = st & it
    rel = rel_x_err(got.vars, ref.vars, n)
    conv = (got.status == 0) & (ref.status == 0)
    res = {
        "problems": int(len(st)),
        "status_equal": float(st.mean()),
        "iters_equal": float(it.mean()),
        "status_and_iters_equal": float(both.mean()),
        "x_rel_1e-9": float(np.mean(rel[conv] < 1e-9)) if conv.any() else None,
        "x_rel_1e-9_same_trajectory": float(np.mean(rel[both & conv] < 1e-9)) if (both & conv).any() else None,
        "x_rel_max_converged": float(rel[conv].max()) if conv.any() else None,
        "converged_gpu": float((got.status == 0).mean()),
        "converged_ref": float((ref.status == 0).mean()),
        "newton_steps_gpu": int(got.iters.sum()),
        "newton_steps_ref": int(ref.iters.sum()),
    }
    if got.kkt_final is not None and ref.kkt_final is not None:
        kg, kr = got.kkt_final[both], ref.kkt_final[both]
        if len(kg):
            # final norms sit between the rounding floor and the tolerance: relative 1e-6 OR absolute 1e-9
            ok = np.abs(kg - kr) <= np.maximum(1e-6 * np.abs(kr), 1e-9)
            res["kkt_rel_1e-6_or_abs_1e-9"] = float(ok.all(axis=1).mean())
    mism = np.flatnonzero(~st)
    if len(mism):
        pairs = {}
        for g, r in zip(got.status[mism], ref.status[mism]):
            pairs[f"gpu{int(g)}_ref{int(r)}"] = pairs.get(f"gpu{int(g)}_ref{int(r)}", 0) + 1
        res["status_mismatch_pairs"] = pairs
    if clamped is not None:
        for name, mask in (("unclamped", ~clamped), ("clamped", clamped)):
            if mask.any():
                res[name] = {"problems": int(mask.sum()), "status_equal": float(st[mask].mean()),
                             "iters_equal": float(it[mask].mean()),
                             "converged_gpu": float((got.status[mask] == 0).mean()),
                             "converged_ref": float((ref.status[mask] == 0).mean())}
    return res


def compare_nls(got: abi.NLSOutArrays, ref: abi.NLSOutArrays) -> dict:
    term = got.termination == ref.termination
    same = term & (got.iterations == ref.iterations) & (got.qp_iterations == ref.qp_iterations)
    d = np.abs(got.x - ref.x).max(axis=1) / np.maximum(1.0, np.abs(ref.x).max(axis=1))
    res = {"problems": int(len(term)), "termination_equal": float(term.mean()),
           "termination_and_counts_equal": float(same.mean()),
           "x_rel_1e-9_same_trajectory": float(np.mean(d[same] < 1e-9)) if same.any() else None,
           "x_rel_1e-6_same_trajectory": float(np.mean(d[same] < 1e-6)) if same.any() else None,
           "qp_iterations_gpu": int(got.qp_iterations.sum()), "qp_iterations_ref": int(ref.qp_iterations.sum())}
    hist = {}
    for name, arr in (("gpu", got.termination), ("ref", ref.termination)):
        v, c = np.unique(arr, return_counts=True)
        hist[name] = {int(a): int(b) for a, b in zip(v, c)}
    res["termination_histogram"] = hist
    if got.errors is not None and ref.errors is not None:
        # trajectory-independent quality: final cost f no worse than the reference's (x1.001 + 1e-9)
        fg, fr = got.errors[:, 0], ref.errors[:, 0]
        res["final_cost_not_worse"] = float(np.mean(fg <= fr * 1.001 + 1e-9))
        res["final_cost_ratio_p99"] = float(np.quantile((fg + 1e-12) / (fr + 1e-12), 0.99))
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=8)
    ap.add_argument("--k", type=int, default=2)
    ap.add_argument("--batch", type=int, default=1 << 20)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--guess", default="eq", choices=["eq", "naive"])
    ap.add_argument("--triplets", action="store_true", help="explicit (var,a,b) constraints: generic kernel")
    ap.add_argument("--torch-gen", action="store_true", help="generate on the device with bench.py's generator")
    ap.add_argument("--which", default="ref", choices=["ref", "oracle"])
    ap.add_argument("--ik", action="store_true")
    ap.add_argument("--links", type=int, default=6)
    ap.add_argument("--n-eq", type=int, default=0)
    ap.add_argument("--generic", action="store_true", help="--ik: MO_NLS_QP_GENERIC inner solver (audit path)")
    ap.add_argument("--max-qp-iterations", type=int, default=20)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()

    import torch
    from mini_opt_b200 import lib, nls, qp
    from oracle import oracle_py as O
    lib.load()
    dev = torch.device("cuda:0")
    threads = len(__import__("os").sched_getaffinity(0))
    if args.ik:
        ch = synth.chain_ik_batch_numpy(args.links, args.batch, seed=args.seed, n_eq=args.n_eq)
        prm = synth.chain_ik_params(args.max_qp_iterations)
        if args.generic:
            prm.inner_qp_solver = 1
        out = nls.alloc_nls_out_torch(args.batch, ch.n, dev)
        nls.solve_chain_ik_device(nls.chain_to_device(ch, dev), prm, out, dev)
        torch.cuda.synchronize()
        got = nls.nls_out_to_numpy(out)
        prm_ref = synth.chain_ik_params(args.max_qp_iterations)
        t0 = time.perf_counter()
        ref, _ = O.nls_solve_chain(args.which, ch, prm_ref, nthreads=threads)
        res = compare_nls(got, ref)
        res.update(kind="ik", links=args.links, n_eq=args.n_eq, generic=args.generic, ref_seconds=time.perf_counter() - t0,
                   which=args.which)
    else:
        guess = abi.MO_GUESS_SOLVE_EQUALITY_CONSTRAINED if args.guess == "eq" else abi.MO_GUESS_NAIVE
        prm = synth.solve_params(guess)
        if args.torch_gen:
            darrs = synth.box_qp_batch_torch(args.n, args.k, args.batch, seed=args.seed, device=dev)
            arrs = synth.to_numpy(darrs)
        else:
            arrs = synth.box_qp_batch_numpy(args.n, args.k, args.batch, seed=args.seed, explicit_constraints=args.triplets)
            if args.triplets:
                arrs.flags = 0
            darrs = qp.to_device(arrs, dev)
        out = qp.alloc_out_torch(args.batch, args.n, args.k, 2 * args.n, dev)
        qp.solve_batch_device(darrs, prm, out, dev)
        torch.cuda.synchronize()
        got = qp.out_to_numpy(out)
        t0 = time.perf_counter()
        ref, _ = O.qp_solve(args.which, arrs, prm, nthreads=threads)
        dt = time.perf_counter() - t0
        res = compare_qp(got, ref, args.n, clamped_start_mask(arrs, guess))
        res.update(kind="qp", n=args.n, k=args.k, guess=args.guess, triplets=args.triplets, ref_seconds=dt,
                   which=args.which, variant=lib.kernel_variant(args.n, args.k, 2 * args.n, arrs.flags))
        bad = np.flatnonzero(got.status != ref.status)[:8]
        res["first_status_mismatches"] = [
            {"p": int(p), "gpu": [int(got.status[p]), int(got.iters[p])], "ref": [int(ref.status[p]), int(ref.iters[p])],
             "kkt_gpu": [float(v) for v in got.kkt_final[p]], "kkt_ref": [float(v) for v in ref.kkt_final[p]]} for p in bad]
    line = json.dumps(res)
    print(line, flush=True)
    if args.out:
        with open(args.out, "a") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
