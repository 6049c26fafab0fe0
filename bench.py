#!/usr/bin/env python
"""Benchmark of the hot path: batched interior point QP solves (BASELINE config 2 by default).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--n 8 --k 2 --batch 1048576]

A "step" is one pass of mo_qp_ip_solve_batch over one synthetic batch (SURVEY 8(d) recipe) of `--batch`
box-constrained QPs per GPU. Metric: IP Newton steps/s (sum of per-problem iteration counts / time), whole job.
  value     inputs resident in HBM, device time by CUDA events on the launch stream, max over ranks
  e2e       same batch through mo_qp_ip_solve_batch_host from PINNED HOST buffers (H2D + solve + D2H inside)
  roofline  algorithmic FP64 flops per launch (F(n,k,m) per Newton step) / kernel time, against an
            FP64 DFMA peak measured live on this GPU (MEASURED_PEAKS.json has no FP64 entry), plus the HBM view
  cpu_baseline  the reference's own QPInteriorPointSolver (oracle/_ref, unmodified sources) one solver per
            thread on all host cores, bounded sample of the same workload (rank 0, N=1 only)
`--impl reference` times that CPU implementation alone (rank 0) and prints the same JSON shape.
Multi-GPU: the batch shards by rank with no collective on the solve path (weak scaling: per-GPU batch fixed).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

from mini_opt_b200 import abi, synth  # noqa: E402

METRIC = "ip_newton_steps_per_sec"
UNIT = "steps/s"


def flops_per_step(n, k, m):
    """one factorization of the reduced KKT, one two-sided solve, one residual, O(m) work."""
    p = n + k
    return p ** 3 / 3 + 2 * p ** 2 + 2 * n ** 2 + 4 * n * k + 30 * m


def algorithmic_bytes_per_qp(n, k, m):
    """Compulsory HBM traffic per QP: inputs read once (full G as stored) + [x|s|y|z], status, iters, kkt written."""
    return 8 * (n * n + n + k * n + k + m) + 8 * (n + 2 * m + k) + 8 + 32


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self.stop, self.th = index, [], threading.Event(), None

    def _run(self):
        while not self.stop.is_set():
            try:
                r = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                    "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if r.returncode == 0 and r.stdout.strip():
                    self.samples.append([f.strip() for f in r.stdout.strip().splitlines()[0].split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for s in self.samples for i in range(4) if len(s) > 2 + i and s[2 + i] == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples)}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference_cpu(n, k, sample, threads, seed=1234, repeats=1):
    """Time the UNMODIFIED reference (oracle/_ref) on `sample` problems with `threads` host threads."""
    from oracle import oracle_py as O  # the one place bench.py touches oracle/: CPU baseline / reference arm
    kind = "reference" if O.REF_SO.exists() else "port"
    arrs = synth.box_qp_batch_numpy(n, k, sample, seed=seed, explicit_constraints=False)
    prm = synth.solve_params()
    best, steps, solved = None, 0, 0
    for _ in range(repeats):
        t0 = time.perf_counter()
        out, steps = O.qp_solve("ref" if kind == "reference" else "oracle", arrs, prm,
                                nthreads=threads if kind == "reference" else 1)
        dt = time.perf_counter() - t0
        solved = int((out.status == 0).sum())
        best = dt if best is None or dt < best else best
    return dict(kind=kind, seconds=best, steps=steps, solved=solved, cores=threads if kind == "reference" else 1,
                sample=f"{sample} QPs n={n} k={k} m={2 * n} (same recipe, seed {seed}), "
                       f"{'one QPInteriorPointSolver per std::thread' if kind == 'reference' else 'scalar C port'}")


def pinned_copy(L, arr):
    """Copy a numpy array into pinned host memory obtained from the library; returns (ndarray view, ptr)."""
    import ctypes as C
    ptr = L.mo_host_alloc(arr.nbytes)
    if not ptr:
        raise RuntimeError("mo_host_alloc failed")
    buf = (C.c_char * arr.nbytes).from_address(ptr)
    view = np.frombuffer(buf, dtype=arr.dtype).reshape(arr.shape)
    view[...] = arr
    return view, ptr


def reference_arm(args, rank, world):
    """--impl reference: the reference's CPU implementation, all host threads, bounded sample per step."""
    if rank != 0:
        return
    threads = host_cores()
    sample = args.cpu_sample
    run_reference_cpu(args.n, args.k, min(sample, 4096), threads)  # page-in / thread start
    times, steps_per = [], 0
    res = None
    for i in range(args.warmup + args.steps):
        res = run_reference_cpu(args.n, args.k, sample, threads, seed=1234 + i)
        if i >= args.warmup:
            times.append(res["seconds"])
            steps_per += res["steps"]
    total = sum(times)
    value = steps_per / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args), "n": args.n, "k": args.k, "m": 2 * args.n,
                   "cpu_sample_per_step": sample, "timing": "host wall clock around the threaded solve"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": res["cores"], "kind": res["kind"],
                         "sample": res["sample"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_name(args):
    return (f"config2-like: {args.batch} box-constrained QPs per GPU, n={args.n}, k={args.k}, m={2 * args.n}, "
            f"SOLVE_EQUALITY_CONSTRAINED guess, sigma=0.1, kkt_tol=1e-9, max 30 iterations")


def ik_main(args, rank, world, local):
    """BASELINE config 3: batch of transform-chain IK problems through mo_nls_solve_chain_ik_batch (device SQP + LM loop).
    Metric: IP Newton steps/s inside the NLS solves (sum of QP iterations / time) and NLS problems/s."""
    import torch
    import torch.distributed as dist
    from mini_opt_b200 import lib, nls
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib.load()
    B = args.batch if args.batch != (1 << 20) else 100000
    prm = synth.chain_ik_params()
    if args.impl == "reference":
        if rank == 0:
            from oracle import oracle_py as O
            sample = min(B, 20000)
            ch = synth.chain_ik_batch_numpy(args.links, sample, seed=4321)
            threads = host_cores()
            O.nls_solve_chain("ref", synth.chain_ik_batch_numpy(args.links, 512, seed=1), prm, nthreads=threads)
            t0 = time.perf_counter()
            out, total = O.nls_solve_chain("ref", ch, prm, nthreads=threads)
            dt = time.perf_counter() - t0
            print(json.dumps({"impl": "reference", "metric": METRIC, "value": total / dt, "unit": UNIT, "n_gpus": args.gpus,
                              "steps": 1, "warmup": 0, "ms_per_step": 1e3 * dt, "higher_is_better": True, "scaling": "weak",
                              "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                              "config": {"workload": f"config3: {sample} chain-IK problems, {args.links} links (CPU sample)"},
                              "nls_problems_per_sec": sample / dt,
                              "cpu_baseline": {"value": total / dt, "unit": UNIT, "cores": threads, "kind": "reference",
                                               "sample": f"{sample} problems, one ConstrainedNonlinearLeastSquares per std::thread"},
                              "e2e": {"value": total / dt, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                              "gpu_launches": 0}), flush=True)
        return
    ch = nls.chain_to_device(synth.chain_ik_batch_numpy(args.links, B, seed=4321 + rank), dev)
    out = nls.alloc_nls_out_torch(B, ch.n, dev)
    stream = torch.cuda.current_stream(dev)
    for _ in range(max(args.warmup, 3)):
        nls.solve_chain_ik_device(ch, prm, out, dev, stream)
    torch.cuda.synchronize()
    l0 = lib.launch_count()
    if world > 1:
        dist.barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        t0.record(stream)
        for _ in range(args.steps):
            nls.solve_chain_ik_device(ch, prm, out, dev, stream)
        t1.record(stream)
        torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / args.steps
    qp_steps = float(out.qp_iterations.sum().item())
    ok = float((out.termination == abi.MO_NLS_SATISFIED_ABSOLUTE_TOL).sum().item())
    stats = torch.tensor([ms, qp_steps, ok], dtype=torch.float64, device=dev)
    if world > 1:
        mx, sm_ = stats.clone(), stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(sm_, op=dist.ReduceOp.SUM)
        ms, qp_steps, ok = float(mx[0]), float(sm_[1]), float(sm_[2])
    nls_iters = float(out.iterations.sum().item())
    # algorithmic flops: F(n, k, m) per inner Newton step + per outer iteration the chain FK / Jacobians
    # (~300 per link) and the linearization R N (N + 1) + 2 R N + N with R = cost rows
    n_, m_, R_ = ch.n, ch.m, ch.n_cost
    flops_local = qp_steps * flops_per_step(n_, ch.n_eq, m_) + nls_iters * (300.0 * args.links + R_ * n_ * (n_ + 1) + 2 * R_ * n_ + n_)
    fl = torch.tensor([flops_local], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(fl, op=dist.ReduceOp.SUM)
    fp64_peak = lib.measure_fp64_peak(local)
    e2e = None
    if not args.no_e2e:
        pipe = nls.ChainIKHostPipeline(synth.chain_ik_batch_numpy(args.links, B, seed=4321 + rank), dev)
        for _ in range(2):
            pipe.solve(prm)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        n_e2e = max(2, min(args.steps, 3))
        ts = time.perf_counter()
        for _ in range(n_e2e):
            hres = pipe.solve(prm)
        sec = (time.perf_counter() - ts) / n_e2e
        assert int(hres["qp_iterations"].sum()) == int(out.qp_iterations.sum().item()), "host pipeline disagrees with device path"
        tsec = torch.tensor([sec], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tsec, op=dist.ReduceOp.MAX)
        e2e = {"value": qp_steps / float(tsec[0]), "unit": UNIT, "h2d_bytes_per_step": int(pipe.h2d_bytes) * world,
               "d2h_bytes_per_step": int(pipe.d2h_bytes) * world, "ms_per_step": 1e3 * float(tsec[0]),
               "nls_problems_per_sec": B * world / float(tsec[0])}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import oracle_py as O  # bench cpu_baseline leg: the unmodified reference, timed as the checker's twin
        sample = min(B, 20000)
        chs = synth.chain_ik_batch_numpy(args.links, sample, seed=4321)
        threads = host_cores()
        O.nls_solve_chain("ref", synth.chain_ik_batch_numpy(args.links, 512, seed=1), prm, nthreads=threads)
        tc = time.perf_counter()
        _, total = O.nls_solve_chain("ref", chs, prm, nthreads=threads)
        dtc = time.perf_counter() - tc
        cpu = {"value": total / dtc, "unit": UNIT, "cores": threads, "kind": "reference",
               "sample": f"{sample} problems of the same recipe, one ConstrainedNonlinearLeastSquares per std::thread",
               "nls_problems_per_sec": sample / dtc, "seconds": dtc}
    if rank == 0:
        tfl = float(fl[0]) / (ms * 1e-3) / 1e12
        roofline = {"bound": "tensor", "pipe": "fp64 (DFMA; the n <= 12 systems have no 8x8x4 contraction for DMMA)",
                    "achieved": tfl, "peak": fp64_peak, "unit": "TFLOP/s", "frac": tfl / fp64_peak, "traffic": None,
                    "kernel": "nls_chain_ik_kernel", "kernel_ms": ms,
                    "algorithmic_flops_per_launch": float(fl[0]),
                    "peak_source": "DFMA microbenchmark measured live on this GPU (mo_measure_fp64_peak)",
                    "note": "one warp per problem, dependent scalar chains: latency-bound, see DESIGN.md 3.2"}
        print(json.dumps({"roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
                          "metric": METRIC, "value": qp_steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                          "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                          "config": {"workload": f"config3: {B} transform-chain IK problems per GPU, {args.links} Z-revolute links, "
                                                 "2 cost rows, joint limits +-pi/2, ARMIJO, max_qp_iterations=20, ModPi retraction",
                                     "links": args.links, "batch_per_gpu": B},
                          "nls_problems_per_sec": B * world / (ms * 1e-3), "absolute_tol_fraction": ok / (B * world),
                          "gpu_launches": int(lib.launch_count() - l0), "clocks": clocks.summary()}), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=8)
    ap.add_argument("--k", type=int, default=2)
    ap.add_argument("--batch", type=int, default=1 << 20, help="problems per GPU")
    ap.add_argument("--cpu-sample", type=int, default=262144)
    ap.add_argument("--workload", default="qp", choices=["qp", "ik"],
                    help="qp: BASELINE config 2/4 (default); ik: config 3, transform-chain IK via the device SQP+LM loop")
    ap.add_argument("--links", type=int, default=6, help="--workload ik: chain links (6-12)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "ik":
        ik_main(args, rank, world, local)
        return
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from mini_opt_b200 import lib, qp

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = lib.load()
    n, k, m, B = args.n, args.k, 2 * args.n, args.batch
    prm = synth.solve_params()
    arrs = synth.box_qp_batch_torch(n, k, B, seed=1234 + rank, device=dev)
    out = qp.alloc_out_torch(B, n, k, m, dev, lagrange=False)
    stream = torch.cuda.current_stream(dev)
    fp64_peak = lib.measure_fp64_peak(local)

    def step():
        qp.solve_batch_device(arrs, prm, out, dev, stream)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    launches0 = lib.launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    with ClockSampler(local) as clocks:
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record(stream)
        for a, b in evs:
            a.record(stream)
            step()
            b.record(stream)
        t1.record(stream)
        torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    launches = lib.launch_count() - launches0
    total_ms = t0.elapsed_time(t1)
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in evs]))
    steps_local = int(out.iters.sum().item())
    solved_local = int((out.status == 0).sum().item())
    stats = torch.tensor([total_ms, float(steps_local), float(solved_local)], dtype=torch.float64, device=dev)
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        total_ms, newton_steps, solved = float(mx[0]), float(sm[1]), float(sm[2])
    else:
        newton_steps, solved = float(steps_local), float(solved_local)
    sec_per_step = total_ms * 1e-3 / args.steps
    value = newton_steps / sec_per_step

    # roofline of the dominant (only) kernel on this rank
    F = flops_per_step(n, k, m)
    tflops = F * steps_local / (kernel_ms * 1e-3) / 1e12
    alg_bytes = algorithmic_bytes_per_qp(n, k, m) * B
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    # ncu dram__bytes_read.sum + dram__bytes_write.sum per pass of the default workload (profiles/r01_tpp_n8_ncu_full_summary.txt:
    # guess kernel 0.894 + 0.328 GB, iterate kernel 1.247 + 0.364 GB); other shapes: not captured -> null
    # dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel from `ncu --set full`, scaled to this launch:
    # n=8: profiles/r01_tpp_n8_ncu_full_summary.txt; n=64: profiles/r01_wd64_ncu_full_summary.txt (188.7 + 12.4 MB / 6216 QPs,
    # below the algorithmic bytes because only the lower block triangle of G is fetched; re-reads per Newton step hit L2)
    traffic = {(8, 2, 1 << 20): 2.833e9, (64, 16, 65536): 201.06e6 / 6216 * 65536}.get((n, k, B))
    roofline = {
        "bound": "tensor", "pipe": "fp64: DFMA and DMMA.8x8x4 share one 64 FMA/clk/SM datapath (profiles/r01_dmma_vs_dfma_microbench.txt)",
        "traffic": traffic, "achieved": tflops, "peak": fp64_peak, "unit": "TFLOP/s", "frac": tflops / fp64_peak,
        "kernel": lib.kernel_variant(n, k, m, arrs.flags), "kernel_ms": kernel_ms,
        "flops_per_newton_step": F, "newton_steps_per_launch": steps_local,
        "peak_source": "DFMA microbenchmark measured live on this GPU (mo_measure_fp64_peak)",
        "hbm": {"achieved": alg_bytes / (kernel_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                "frac": alg_bytes / (kernel_ms * 1e-3) / 1e9 / hbm_peak, "algorithmic_bytes_per_launch": alg_bytes,
                "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
    }

    # e2e: pinned host buffers through the host entry point (H2D + solve + D2H inside the timed region)
    e2e = None
    if not args.no_e2e:
        host = synth.to_numpy(arrs)
        keep = []
        def pin(a):
            v, p = pinned_copy(L, a)
            keep.append(p)
            return v
        harrs = abi.QPBatchArrays(n=n, k=k, m=m, batch=B, G=pin(host.G), c=pin(host.c), A_eq=pin(host.A_eq) if k else None,
                                  b_eq=pin(host.b_eq) if k else None, ineq_b=pin(host.ineq_b), flags=host.flags)
        hout = abi.QPOutArrays(vars=pin(np.zeros((B, n + 2 * m + k))), status=pin(np.zeros(B, dtype=np.int32)),
                               iters=pin(np.zeros(B, dtype=np.int32)), kkt_final=pin(np.zeros((B, 4))))
        h2d = sum(a.nbytes for a in (harrs.G, harrs.c, harrs.A_eq, harrs.b_eq, harrs.ineq_b) if a is not None)
        d2h = hout.vars.nbytes + hout.status.nbytes + hout.iters.nbytes + hout.kkt_final.nbytes
        for _ in range(2):
            qp.solve_batch_host(harrs, prm, hout, local)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t_start = time.perf_counter()
        e2e_steps = max(3, min(args.steps, 5))
        for _ in range(e2e_steps):
            qp.solve_batch_host(harrs, prm, hout, local)
        torch.cuda.synchronize()
        e2e_sec = (time.perf_counter() - t_start) / e2e_steps
        ts = torch.tensor([e2e_sec], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ts, op=dist.ReduceOp.MAX)
        assert np.array_equal(hout.iters, out.iters.cpu().numpy()), "host entry point disagrees with device path"
        e2e = {"value": newton_steps / float(ts[0]), "unit": UNIT, "h2d_bytes_per_step": int(h2d) * world,
               "d2h_bytes_per_step": int(d2h) * world, "ms_per_step": 1e3 * float(ts[0]),
               "qps_per_sec": B * world / float(ts[0])}
        for p in keep:
            L.mo_host_free(p)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = host_cores()
        run_reference_cpu(n, k, 4096, threads)
        r = run_reference_cpu(n, k, args.cpu_sample, threads)
        cpu = {"value": r["steps"] / r["seconds"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
               "sample": r["sample"], "qps_per_sec": r["solved"] / r["seconds"], "seconds": r["seconds"]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": 1e3 * sec_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "n": n, "k": k, "m": m, "batch_per_gpu": B,
                       "l2": "inputs (%.0f MB per pass) exceed the 126 MB L2" % (alg_bytes / 1e6),
                       "newton_steps_per_qp": newton_steps / (B * world)},
            "solved_qps_per_sec": solved / sec_per_step, "converged_fraction": solved / (B * world),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clocks.summary(),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
