#!/usr/bin/env python
"""Benchmark of the hot path: batched interior point QP solves (BASELINE config 2 by default).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--n 8 --k 2 --batch 1048576]
  python bench.py --sweep            # BASELINE config 5: one JSON line per (n, batch), at the --gpus given
  python bench.py --workload ik      # BASELINE config 3 as the main line

A "step" is one pass of mo_qp_ip_solve_batch over one synthetic batch (SURVEY 8(d) recipe) of `--batch`
box-constrained QPs per GPU. Metric: IP Newton steps/s (sum of per-problem iteration counts / time), whole job.
  value     inputs resident in HBM, device time by CUDA events on the launch stream, max over ranks
  e2e       same batch through mo_qp_ip_solve_batch_host from PINNED HOST buffers (H2D + solve + D2H inside)
  roofline  algorithmic FP64 flops per launch (F(n,k,m) per Newton step) / kernel time, against the
            FP64 peak measured live on this GPU: the DFMA peak for the n = 8 kernels, the DMMA.8x8x4 peak for the kernels
            whose O(p^3) work runs on DMMA (MEASURED_PEAKS.json has no FP64 entry), plus the HBM view
  cpu_baseline  the reference's own QPInteriorPointSolver (oracle/_ref, unmodified sources), one solver per std::thread
            on all host cores, on the FIRST `--cpu-sample` problems of the GPU's own batch (rank 0, N=1 only)
  parity    GPU results vs the reference's on those same bytes: identical status / iteration count fractions, x to 1e-9
  also      the same record for BASELINE config 4 (64K QPs n=64, DMMA kernel) and config 3 (100K 6-link chain IK),
            measured in the same process after the main line
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
sys.path.insert(0, str(ROOT / "tools"))

from mini_opt_b200 import abi, synth  # noqa: E402

E2E_FORMAT = "compact"  # --e2e-format
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


class Ctx:
    """Rank / device plumbing shared by the workloads (torch = device memory, streams, torch.distributed only)."""

    def __init__(self):
        import torch
        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        self.dev = None

    def init_gpu(self):
        torch = self.torch
        assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        from mini_opt_b200 import lib
        self.lib = lib
        self.L = lib.load()
        self.peaks = {}
        try:
            self.peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except Exception:
            pass
        self.dfma_peak = lib.measure_fp64_peak(self.local)
        self.dmma_peak = lib.measure_dmma_peak(self.local)

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def reduce(self, vals, op):
        """all-reduce a list of python floats (MAX or SUM) over ranks: timing scalars only, never problem data."""
        if self.dist is None:
            return list(vals)
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return [float(v) for v in t]

    def finish(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def gen_qp_batch(n, k, batch, seed, dev):
    """The synthetic batch of SURVEY 8(d). Generated on the device when there is one (bench: straight into HBM; the
    reference arm copies it back so both arms see the same stream), else with numpy."""
    if dev is not None:
        return synth.box_qp_batch_torch(n, k, batch, seed=seed, device=dev)
    return synth.box_qp_batch_numpy(n, k, batch, seed=seed, explicit_constraints=False)


def run_reference_qp(arrs_host, prm, threads):
    """Time the UNMODIFIED reference (oracle/_ref) on a host batch. The timed region is the threaded solve only:
    unpacking the flat arrays into mini_opt::QP objects and allocating outputs happen before it."""
    from oracle import oracle_py as O  # bench.py touches oracle/ only here: CPU baseline / reference arm / parity checker
    kind = "reference" if O.REF_SO.exists() else "port"
    out, steps, seconds = O.qp_solve_timed("ref" if kind == "reference" else "oracle", arrs_host, prm,
                                           nthreads=threads if kind == "reference" else 1)
    return dict(kind=kind, seconds=seconds, steps=steps, solved=int((out.status == 0).sum()), out=out,
                cores=threads if kind == "reference" else 1)


def pinned_copy(L, arr, interleaved=False):
    """Copy a numpy array into pinned host memory obtained from the library; returns (ndarray view, ptr).
    interleaved: pages spread over the host's NUMA nodes (buffers that feed several GPUs from one process)."""
    import ctypes as C
    ptr = (L.mo_host_alloc_interleaved if interleaved else L.mo_host_alloc)(arr.nbytes)
    if not ptr:
        raise RuntimeError("mo_host_alloc failed")
    buf = (C.c_char * arr.nbytes).from_address(ptr)
    view = np.frombuffer(buf, dtype=arr.dtype).reshape(arr.shape)
    view[...] = arr
    return view, ptr


def qp_workload_name(n, k, B):
    return (f"{'config2' if n == 8 else 'config4' if n == 64 else 'config5'}-like: {B} box-constrained QPs per GPU, "
            f"n={n}, k={k}, m={2 * n}, SOLVE_EQUALITY_CONSTRAINED guess, sigma=0.1, kkt_tol=1e-9, max 30 iterations")


def traffic_record(kernel, n, k):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel per QP, from the committed `ncu --set full`
    capture named in profiles/traffic.json (source file and commit are recorded there, not hard-coded here)."""
    try:
        tab = json.loads((ROOT / "profiles" / "traffic.json").read_text())
        return tab.get(f"{kernel}:{n}:{k}")
    except Exception:
        return None


def qp_pass(ctx: Ctx, n, k, B, steps, warmup, with_e2e=True, cpu_sample=0, seed=1234):
    """One measured workload: device-resident value, roofline, e2e, and (N=1, rank 0) cpu_baseline + parity."""
    import parity as par
    torch, dev, lib, L = ctx.torch, ctx.dev, ctx.lib, ctx.L
    from mini_opt_b200 import qp
    m = 2 * n
    prm = synth.solve_params()
    arrs = gen_qp_batch(n, k, B, seed + ctx.rank, dev)
    out = qp.alloc_out_torch(B, n, k, m, dev, lagrange=False)
    stream = torch.cuda.current_stream(dev)

    def step():
        qp.solve_batch_device(arrs, prm, out, dev, stream)

    warmup = max(warmup, 3)
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    launches0 = lib.launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    ctx.barrier()
    torch.cuda.synchronize()
    with ClockSampler(ctx.local) as clocks:
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record(stream)
        for a, b in evs:
            a.record(stream)
            step()
            b.record(stream)
        t1.record(stream)
        torch.cuda.synchronize()
    ctx.barrier()
    launches = lib.launch_count() - launches0
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in evs]))
    steps_local = int(out.iters.sum().item())
    solved_local = int((out.status == 0).sum().item())
    total_ms = ctx.reduce([t0.elapsed_time(t1)], "max")[0]
    newton_steps, solved = ctx.reduce([float(steps_local), float(solved_local)], "sum")
    sec_per_step = total_ms * 1e-3 / steps
    value = newton_steps / sec_per_step

    # roofline of the dominant kernel on this rank
    F = flops_per_step(n, k, m)
    tflops = F * steps_local / (kernel_ms * 1e-3) / 1e12
    alg_bytes = algorithmic_bytes_per_qp(n, k, m) * B
    hbm_peak = float(ctx.peaks.get("hbm_gbs", 6650.0))
    variant = lib.kernel_variant(n, k, m, arrs.flags)
    dmma = variant.startswith("dmma")
    peak = ctx.dmma_peak if dmma else ctx.dfma_peak
    tr = traffic_record(variant, n, k)
    roofline = {
        "bound": "tensor",
        "pipe": ("fp64 tensor path: mma.sync.m8n8k4.f64 = DMMA.8x8x4" if dmma else "fp64 DFMA (no 8x8x4 contraction at this size)"),
        "traffic": (tr["bytes_per_qp"] * B if tr else None), "traffic_source": (tr["source"] if tr else None),
        "achieved": tflops, "peak": peak, "unit": "TFLOP/s", "frac": tflops / peak,
        "kernel": variant, "kernel_ms": kernel_ms, "flops_per_newton_step": F, "newton_steps_per_launch": steps_local,
        "peak_source": ("DMMA.8x8x4" if dmma else "DFMA") + " microbenchmark measured live on this GPU "
                       "(mo_measure_dmma_peak / mo_measure_fp64_peak); MEASURED_PEAKS.json has no FP64 entry",
        "dfma_peak": ctx.dfma_peak, "dmma_peak": ctx.dmma_peak,
        "hbm": {"achieved": alg_bytes / (kernel_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                "frac": alg_bytes / (kernel_ms * 1e-3) / 1e9 / hbm_peak, "algorithmic_bytes_per_launch": alg_bytes,
                "peak_source": "MEASURED_PEAKS.json" if ctx.peaks else "fallback"},
    }

    # e2e: pinned host buffers through the host entry point (H2D + solve + D2H inside the timed region)
    e2e = None
    if with_e2e:
        host = synth.to_numpy(arrs)
        keep = []

        def pin(a):
            v, p = pinned_copy(L, a)
            keep.append(p)
            return v
        if E2E_FORMAT == "compact":
            # what a host caller holding Eigen objects hands over: the lower triangle of G (all the reference reads,
            # qp.cc:289,404), one box for the batch, and x_block() + status + iteration count + KKT norms back
            assert (host.ineq_b == host.ineq_b[0]).all()
            harrs = abi.QPBatchArrays(n=n, k=k, m=m, batch=B, G=pin(qp.pack_lower(host.G)), c=pin(host.c),
                                      A_eq=pin(host.A_eq) if k else None, b_eq=pin(host.b_eq) if k else None,
                                      ineq_b=pin(np.ascontiguousarray(host.ineq_b[0])),
                                      flags=host.flags | abi.MO_QP_FLAG_PACKED_G | abi.MO_QP_FLAG_UNIFORM_BOX)
            hout = abi.QPOutArrays(vars=None, status=pin(np.zeros(B, dtype=np.int32)), iters=pin(np.zeros(B, dtype=np.int32)),
                                   kkt_final=pin(np.zeros((B, 4))), x=pin(np.zeros((B, n))))
        else:
            harrs = abi.QPBatchArrays(n=n, k=k, m=m, batch=B, G=pin(host.G), c=pin(host.c), A_eq=pin(host.A_eq) if k else None,
                                      b_eq=pin(host.b_eq) if k else None, ineq_b=pin(host.ineq_b), flags=host.flags)
            hout = abi.QPOutArrays(vars=pin(np.zeros((B, n + 2 * m + k))), status=pin(np.zeros(B, dtype=np.int32)),
                                   iters=pin(np.zeros(B, dtype=np.int32)), kkt_final=pin(np.zeros((B, 4))))
        h2d = sum(a.nbytes for a in (harrs.G, harrs.c, harrs.A_eq, harrs.b_eq, harrs.ineq_b) if a is not None)
        d2h = sum(a.nbytes for a in (hout.vars, hout.x, hout.status, hout.iters, hout.kkt_final) if a is not None)
        for _ in range(2):
            qp.solve_batch_host(harrs, prm, hout, ctx.local)
        ctx.barrier()
        torch.cuda.synchronize()
        t_start = time.perf_counter()
        e2e_steps = max(3, min(steps, 5))
        for _ in range(e2e_steps):
            qp.solve_batch_host(harrs, prm, hout, ctx.local)
        torch.cuda.synchronize()
        e2e_sec = ctx.reduce([(time.perf_counter() - t_start) / e2e_steps], "max")[0]
        assert np.array_equal(hout.iters, out.iters.cpu().numpy()), "host entry point disagrees with device path"
        if hout.x is not None:
            assert np.array_equal(hout.x, out.vars[:, :n].cpu().numpy()), "host entry point x disagrees with device path"
        e2e = {"format": ("packed lower-triangle G + one box for the batch in; x, status, iters, kkt_final out"
                          if E2E_FORMAT == "compact" else "full G, per-problem bounds in; [x|s|y|z], status, iters, kkt_final out"),
               "value": newton_steps / e2e_sec, "unit": UNIT, "h2d_bytes_per_step": int(h2d) * ctx.world,
               "d2h_bytes_per_step": int(d2h) * ctx.world, "ms_per_step": 1e3 * e2e_sec,
               "qps_per_sec": B * ctx.world / e2e_sec}
        for p in keep:
            L.mo_host_free(p)

    # cpu_baseline + parity: the reference on the first cpu_sample problems of THIS batch (same bytes)
    cpu, parity = None, None
    if ctx.rank == 0 and ctx.world == 1 and cpu_sample > 0:
        S = min(cpu_sample, B)
        threads = host_cores()
        sample = synth.to_numpy(arrs, 0, S)
        run_reference_qp(synth.to_numpy(arrs, 0, min(S, 2048)), prm, threads)  # page-in / thread start
        r = run_reference_qp(sample, prm, threads)
        cpu = {"value": r["steps"] / r["seconds"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
               "sample": f"first {S} QPs of the GPU's own batch (n={n} k={k} m={m}, same bytes), "
                         f"{'one QPInteriorPointSolver per std::thread, solve loop only' if r['kind'] == 'reference' else 'scalar C port'}, "
                         "g++ -O2 -march=x86-64-v3 (prebuilt in the CPU container; /root/reference is absent on the GPU box)",
               "qps_per_sec": r["solved"] / r["seconds"], "seconds": r["seconds"]}
        got = abi.QPOutArrays(vars=out.vars[:S].cpu().numpy(), status=out.status[:S].cpu().numpy(),
                              iters=out.iters[:S].cpu().numpy(), kkt_final=out.kkt_final[:S].cpu().numpy())
        parity = par.compare_qp(got, r["out"], n, par.clamped_start_mask(sample, prm.initial_guess_method))
        parity["against"] = r["kind"]

    return {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": ctx.world, "steps": steps, "warmup": warmup,
        "ms_per_step": 1e3 * sec_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": qp_workload_name(n, k, B), "n": n, "k": k, "m": m, "batch_per_gpu": B,
                   "l2": "inputs (%.0f MB per pass) %s the 126 MB L2" % (alg_bytes / 1e6, "exceed" if alg_bytes > 126e6 else "fit in"),
                   "newton_steps_per_qp": newton_steps / (B * ctx.world)},
        "solved_qps_per_sec": solved / sec_per_step, "converged_fraction": solved / (B * ctx.world),
        "roofline": roofline, "cpu_baseline": cpu, "parity": parity, "e2e": e2e, "gpu_launches": int(launches),
        "clocks": clocks.summary(),
    }


def ik_pass(ctx: Ctx, links, B, steps, warmup, with_e2e=True, cpu_sample=0):
    """BASELINE config 3: batch of transform-chain IK problems through mo_nls_solve_chain_ik_batch (device SQP + LM loop).
    Metric: IP Newton steps/s inside the NLS solves (sum of QP iterations / time) and NLS problems/s."""
    import parity as par
    torch, dev, lib = ctx.torch, ctx.dev, ctx.lib
    from mini_opt_b200 import nls
    prm = synth.chain_ik_params()
    host_ch = synth.chain_ik_batch_numpy(links, B, seed=4321 + ctx.rank)
    ch = nls.chain_to_device(host_ch, dev)
    out = nls.alloc_nls_out_torch(B, ch.n, dev)
    stream = torch.cuda.current_stream(dev)
    warmup = max(warmup, 3)
    for _ in range(warmup):
        nls.solve_chain_ik_device(ch, prm, out, dev, stream)
    torch.cuda.synchronize()
    l0 = lib.launch_count()
    ctx.barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(ctx.local) as clocks:
        t0.record(stream)
        for _ in range(steps):
            nls.solve_chain_ik_device(ch, prm, out, dev, stream)
        t1.record(stream)
        torch.cuda.synchronize()
    launches = lib.launch_count() - l0
    ms_local = t0.elapsed_time(t1) / steps
    qp_local = float(out.qp_iterations.sum().item())
    ok_local = float((out.termination == abi.MO_NLS_SATISFIED_ABSOLUTE_TOL).sum().item())
    nls_iters = float(out.iterations.sum().item())
    # algorithmic flops: F(n, k, m) per inner Newton step + per outer iteration the chain FK / Jacobians
    # (~300 per link) and the linearization R N (N + 1) + 2 R N + N with R = cost rows
    n_, m_, R_ = ch.n, ch.m, ch.n_cost
    flops_local = qp_local * flops_per_step(n_, ch.n_eq, m_) + nls_iters * (300.0 * links + R_ * n_ * (n_ + 1) + 2 * R_ * n_ + n_)
    ms = ctx.reduce([ms_local], "max")[0]
    qp_steps, ok, flops = ctx.reduce([qp_local, ok_local, flops_local], "sum")
    e2e = None
    if with_e2e:
        pipe = nls.ChainIKHostPipeline(host_ch, dev)
        for _ in range(2):
            pipe.solve(prm)
        ctx.barrier()
        torch.cuda.synchronize()
        n_e2e = max(2, min(steps, 3))
        ts = time.perf_counter()
        for _ in range(n_e2e):
            hres = pipe.solve(prm)
        sec = ctx.reduce([(time.perf_counter() - ts) / n_e2e], "max")[0]
        assert int(hres["qp_iterations"].sum()) == int(out.qp_iterations.sum().item()), "host pipeline disagrees with device path"
        e2e = {"value": qp_steps / sec, "unit": UNIT, "h2d_bytes_per_step": int(pipe.h2d_bytes) * ctx.world,
               "d2h_bytes_per_step": int(pipe.d2h_bytes) * ctx.world, "ms_per_step": 1e3 * sec,
               "nls_problems_per_sec": B * ctx.world / sec}
    cpu, parity = None, None
    if ctx.rank == 0 and ctx.world == 1 and cpu_sample > 0:
        from oracle import oracle_py as O  # cpu_baseline leg: the unmodified reference, also the parity checker
        S = min(B, cpu_sample)
        chs = synth.chain_slice(host_ch, 0, S)
        threads = host_cores()
        O.nls_solve_chain("ref", synth.chain_slice(host_ch, 0, min(S, 512)), prm, nthreads=threads)
        tc = time.perf_counter()
        ref, total = O.nls_solve_chain("ref", chs, prm, nthreads=threads)
        dtc = time.perf_counter() - tc
        cpu = {"value": total / dtc, "unit": UNIT, "cores": threads, "kind": "reference",
               "sample": f"first {S} problems of the GPU's own batch, one ConstrainedNonlinearLeastSquares per std::thread",
               "nls_problems_per_sec": S / dtc, "seconds": dtc}
        got = nls.nls_out_to_numpy(out)
        got = abi.NLSOutArrays(**{f: getattr(got, f)[:S] for f in ("x", "termination", "iterations", "qp_iterations", "ls_steps", "errors")})
        parity = par.compare_nls(got, ref)
        parity["against"] = "reference"
    tfl = flops / ctx.world / (ms * 1e-3) / 1e12
    roofline = {"bound": "tensor", "pipe": "fp64 (DFMA; the n <= 12 systems have no 8x8x4 contraction for DMMA)",
                "achieved": tfl, "peak": ctx.dfma_peak, "unit": "TFLOP/s", "frac": tfl / ctx.dfma_peak, "traffic": None,
                "kernel": "nls_chain_ik_kernel", "kernel_ms": ms, "algorithmic_flops_per_launch": flops / ctx.world,
                "peak_source": "DFMA microbenchmark measured live on this GPU (mo_measure_fp64_peak)",
                "note": "one warp per problem, dependent scalar chains: latency-bound, see DESIGN.md 3.2"}
    return {"metric": METRIC, "value": qp_steps / (ms * 1e-3), "unit": UNIT, "n_gpus": ctx.world, "steps": steps,
            "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"config3: {B} transform-chain IK problems per GPU, {links} Z-revolute links, "
                                   "2 cost rows, joint limits +-pi/2, ARMIJO, max_qp_iterations=20, ModPi retraction",
                       "links": links, "batch_per_gpu": B},
            "nls_problems_per_sec": B * ctx.world / (ms * 1e-3), "absolute_tol_fraction": ok / (B * ctx.world),
            "roofline": roofline, "cpu_baseline": cpu, "parity": parity, "e2e": e2e,
            "gpu_launches": int(launches), "clocks": clocks.summary()}


ALSO_KEYS = ("value", "unit", "ms_per_step", "config", "roofline", "e2e", "parity", "cpu_baseline", "converged_fraction",
             "solved_qps_per_sec", "nls_problems_per_sec", "absolute_tol_fraction", "gpu_launches")


def short(rec):
    return {k: rec[k] for k in ALSO_KEYS if k in rec}


# ---------------------------------------------------------------------------------------------------------
# reference arm: the reference's CPU implementation, all host threads, bounded sample per step
# ---------------------------------------------------------------------------------------------------------
def reference_qp(n, k, sample, steps, warmup, dev):
    threads = host_cores()
    prm = synth.solve_params()
    run_reference_qp(synth.to_numpy(gen_qp_batch(n, k, min(sample, 4096), 7, dev)), prm, threads)  # page-in / thread start
    # the same stream as the GPU arm's rank 0 (seed 1234, generated on the device when there is one)
    host = synth.to_numpy(gen_qp_batch(n, k, sample, 1234, dev))
    times, steps_per, res = [], 0, None
    for i in range(warmup + steps):
        res = run_reference_qp(host, prm, threads)
        if i >= warmup:
            times.append(res["seconds"])
            steps_per += res["steps"]
    total = sum(times)
    value = steps_per / total
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "steps": steps, "warmup": warmup,
        "ms_per_step": 1e3 * total / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": qp_workload_name(n, k, sample), "n": n, "k": k, "m": 2 * n, "cpu_sample_per_step": sample,
                   "timing": "std::chrono around the threaded solve loop (unpacking into mini_opt::QP outside)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": res["cores"], "kind": res["kind"],
                         "sample": f"{sample} QPs n={n} k={k} per step: the first problems of the GPU arm's rank-0 batch, "
                                   "one QPInteriorPointSolver per std::thread, g++ -O2 -march=x86-64-v3"},
        "converged_fraction": res["solved"] / sample,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def reference_ik(links, sample):
    from oracle import oracle_py as O
    prm = synth.chain_ik_params()
    ch = synth.chain_ik_batch_numpy(links, sample, seed=4321)
    threads = host_cores()
    O.nls_solve_chain("ref", synth.chain_ik_batch_numpy(links, 512, seed=1), prm, nthreads=threads)
    t0 = time.perf_counter()
    _, total = O.nls_solve_chain("ref", ch, prm, nthreads=threads)
    dt = time.perf_counter() - t0
    return {"impl": "reference", "metric": METRIC, "value": total / dt, "unit": UNIT, "steps": 1, "warmup": 0,
            "ms_per_step": 1e3 * dt, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"config3: {sample} chain-IK problems, {links} links (CPU sample)"},
            "nls_problems_per_sec": sample / dt,
            "cpu_baseline": {"value": total / dt, "unit": UNIT, "cores": threads, "kind": "reference",
                             "sample": f"{sample} problems, one ConstrainedNonlinearLeastSquares per std::thread"},
            "e2e": {"value": total / dt, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    dev = None
    try:
        import torch
        if torch.cuda.is_available():
            dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    except Exception:
        pass
    if args.workload == "ik":
        line = reference_ik(args.links, min(args.batch if args.batch != (1 << 20) else 100000, 20000))
    else:
        line = reference_qp(args.n, args.k, args.cpu_sample, args.steps, args.warmup, dev)
        if args.n == 8 and args.k == 2 and not args.no_also:
            line["also"] = {"n64": short(reference_qp(64, 16, 8192, 1, 0, dev)), "ik6": short(reference_ik(6, 20000))}
    line["n_gpus"] = args.gpus
    print(json.dumps(line), flush=True)


SWEEP = {8: (2, [1 << 10, 1 << 14, 1 << 18, 1 << 20, 1 << 24]), 32: (8, [1 << 10, 1 << 14, 1 << 18, 1 << 20]),
         64: (16, [1 << 10, 1 << 14, 1 << 18, 1 << 20])}


def sweep(args, ctx: Ctx):
    """BASELINE config 5: batch-size sweep x n in {8, 32, 64} at this world size; one JSON line per point. The CPU
    baseline (N=1) is measured once per n on a capped sample of the largest batch and repeated on every line of that n."""
    for n, (k, batches) in SWEEP.items():
        cpu = None
        for B in batches:
            if args.sweep_max_batch and B > args.sweep_max_batch:
                continue
            steps = 20 if B <= (1 << 18) else (5 if B <= (1 << 20) else 3)
            sample = {8: 131072, 32: 16384, 64: 4096}[n] if (cpu is None and B >= 16384) else 0
            rec = qp_pass(ctx, n, k, B, steps, 3, with_e2e=(B <= (1 << 18) and not args.no_e2e),
                          cpu_sample=0 if args.no_cpu else sample, seed=1234)
            if rec["cpu_baseline"]:
                cpu = rec["cpu_baseline"]
            elif cpu:
                rec["cpu_baseline"] = cpu
            rec["sweep"] = True
            ctx.torch.cuda.empty_cache()
            if ctx.rank == 0:
                print(json.dumps(rec), flush=True)


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
    ap.add_argument("--no-also", action="store_true", help="skip the config-4 (n=64) and config-3 (IK) records")
    ap.add_argument("--sweep", action="store_true", help="BASELINE config 5: one JSON line per (n, batch)")
    ap.add_argument("--sweep-max-batch", type=int, default=0)
    ap.add_argument("--e2e-format", default="compact", choices=["compact", "full"],
                    help="host buffers of the e2e leg: compact = packed lower-triangle G, one box per batch, x-only output")
    args = ap.parse_args()
    global E2E_FORMAT
    E2E_FORMAT = args.e2e_format
    if args.impl == "reference":
        reference_arm(args)
        return
    ctx = Ctx()
    ctx.init_gpu()
    if args.sweep:
        sweep(args, ctx)
        ctx.finish()
        return
    default_line = args.workload == "qp" and args.n == 8 and args.k == 2 and args.batch == (1 << 20)
    if args.workload == "ik":
        B = args.batch if args.batch != (1 << 20) else 100000
        line = ik_pass(ctx, args.links, B, args.steps, args.warmup, not args.no_e2e, 0 if args.no_cpu else 20000)
    else:
        line = qp_pass(ctx, args.n, args.k, args.batch, args.steps, args.warmup, not args.no_e2e,
                       0 if args.no_cpu else args.cpu_sample)
    if default_line and not args.no_also:
        # BASELINE.json's metric is quoted on n=8 AND n=64; config 3 rides along (same process, same GPU(s))
        ctx.torch.cuda.empty_cache()
        n64 = qp_pass(ctx, 64, 16, 65536, max(3, min(args.steps, 10)), 3, not args.no_e2e, 0 if args.no_cpu else 8192)
        ctx.torch.cuda.empty_cache()
        ik6 = ik_pass(ctx, 6, 100000, max(2, min(args.steps, 5)), 3, not args.no_e2e, 0 if args.no_cpu else 20000)
        line["also"] = {"n64": short(n64), "ik6": short(ik6)}
        line["gpu_launches"] = line["gpu_launches"]  # the main line's own launches; `also` carries its own counts
    if ctx.rank == 0:
        print(json.dumps(line), flush=True)
    ctx.finish()


if __name__ == "__main__":
    main()
