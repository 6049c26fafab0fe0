"""Bench of mo_nls_linearize_batch on dense per-problem Jacobians (north_star item 4): n = 64, R = 128 runs the DMMA kernel
(nls_linearize_dmma.cuh), n = 12 the warp DFMA kernel. Reports time, algorithmic TFLOP/s (R N (N + 1) + 2 R N + N per problem,
SURVEY 8(d)) against the measured DMMA / DFMA peak and algorithmic GB/s against MEASURED_PEAKS.json's HBM copy bandwidth.
Usage (GPU box): python tools/lin_bench.py [--n 64 --r 128 --batch 65536]"""
import argparse, ctypes as C, json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch
from mini_opt_b200 import abi, lib

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=64)
ap.add_argument("--r", type=int, default=128)
ap.add_argument("--k", type=int, default=16)
ap.add_argument("--batch", type=int, default=65536)
ap.add_argument("--reps", type=int, default=10)
a = ap.parse_args()
L = lib.load()
dev = torch.device("cuda:0")
n, R, k, B, m = a.n, a.r, a.k, a.batch, 2 * a.n
g = torch.Generator(device=dev); g.manual_seed(3)
f64 = dict(dtype=torch.float64, device=dev)
rnd = lambda *s: torch.rand(*s, generator=g, **f64) * 2 - 1
inp = abi.LinearizeArrays(n=n, r_cost=R, cost_block=R, r_eq=k, m=m, batch=B, x=rnd(B, n), J_cost=rnd(B, n, R), r_cost_v=rnd(B, R),
                          J_eq=rnd(B, n, k), r_eq_v=rnd(B, k), lam=torch.full((B,), 1e-3, **f64),
                          ineq_var=torch.arange(n, dtype=torch.int32, device=dev).repeat_interleave(2).repeat(B, 1).contiguous(),
                          ineq_a=torch.tensor([1.0, -1.0], **f64).repeat(B, n).contiguous(), ineq_b=torch.ones(B, m, **f64))
out = abi.LinearizeOut(G=torch.empty(B, n, n, **f64), c=torch.empty(B, n, **f64), A_eq=torch.empty(B, n, k, **f64),
                       b_eq=torch.empty(B, k, **f64), ineq_b_shifted=torch.empty(B, m, **f64), errors=torch.empty(B, 2, **f64))
args = abi.linearize_to_c(inp, out)
st = torch.cuda.current_stream(dev)
call = lambda: lib.check(L.mo_nls_linearize_batch(C.byref(args), 0, st.cuda_stream))
for _ in range(3): call()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.reps): call()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.reps
# spot check against torch on the first problems (lower triangle of J'J + lambda I)
J = inp.J_cost[:64].transpose(1, 2)  # [b, R, n]
want = torch.tril(J.transpose(1, 2) @ J + 1e-3 * torch.eye(n, **f64))
got = out.G[:64].transpose(1, 2)  # column-major -> [row, col]
err = float((got - want).abs().max())
flops = B * (R * n * (n + 1) + 2 * R * n + n)
bytes_ = 8 * B * (R * n + R + n * n + n + k * n * 2 + k * 2 + m * 2 + n + 2)
peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
dmma, dfma = lib.measure_dmma_peak(0), lib.measure_fp64_peak(0)
use_dmma = n % 8 == 0 and 32 <= n <= 64 and R % 2 == 0
peak = dmma if use_dmma else dfma
print(json.dumps({"kernel": "nls_linearize_dmma_kernel" if use_dmma else "nls_linearize_kernel", "n": n, "r_cost": R, "k": k, "batch": B,
                  "ms": ms, "problems_per_s": B / (ms * 1e-3), "max_abs_err_vs_torch": err,
                  "roofline": {"bound": "tensor+hbm", "achieved_tflops": flops / (ms * 1e-3) / 1e12, "peak_tflops": peak,
                               "frac_fp64": flops / (ms * 1e-3) / 1e12 / peak, "achieved_gbs": bytes_ / (ms * 1e-3) / 1e9,
                               "hbm_peak_gbs": peaks.get("hbm_gbs"), "frac_hbm": bytes_ / (ms * 1e-3) / 1e9 / peaks.get("hbm_gbs", 6553.6),
                               "algorithmic_bytes": bytes_, "algorithmic_flops": flops}}))
