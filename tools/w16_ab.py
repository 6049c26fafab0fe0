"""A/B of the warp16 kernel's staging: TMA bulk copies (cp.async.bulk + mbarrier) vs per-lane loads, same launch otherwise.
Usage (GPU box): python tools/w16_ab.py [--n 12 --k 3 --batch 262144]"""
import argparse, json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
from mini_opt_b200 import abi, lib, qp, synth

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=12)
ap.add_argument("--k", type=int, default=3)
ap.add_argument("--batch", type=int, default=262144)
ap.add_argument("--reps", type=int, default=10)
args = ap.parse_args()
L = lib.load()
dev = torch.device("cuda:0")
n, k, B = args.n, args.k, args.batch
arrs = synth.box_qp_batch_torch(n, k, B, seed=77, device=dev)
prm = synth.solve_params()
out = qp.alloc_out_torch(B, n, k, 2 * n, dev, lagrange=False)
res = {"kernel": lib.kernel_variant(n, k, 2 * n, arrs.flags), "n": n, "k": k, "batch": B}
ref_iters = None
for mode in (1, 0, 1, 0):
    L.mo_debug_set_tma(mode)
    for _ in range(3):
        qp.solve_batch_device(arrs, prm, out, dev)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.reps):
        qp.solve_batch_device(arrs, prm, out, dev)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.reps
    it = out.iters.cpu().numpy().copy()
    if ref_iters is None:
        ref_iters = it
    assert np.array_equal(it, ref_iters), "staging mode changed the results"
    steps = int(it.sum())
    res.setdefault("tma" if mode else "lanes", []).append({"ms": ms, "steps_per_s": steps / (ms * 1e-3)})
L.mo_debug_set_tma(1)
F = (n + k) ** 3 / 3 + 2 * (n + k) ** 2 + 2 * n * n + 4 * n * k + 30 * 2 * n
res["flops_per_step"] = F
res["tflops_tma"] = F * res["tma"][-1]["steps_per_s"] / 1e12
res["converged"] = float((out.status == 0).float().mean().item())
print(json.dumps(res))
