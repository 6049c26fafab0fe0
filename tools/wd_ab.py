"""A/B of the thread-per-problem kernel's refill: cp.async one round ahead (mo_debug_set_wd_lock(1), default) vs LDG/STS
through registers (0, the round-1 schedule). Same launch otherwise; every output array must be bit-identical.
Usage (GPU box): python tools/tpp_ab.py [--n 8 --k 2 --batch 1048576]"""
import argparse, json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from mini_opt_b200 import lib, qp, synth

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=64)
ap.add_argument("--k", type=int, default=16)
ap.add_argument("--batch", type=int, default=65536)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--guess", default="eq")
args = ap.parse_args()
L = lib.load()
dev = torch.device("cuda:0")
n, k, B = args.n, args.k, args.batch
arrs = synth.box_qp_batch_torch(n, k, B, seed=1234, device=dev)
prm = synth.solve_params()
res = {"kernel": lib.kernel_variant(n, k, 2 * n, arrs.flags), "n": n, "k": k, "batch": B}
ref = None
for mode in (1, 0, 1, 0):
    L.mo_debug_set_wd_lock(mode)
    out = qp.alloc_out_torch(B, n, k, 2 * n, dev, lagrange=True)
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
    cur = {f: getattr(out, f).clone() for f in ("vars", "status", "iters", "kkt_final", "lagrange") if getattr(out, f, None) is not None}
    if ref is None:
        ref = cur
    same = all(torch.equal(cur[f].view(torch.int64) if cur[f].dtype == torch.float64 else cur[f],
                           ref[f].view(torch.int64) if ref[f].dtype == torch.float64 else ref[f]) for f in ref)
    steps = int(out.iters.sum().item())
    res.setdefault("lock" if mode else "single", []).append({"ms": ms, "steps_per_s": steps / (ms * 1e-3), "bit_identical_to_first": bool(same)})
L.mo_debug_set_wd_lock(1)
res["converged"] = float((out.status == 0).float().mean().item())
print(json.dumps(res))
