"""A/B of the lock-step chain-IK kernel: warps per CTA 1 (round-1 schedule: every warp on its own), 16, 24 through
mo_debug_set_ik_warps. Results must be bit-identical; only the time may change.
Usage (GPU box): python tools/ik_ab.py [--links 6 --batch 100000]"""
import argparse, json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from mini_opt_b200 import lib, nls, synth

ap = argparse.ArgumentParser()
ap.add_argument("--links", type=int, default=6)
ap.add_argument("--batch", type=int, default=100000)
ap.add_argument("--n-eq", type=int, default=0)
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
L = lib.load()
dev = torch.device("cuda:0")
ch = nls.chain_to_device(synth.chain_ik_batch_numpy(a.links, a.batch, seed=4321, n_eq=a.n_eq), dev)
prm = synth.chain_ik_params()
res = {"links": a.links, "batch": a.batch, "n_eq": a.n_eq, "runs": []}
ref = None
for w, mask in ((1, 3), (24, 1), (24, 5), (24, 13), (24, 4), (24, 1)):
    assert L.mo_debug_set_ik_warps(w) >= 1
    L.mo_debug_set_ik_sync(mask)
    out = nls.alloc_nls_out_torch(a.batch, ch.n, dev)
    nls.solve_chain_ik_device(ch, prm, out, dev)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.reps):
        nls.solve_chain_ik_device(ch, prm, out, dev)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.reps
    cur = nls.nls_out_to_numpy(out)
    if ref is None:
        ref = cur
    same = all((getattr(cur, f) == getattr(ref, f)).all() for f in ("x", "termination", "iterations", "qp_iterations", "ls_steps"))
    res["runs"].append({"warps_per_cta": w, "sync_mask": mask, "ms": ms, "problems_per_s": a.batch / (ms * 1e-3),
                        "qp_steps_per_s": float(cur.qp_iterations.sum()) / (ms * 1e-3), "bit_identical_to_first": bool(same)})
L.mo_debug_set_ik_warps(24)
L.mo_debug_set_ik_sync(1)
print(json.dumps(res))
