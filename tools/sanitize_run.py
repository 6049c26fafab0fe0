"""Small solves of every kernel family in one process (a quick end-to-end sanity run; also the input for a sanitizer where one is allowed): tpp (both refill schedules), warp16,
dmma_warp, generic, chain IK, a functor family. Usage (GPU box): compute-sanitizer --tool memcheck python tools/sanitize_run.py"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
from mini_opt_b200 import abi, lib, nls, qp, synth

L = lib.load()
dev = torch.device("cuda:0")
prm = synth.solve_params()
for n, k, B in ((8, 2, 3000), (4, 1, 1500), (6, 2, 1500), (12, 3, 500), (16, 8, 256), (64, 16, 21), (20, 4, 64)):
    arrs = qp.to_device(synth.box_qp_batch_numpy(n, k, B, seed=7, explicit_constraints=False), dev)
    for mode in ((1, 0) if lib.kernel_variant(n, k, 2 * n, arrs.flags) == "tpp" else (1,)):
        L.mo_debug_set_tpp_async(mode)
        out = qp.alloc_out_torch(B, n, k, 2 * n, dev)
        qp.solve_batch_device(arrs, prm, out, dev)
        torch.cuda.synchronize()
        print(lib.kernel_variant(n, k, 2 * n, arrs.flags), n, k, B, "converged", float((out.status == 0).float().mean()))
L.mo_debug_set_tpp_async(1)
ch = nls.chain_to_device(synth.chain_ik_batch_numpy(6, 600, seed=4321), dev)
o = nls.alloc_nls_out_torch(600, ch.n, dev)
nls.solve_chain_ik_device(ch, synth.chain_ik_params(), o, dev)
torch.cuda.synchronize()
print("chain ik", int((o.termination > 0).sum()))
x0 = torch.tensor([[-5.0, -3.0], [10.0, 8.0], [0.8, -0.3]] * 40, dtype=torch.float64, device=dev)
o = nls.alloc_nls_out_torch(x0.shape[0], 2, dev)
nls.solve_family_device(abi.MO_NLS_FAMILY_ROSENBROCK, x0, abi.NLSParams(max_iterations=5, max_qp_iterations=1), o, dev)
torch.cuda.synchronize()
print("rosenbrock family", o.x[:3].cpu().numpy().round(6).tolist())
