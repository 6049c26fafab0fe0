import sys, torch
sys.path.insert(0, "/root/repo")
from mini_opt_b200 import synth, qp
dev = torch.device("cuda:0")
for (n, k, B) in [(64, 16, 24), (32, 8, 24), (16, 8, 24)]:
    arrs = synth.box_qp_batch_torch(n, k, B, seed=3, device=dev)
    out = qp.alloc_out_torch(B, n, k, 2 * n, dev)
    qp.solve_batch_device(arrs, synth.solve_params(), out, dev); torch.cuda.synchronize()
    print(n, k, "status", out.status.cpu().tolist()[:8], "iters", out.iters.cpu().tolist()[:8])
