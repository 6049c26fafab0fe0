"""One process, several GPUs: mo_qp_ip_solve_batch_host_multi on pinned host buffers (compact formats) vs the same batch
on one device. Reports e2e Newton steps/s and the time of the gather (the D2H copies are the gather: each shard's results
land in its slice of the caller's arrays). Usage (GPU box): python tools/e2e_multi.py --devices 0,1[,2,...]"""
import argparse, json, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
from mini_opt_b200 import abi, lib, qp, synth
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--devices", default="0,1")
ap.add_argument("--n", type=int, default=8)
ap.add_argument("--k", type=int, default=2)
ap.add_argument("--per-gpu", type=int, default=1 << 20)
ap.add_argument("--interleave", type=int, default=1, help="pinned buffers interleaved over the NUMA nodes (multi-device runs)")
a = ap.parse_args()
devs = [int(d) for d in a.devices.split(",")]
L = lib.load()
n, k, m = a.n, a.k, 2 * a.n
res = {"n": n, "k": k, "per_gpu": a.per_gpu, "interleave": a.interleave, "runs": []}
try:
    res["numa_nodes_online"] = open("/sys/devices/system/node/online").read().strip()
except OSError:
    pass
for nd in sorted({1, len(devs)}):
    B = a.per_gpu * nd
    host = synth.box_qp_batch_numpy(n, k, B, seed=1234, explicit_constraints=False)
    keep = []
    def pin(x):
        v, p = bench.pinned_copy(L, x, interleaved=bool(a.interleave) and nd > 1); keep.append(p); return v
    harrs = abi.QPBatchArrays(n=n, k=k, m=m, batch=B, G=pin(qp.pack_lower(host.G)), c=pin(host.c), A_eq=pin(host.A_eq), b_eq=pin(host.b_eq),
                              ineq_b=pin(np.ascontiguousarray(host.ineq_b[0])),
                              flags=host.flags | abi.MO_QP_FLAG_PACKED_G | abi.MO_QP_FLAG_UNIFORM_BOX)
    hout = abi.QPOutArrays(vars=None, status=pin(np.zeros(B, dtype=np.int32)), iters=pin(np.zeros(B, dtype=np.int32)),
                           kkt_final=pin(np.zeros((B, 4))), x=pin(np.zeros((B, n))))
    prm = synth.solve_params()
    for _ in range(2):
        qp.solve_batch_host_multi(harrs, prm, hout, devs[:nd])
    t0 = time.perf_counter()
    reps = 4
    for _ in range(reps):
        qp.solve_batch_host_multi(harrs, prm, hout, devs[:nd])
    sec = (time.perf_counter() - t0) / reps
    steps = int(hout.iters.sum())
    res["runs"].append({"devices": nd, "batch": B, "ms": 1e3 * sec, "steps_per_s": steps / sec, "qps_per_s": B / sec,
                        "h2d_bytes": int(sum(x.nbytes for x in (harrs.G, harrs.c, harrs.A_eq, harrs.b_eq, harrs.ineq_b))),
                        "d2h_bytes": int(sum(x.nbytes for x in (hout.x, hout.status, hout.iters, hout.kkt_final)))})
    for p in keep:
        L.mo_host_free(p)
r = res["runs"]
if len(r) == 2:
    res["e2e_scaling_efficiency"] = r[1]["steps_per_s"] / (r[0]["steps_per_s"] * r[1]["devices"])
print(json.dumps(res))
