#!/usr/bin/env python
"""Phase timeline of qp_dm_kernel (debug build: make -C mini_opt_b200/csrc OUT=../libminiopt_trace.so EXTRA=-DMO_DM_TRACE).
Solves a full batch with the trace library and prints per-phase cycle totals of CTA 0's first Newton steps."""
import ctypes as C, sys, collections
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from mini_opt_b200 import abi, lib, synth, qp
lib.LIB_PATH = ROOT / "mini_opt_b200" / "libminiopt_trace.so"
L = lib.load()
n, k, B = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
dev = torch.device("cuda:0")
arrs = synth.box_qp_batch_torch(n, k, B, seed=1, device=dev)
out = qp.alloc_out_torch(B, n, k, 2 * n, dev)
qp.solve_batch_device(arrs, synth.solve_params(), out, dev); torch.cuda.synchronize()
buf = (C.c_longlong * 2048)()  # first 1024 events of CTA 0
L.mo_debug_read_trace.argtypes = [C.c_void_p, C.c_int]
L.mo_debug_read_trace(buf, 2048)  # discard the first (cold) run
qp.solve_batch_device(arrs, synth.solve_params(), out, dev); torch.cuda.synchronize()
cnt = L.mo_debug_read_trace(buf, 2048)
ev = [(buf[i] >> 8, buf[i] & 255) for i in range(cnt)]
names = {(4, 8): "P4 pass A (warp0)", (8, 9): "look-ahead diag block (warp0 owner)", (9, 1): "P4 pass B after diag (warp0 owner)",
         (8, 1): "P4 pass B (warp0 not owner)", (1, 2): "P4 barrier wait", (2, 3): "P3 trsm (warp0)", (3, 4): "P3 barrier wait",
         (2, 8): "prologue pass A", (15, 8): "assemble->prologue spill", (4, 6): "last panel exit", (6, 16): "backsubst: inv(L_JJ) columns",
         (16, 7): "backsubst: block loop", (10, 11): "stage wait", (11, 12): "residual matvec+per-var", (12, 13): "reduce",
         (13, 14): "bookkeeping", (14, 15): "assemble", (7, 10): "update+loop", (11, 14): "guess rhs"}
tot = collections.Counter(); num = collections.Counter()
for (t0, a), (t1, b) in zip(ev, ev[1:]):
    tot[(a, b)] += t1 - t0; num[(a, b)] += 1
steps = num[(6, 7)]
print(f"events {cnt}, factor-solves traced {steps}, total cycles {ev[-1][0] - ev[0][0]}")
for key, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print(f"{names.get(key, str(key)):28s} total {v:9d}  per factor-solve {v / max(steps, 1):9.0f}  n={num[key]}  avg {v / num[key]:7.0f}")
