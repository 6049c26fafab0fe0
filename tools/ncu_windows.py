#!/usr/bin/env python
"""Warp-stall samples of a kernel in a .ncu-rep per window of W SASS instructions, with an opcode signature
(for kernels without CTA barriers). Usage: ncu_windows.py report.ncu-rep [W]"""
import csv, subprocess, sys, collections
W = int(sys.argv[2]) if len(sys.argv) > 2 else 250
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines())); hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
recs = [r for r in rows[2:] if len(r) >= len(hdr)]
tot = sum(int(r[ix["# Samples"]] or 0) for r in recs)
print("total samples", tot, "instructions", len(recs))
for s in range(0, len(recs), W):
    ops = collections.Counter(); samp = 0; exe = 0
    for r in recs[s:s + W]:
        src = r[ix["Source"]].split(); op = (src[1] if src[0].startswith("@") else src[0]).split(".")[0]
        samp += int(r[ix["# Samples"]] or 0); exe += int(r[ix["Instructions Executed"]] or 0)
        if op in ("DMMA", "SHFL", "MUFU", "LDS", "STS", "DFMA", "LDGSTS", "LDG", "STG", "WARPSYNC", "CALL", "NOP"): ops[op] += 1
    print(f"sass[{s:5d}] samples {samp:6d} ({100*samp/tot:5.1f}%) exec {exe/1e6:7.1f}M {dict(ops)}")
