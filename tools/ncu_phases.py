#!/usr/bin/env python
"""Split the SASS of a kernel in a .ncu-rep at its CTA barriers and print warp-stall samples per segment
(barrier-delimited phase): total, barrier-wait share, opcode signature. Usage: ncu_phases.py report.ncu-rep"""
import csv, subprocess, sys, collections
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines())); hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
seg = []; cur = dict(start=0, n=0, samp=0, bar=0, noinst=0, ops=collections.Counter(), exe=0)
for n, r in enumerate(rows[2:]):
    if len(r) < len(hdr): continue
    src = r[ix["Source"]].split(); op = (src[1] if src[0].startswith("@") else src[0]).split(".")[0]
    s = int(r[ix["# Samples"]] or 0)
    cur["n"] += 1; cur["samp"] += s; cur["bar"] += int(r[ix["stall_barrier"]] or 0); cur["noinst"] += int(r[ix["stall_no_inst"]] or 0)
    cur["exe"] += int(r[ix["Instructions Executed"]] or 0)
    if op in ("DMMA", "SHFL", "MUFU", "LDS", "STS", "DFMA", "LDGSTS", "BRX", "LDG", "STG"): cur["ops"][op] += 1
    if op == "BAR":
        seg.append(cur); cur = dict(start=n + 1, n=0, samp=0, bar=0, noinst=0, ops=collections.Counter(), exe=0)
seg.append(cur)
tot = sum(c["samp"] for c in seg)
print(f"total samples {tot}")
for c in seg:
    print(f"sass[{c['start']:5d}+{c['n']:4d}] samples {c['samp']:6d} ({100*c['samp']/tot:5.1f}%)  non-barrier {c['samp']-c['bar']:6d} ({100*(c['samp']-c['bar'])/tot:5.1f}%) no_inst {c['noinst']:5d} exec {c['exe']/1e6:7.1f}M  {dict(c['ops'])}")
