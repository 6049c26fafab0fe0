#!/usr/bin/env python
"""Aggregate the ncu source page (--page source --csv) of a .ncu-rep: top stalled SASS instructions and
stall totals per opcode. Usage: ncu_stalls.py report.ncu-rep [topN]"""
import csv, subprocess, sys, collections
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
recs = []
for n, r in enumerate(rows[2:]):
    if len(r) < len(hdr): continue
    recs.append((n, r))
tot = sum(int(r[ix["# Samples"]] or 0) for _, r in recs)
print("instructions:", len(recs), "total samples:", tot)
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("stall totals:", {h: sum(int(r[ix[h]] or 0) for _, r in recs) for h in stall_cols if sum(int(r[ix[h]] or 0) for _, r in recs)})
byop = collections.Counter(); exe = collections.Counter()
for n, r in recs:
    op = r[ix["Source"]].split()[0] if not r[ix["Source"]].strip().startswith("@") else r[ix["Source"]].split()[1]
    op = op.split(".")[0]
    byop[op] += int(r[ix["# Samples"]] or 0); exe[op] += int(r[ix["Instructions Executed"]] or 0)
print("samples by opcode:", byop.most_common(14))
print("executed by opcode:", exe.most_common(14))
recs.sort(key=lambda t: -int(t[1][ix["# Samples"]] or 0))
for n, r in recs[:top]:
    st = {h[6:]: int(r[ix[h]] or 0) for h in stall_cols if int(r[ix[h]] or 0) > 0}
    print(f"{n:5d} {r[ix['# Samples']]:>6} exec={r[ix['Instructions Executed']]:>9} thr={r[ix['Avg. Threads Executed']]:>4} {r[ix['Source']].strip()[:70]:70s} {st}")
