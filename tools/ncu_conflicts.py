#!/usr/bin/env python
"""Top shared-memory bank-conflict instructions of a .ncu-rep source page. Usage: ncu_conflicts.py report.ncu-rep [topN]"""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 20
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines())); hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
recs = [r for r in rows[2:] if len(r) >= len(hdr)]
f = lambda r, h: int(float(r[ix[h]] or 0))
tot_ex = sum(f(r, "L1 Wavefronts Shared Excessive") for r in recs); tot = sum(f(r, "L1 Wavefronts Shared") for r in recs)
print("shared wavefronts:", tot, "excessive:", tot_ex)
by = {}
for r in recs:
    op = r[ix["Source"]].split()[0] if not r[ix["Source"]].strip().startswith("@") else r[ix["Source"]].split()[1]
    d = by.setdefault(op, [0, 0, 0]); d[0] += f(r, "L1 Wavefronts Shared"); d[1] += f(r, "L1 Wavefronts Shared Excessive"); d[2] += f(r, "Instructions Executed")
for op, d in sorted(by.items(), key=lambda kv: -kv[1][1])[:10]: print(op, d)
recs.sort(key=lambda r: -f(r, "L1 Wavefronts Shared Excessive"))
for r in recs[:top]:
    print(r[ix["Source"]].strip()[:60].ljust(60), "exec", r[ix["Instructions Executed"]], "thr", r[ix["Avg. Threads Executed"]], "wave", r[ix["L1 Wavefronts Shared"]], "ideal", r[ix["L1 Wavefronts Shared Ideal"]], "excess", r[ix["L1 Wavefronts Shared Excessive"]])
