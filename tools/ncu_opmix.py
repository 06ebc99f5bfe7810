#!/usr/bin/env python
"""Opcode mix of the executed warp-instructions of the first kernel in an .ncu-rep (source page).
usage: python tools/ncu_opmix.py prof.ncu-rep [top]"""
import csv, io, collections, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ie, it, src = h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("Source")
tot, thr, n = collections.Counter(), collections.Counter(), collections.Counter()
for r in rows[hi + 1:]:
    if len(r) <= it or not r[ie].isdigit():
        continue
    t = r[src].split()
    op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    tot[op] += int(r[ie]); thr[op] += int(r[it]); n[op] += 1
T = sum(tot.values())
print(f"total warp-inst {T:,}")
for k, v in tot.most_common(top):
    print(f"{k:12s} {v / T * 100:6.2f}%  act {thr[k] / max(v, 1):5.1f}  ({n[k]} sass)")
