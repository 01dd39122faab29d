#!/usr/bin/env python
"""Split a kernel's SASS (ncu --page source --csv) into runs of equal execution count and print
each run's share of executed instructions / stall samples — shows loop bodies, rare paths and
straight-line phases at a glance.   python tools/ncu_regions.py X.ncu-rep [min_pct]"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
col = {h: i for i, h in enumerate(rows[hi])}
body = rows[hi + 1:]
f = lambda x: float(x) if x else 0.0
I, S, TI, SRC = col["Instructions Executed"], col["# Samples"], col["Thread Instructions Executed"], col["Source"]
ti = sum(f(r[I]) for r in body); ts = sum(f(r[S]) for r in body)
minp = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3
segs = []; prev = None
for i, r in enumerate(body):
    c = f(r[I])
    if prev is None or abs(c - prev) > 0.02 * max(c, prev, 1):
        segs.append([i, i, c, 0.0, 0.0, 0.0, 0])
    g = segs[-1]; g[1] = i; g[3] += c; g[4] += f(r[S]); g[5] += f(r[TI]); g[6] += ("FFMA2" in r[SRC]); prev = c
print(f"total warp-inst {ti:.4g}, samples {ts:.0f}")
for a, b, c, si, ss, st, nf in segs:
    if 100 * si / ti >= minp:
        print(f"rows {a:5d}-{b:5d} ({b-a+1:4d} instr, {nf:3d} FFMA2) exec {c:12.0f} inst {100*si/ti:5.2f}% samples {100*ss/ts:5.2f}% lanes {st/max(si,1):4.1f}  {body[a][SRC].strip()[:44]}")
