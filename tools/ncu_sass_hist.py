#!/usr/bin/env python
"""Histogram of stall samples / executed instructions over 64-row windows of the SASS view
(`ncu -i X.ncu-rep --page source --csv`), with each window's dominant opcodes."""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
W = int(sys.argv[2]) if len(sys.argv) > 2 else 64
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[hi + 1:] if len(r) >= len(hdr) - 2]
S, N, SMP = col["Source"], col["Instructions Executed"], col["# Samples"]
f = lambda x: float(x) if x else 0.0
ts = sum(f(r[SMP]) for r in body)
ti = sum(f(r[N]) for r in body)
for lo in range(0, len(body), W):
    seg = body[lo:lo + W]
    s = sum(f(r[SMP]) for r in seg)
    n = sum(f(r[N]) for r in seg)
    if s / ts < 0.004:
        continue
    ops = collections.Counter()
    for r in seg:
        m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", r[S])
        ops[m.group(1) if m else "?"] += f(r[N])
    print(f"{lo:5d} samples {100 * s / ts:5.2f}% inst {100 * n / ti:5.2f}%  " + ", ".join(k for k, _ in ops.most_common(5)))
