#!/usr/bin/env python
"""Per-CUDA-source-line shares of executed warp instructions and stall samples.
    ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > /tmp/cs.csv
    python tools/ncu_lines.py /tmp/cs.csv [min_pct]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
minp = float(sys.argv[2]) if len(sys.argv) > 2 else 0.4
out, fname, h = [], None, None
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": h = r; ie = h.index("Instructions Executed"); sa = h.index("# Samples"); at = h.index("Avg. Threads Executed"); continue
    if h is None or r[0] in ("", "Function Name"): continue
    try: out.append((fname, int(r[0]), r[1].strip(), float(r[ie] or 0), float(r[sa] or 0), r[at]))
    except ValueError: pass
ti = sum(o[3] for o in out); ts = sum(o[4] for o in out)
print(f"total warp-inst {ti:.4g}, samples {ts:.0f}")
for f, ln, src, i, s, a in out:
    if 100 * i / ti >= minp or 100 * s / ts >= minp:
        print(f"{f}:{ln:<4} inst {100*i/ti:5.2f}% samples {100*s/ts:5.2f}% lanes {a:>4}  {src[:100]}")
