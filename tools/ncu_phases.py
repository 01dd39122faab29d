#!/usr/bin/env python
"""Executed warp instructions, stall samples and active lanes of a kernel per PHASE of the source.
Phases are the `// @phase <name>` markers in csrc/rt_wave.cu and csrc/rt_device.cuh: a source line
belongs to the last marker above it; a SASS instruction is counted once, under the phase of the
innermost user source line it is attributed to.

    ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > /tmp/cs.csv
    python tools/ncu_phases.py /tmp/cs.csv [csrc dir]"""
import csv
import os
import re
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
csrc = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "ray-tracing-in-one-weekend_b200", "csrc")


def phase_map(path):
    out, cur = {}, "other"
    for n, line in enumerate(open(path), 1):
        m = re.search(r"//\s*@phase\s+(\S+)", line)
        if m:
            cur = m.group(1)
        out[n] = cur
    return out


maps = {f: phase_map(os.path.join(csrc, f)) for f in ("rt_wave.cu", "rt_device.cuh") if os.path.exists(os.path.join(csrc, f))}
rows = list(csv.reader(open(sys.argv[1])))
fname, h, cur = None, None, None
sass, locs = {}, defaultdict(list)
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        h = r
        ie, sa, ti = h.index("Instructions Executed"), h.index("# Samples"), h.index("Thread Instructions Executed")
        continue
    if r[0] == "Function Name" or h is None:
        continue
    if r[0] != "":
        cur = (fname, int(r[0]))
        continue
    try:
        sass[r[2]] = (r[3].strip(), float(r[ie] or 0), float(r[ti] or 0), float(r[sa] or 0))
        locs[r[2]].append(cur)
    except (ValueError, IndexError):
        pass
G = defaultdict(lambda: [0.0, 0.0, 0.0, 0])
for a, (txt, i, t, sm) in sass.items():
    ph = [maps[f].get(l, "other") for f, l in locs[a] if f in maps]
    ph = [p for p in ph if p != "other"] or ph or ["library"]
    # helper functions (wffma2, philox, ...) sit under their own marker; prefer the call site's phase
    g = G[ph[-1] if ph[-1] != "helpers" else ph[0]]
    g[0] += i; g[1] += t; g[2] += sm; g[3] += 1
TI = sum(g[0] for g in G.values()); TS = sum(g[2] for g in G.values())
print(f"total warp instructions {TI:.4g}, stall samples {TS:.0f}, SASS instructions {len(sass)}, "
      f"lanes per instruction {sum(g[1] for g in G.values()) / TI:.2f}")
for k, g in sorted(G.items(), key=lambda kv: -kv[1][0]):
    print(f"{k:16s} sass {g[3]:5d}  inst {100 * g[0] / TI:6.2f}%  samples {100 * g[2] / TS:6.2f}%  lanes {g[1] / max(g[0], 1):5.1f}")
