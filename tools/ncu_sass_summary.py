#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --csv` (SASS view): executed warp instructions and
stall samples per opcode, plus the hottest address ranges.  Used to write profiles/*.md.

    ncu -i gpurun_out/prof.ncu-rep --page source --csv > /tmp/src.csv
    python tools/ncu_sass_summary.py /tmp/src.csv [--top 25] [--dump lo:hi]
"""
import argparse
import collections
import csv
import re


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--top", type=int, default=25)
    ap.add_argument("--dump", default=None, help="row range lo:hi to print verbatim")
    a = ap.parse_args()
    rows = list(csv.reader(open(a.csv)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    col = {h: i for i, h in enumerate(hdr)}
    body = [r for r in rows[hdr_i + 1:] if len(r) >= len(hdr) - 2]
    S, N, T, SMP = col["Source"], col["Instructions Executed"], col["Thread Instructions Executed"], col["# Samples"]

    def num(x):
        try:
            return float(x)
        except ValueError:
            return 0.0
    tot_i = sum(num(r[N]) for r in body)
    tot_s = sum(num(r[SMP]) for r in body)
    tot_t = sum(num(r[T]) for r in body)
    print(f"SASS rows {len(body)}  warp-inst {tot_i:.4g}  thread-inst/inst {tot_t / tot_i:.2f}  samples {tot_s:.0f}")
    by_op = collections.defaultdict(lambda: [0.0, 0.0, 0])
    for r in body:
        m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", r[S])
        op = m.group(1) if m else "?"
        base = op.split(".")[0]
        if base in ("LDS", "LDG", "STS", "STG", "LD", "ST", "LDL", "STL"):
            base = op  # keep width
        e = by_op[base]
        e[0] += num(r[N]); e[1] += num(r[SMP]); e[2] += 1
    print(f"{'opcode':<18}{'inst %':>8}{'samples %':>11}{'static':>8}")
    for op, (n, s, c) in sorted(by_op.items(), key=lambda kv: -kv[1][0])[:a.top]:
        print(f"{op:<18}{100 * n / tot_i:8.2f}{100 * s / tot_s:11.2f}{c:8d}")
    # hottest contiguous 32-row windows by samples
    print("hottest windows (rows, inst %, samples %):")
    W = 32
    wins = []
    for lo in range(0, len(body), W):
        seg = body[lo:lo + W]
        wins.append((sum(num(r[SMP]) for r in seg), sum(num(r[N]) for r in seg), lo))
    for s, n, lo in sorted(wins, reverse=True)[:10]:
        print(f"  rows {lo:5d}-{lo + W - 1:5d}  inst {100 * n / tot_i:6.2f}%  samples {100 * s / tot_s:6.2f}%")
    if a.dump:
        lo, hi = (int(x) for x in a.dump.split(":"))
        for k, r in enumerate(body[lo:hi], lo):
            print(f"{k:5d} {num(r[N]):12.4g} {num(r[SMP]):8.0f} thr {r[col['Avg. Threads Executed']]:>5}  {r[S]}")


if __name__ == "__main__":
    main()
