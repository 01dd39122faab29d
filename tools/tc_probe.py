#!/usr/bin/env python
"""The brute-force filter as a tf32 contraction on the tensor cores (csrc/rt_tc.cu): validation and
measurement on real rays of the cover scene.

    python tools/tc_probe.py [--rays 262144] [--repeats 20]

(1) conservativeness: every (ray, sphere) pair whose LINE meets the sphere in f64 must be in the
    tensor-core pass mask; (2) selectivity: passes per ray of the tensor-core filter, of the f32 SIMT
    filter's formula and of the exact line test; (3) numerical error of the contraction against f64,
    in units of kappa * (A_s + B_r); (4) throughput of MMA + sign-mask epilogue in (ray, sphere)
    tests per second, next to the SIMT filter loop's."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import _load_pkg  # noqa: E402


def make_rays(rows, n, seed=1):
    """half camera rays of the cover view, half secondary rays leaving random spheres"""
    rng = np.random.default_rng(seed)
    m = n // 2
    look = np.array([13.0, 2.0, 3.0])
    w = look / np.linalg.norm(look)
    u = np.cross([0, 1, 0], w); u /= np.linalg.norm(u)
    v = np.cross(w, u)
    h = np.tan(np.radians(20.0) / 2)
    s, t = rng.random(m), rng.random(m)
    d = -10.0 * w + (s[:, None] - 0.5) * (2 * h * 16 / 9 * 10.0) * u + (t[:, None] - 0.5) * (2 * h * 10.0) * v
    o = np.tile(look, (m, 1)) + 0.05 * rng.standard_normal((m, 3)) * 0
    k = rng.integers(0, len(rows), n - m)
    nrm = rng.standard_normal((n - m, 3)); nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    nrm[:, 1] = np.abs(nrm[:, 1])  # upper side (the ground is below)
    o2 = rows[k, :3] + np.abs(rows[k, 3:4]) * nrm
    d2 = nrm + rng.standard_normal((n - m, 3)); d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    o = np.concatenate([o, o2]); d = np.concatenate([d, d2])
    ln = np.linalg.norm(d, axis=1, keepdims=True)
    ro = np.concatenate([o, ln], axis=1).astype(np.float32)
    rd = np.concatenate([d / ln, np.zeros((n, 1))], axis=1).astype(np.float32)
    return ro, rd


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=262144)
    ap.add_argument("--repeats", type=int, default=20)
    ap.add_argument("--check", type=int, default=8192, help="rays compared pair by pair against f64")
    a = ap.parse_args()
    rt = _load_pkg()
    rows = np.load(os.path.join(ROOT, "tests", "golden", "cover_small.npz"))["spheres"]
    scene = rt.Scene(rows)
    ro, rd = make_rays(rows, a.rays)
    nv = min(a.check, a.rays)
    bits, vals, ms1 = rt.tc_filter_probe(scene, ro, rd, n_values=nv, repeats=1)
    _, _, msr = rt.tc_filter_probe(scene, ro, rd, repeats=a.repeats)
    n = len(rows)
    bits = bits[:, :n]
    # ---- exact line test in f64 on the f32 rays
    o, d = ro[:nv, :3].astype(np.float64), rd[:nv, :3].astype(np.float64)
    c, r = rows[:, :3], rows[:, 3]
    oc = c[None, :, :] - o[:, None, :]
    b = np.einsum("rsk,rk->rs", oc, d)
    e_exact = b * b - ((oc * oc).sum(axis=2) - r[None, :] ** 2) * (d * d).sum(axis=1)[:, None]
    exact = e_exact >= 0
    missed = exact & ~bits[:nv]
    # ---- the SIMT filter's formula (f32 rows, slack 32 * 2^-24), evaluated in f64
    slack = 32 * 2.0 ** -24
    c2 = (c * c).sum(axis=1)
    wrow = (r * r - c2) + slack * (c2 + r * r)
    o2 = (o * o).sum(axis=1)
    bb = c @ d.T - (o * d).sum(axis=1)[None, :]
    e_simt = (bb * bb + 2 * (c @ o.T) + wrow[:, None]).T >= (o2 * (1 - slack))[:, None]
    # ---- numerical error of the contraction against the same quantity in f64 (without slack)
    kappa = 2.0 ** -16
    A = 2 * c2 + np.abs(r * r - c2)
    B = 2 * o2
    od = (o * d).sum(axis=1)
    e64 = (bb * bb).T + 2 * (o @ c.T) + (r * r - c2)[None, :] - o2[:, None]
    w_slack = kappa * A
    k_slack = (2 * kappa + 4e-7) * o2
    err = vals[:, :n].astype(np.float64) - (e64 + w_slack[None, :] + k_slack[:, None])
    scale = kappa * (A[None, :] + B[:, None])
    pairs = a.rays * n
    out = {
        "rays": a.rays, "spheres": n, "checked_rays": nv,
        "exact_line_passes_missing_from_tc_mask": int(missed.sum()),
        "passes_per_ray": {"exact_line": float(exact.sum(axis=1).mean()), "simt_f32_filter": float(e_simt.sum(axis=1).mean()),
                           "tensor_core_filter": float(bits[:nv].sum(axis=1).mean())},
        "error_over_kappa_bound": {"max": float(np.abs(err / scale).max()), "p999": float(np.quantile(np.abs(err / scale), 0.999)),
                                   "kappa": kappa, "note": "|e_tc - e_f64| / (kappa (A_s + B_r)); must stay below 1"},
        "ms_one_pass": ms1, "ms_per_pass": msr / a.repeats,
        "tests_per_s": pairs / (msr / a.repeats * 1e-3),
        "credited_tflops_17_per_test": pairs * 17 / (msr / a.repeats * 1e-3) / 1e12,
        "simt_filter_loop_tests_per_s_for_comparison": 5.32e11 / 0.118,
    }
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
