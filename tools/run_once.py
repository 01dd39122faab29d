#!/usr/bin/env python
"""Render one frame through the C ABI and print the stats (used under ncu / for quick A-B runs).

    python tools/run_once.py --scene cover --w 1200 --h 675 --spp 64 [--reps 3] [--ppm out.ppm]
"""
import argparse
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import _load_pkg  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="cover", choices=["cover", "three"])
    ap.add_argument("--w", type=int, default=1200)
    ap.add_argument("--h", type=int, default=675)
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--depth", type=int, default=50)
    ap.add_argument("--reps", type=int, default=1)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--ppm", default=None)
    ap.add_argument("--slices", default=None, help="a:b = render only canonical slices [a,b) (one rank's share)")
    ap.add_argument("--bvh", action="store_true", help="closest hit through the host-built BVH")
    ap.add_argument("--flags", type=int, default=0, help="extra RT_FLAG_* bits (16 = v4 model, 32 = f32 hit)")
    ap.add_argument("--rows", default=None, help="scene dump file (f64 [n][9]) instead of a golden scene")
    ap.add_argument("--nspheres", type=int, default=0, help="truncate / tile the scene to n spheres (timing experiments)")
    a = ap.parse_args()
    rt = _load_pkg()
    rows = np.load(os.path.join(ROOT, "tests", "golden", f"{a.scene}_small.npz"))["spheres"]
    if a.rows:
        rows = np.fromfile(a.rows, np.float64).reshape(-1, 9)
    if a.nspheres:
        reps = -(-a.nspheres // len(rows))
        tiles = []
        for k in range(reps):
            t = rows.copy()
            t[1:, 2] += 30.0 * k  # shift copies along z (keep the ground in place)
            tiles.append(t)
        rows = np.concatenate(tiles)[:a.nspheres]
    if a.scene == "cover":
        cam = rt.Camera(aspect_ratio=a.w / a.h)
    else:
        cam = rt.Camera((3.0, 3.0, 2.0), (0.0, 0.0, -1.0), (0.0, 1.0, 0.0), 20.0, a.w / a.h, 2.0,
                        float(np.sqrt(27.0)))
    scene = rt.Scene(rows)
    frame = rt.Frame(a.w, a.h)
    sb, se = (int(x) for x in a.slices.split(":")) if a.slices else (0, rt.SLICES)
    cfg = rt.Config(a.w, a.h, a.spp, a.depth, seed=a.seed, flags=(rt.FLAG_BVH if a.bvh else 0) | a.flags,
                    slice_begin=sb, slice_end=se)
    for _ in range(a.reps):
        st = frame.render(scene, cam, cfg)
        frame.reduce_local(cfg)
        rgb, _, fin = frame.finalize(cfg)
        tests = fin.rays * len(rows)
        print(json.dumps({
            "render_ms": st.render_ms, "finalize_ms": fin.finalize_ms, "samples": st.samples,
            "rays": fin.rays, "msamples_s": st.samples / st.render_ms / 1e3,
            "mrays_s": fin.rays / st.render_ms / 1e3,
            "tflops_algorithmic": tests * 17 / (st.render_ms * 1e-3) / 1e12,
            "image_sha256": hashlib.sha256(rgb.tobytes()).hexdigest()[:16]}))
    if a.ppm:
        rt.write_ppm(a.ppm, rgb)


if __name__ == "__main__":
    main()
