#!/usr/bin/env python
"""Small renders that touch every kernel variant, for compute-sanitizer (memcheck / racecheck):
smoke() (primary gate + brute force), BVH, v4 model + f32 hit, NO_FILTER, candidate-list overflow
(bead row), a 1456-sphere scene (tensor-core filter with the operand streamed through shared memory, then
the FFMA2 loop with rows through L1), the SIMT form of the cover scene, the static-scratch primary gate, the
banded finalise.     compute-sanitizer --tool memcheck python tools/sanitize_cases.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g  # noqa: E402

g.smoke()
rt = g._load_pkg()
W, H = 64, 36
cover = np.load(os.path.join(ROOT, "tests", "golden", "cover_small.npz"))["spheres"]
cam = rt.Camera(aspect_ratio=W / H)
sc = rt.Scene(cover)
base = rt.render_image(sc, cam, rt.Config(W, H, 4, 50, seed=1))
for flags in (rt.FLAG_BVH, rt.FLAG_SIMT_FILTER, rt.FLAG_V4_MODEL, rt.FLAG_V4_MODEL | rt.FLAG_F32_HIT, rt.FLAG_NO_FILTER):
    out = rt.render_image(sc, cam, rt.Config(W, H, 4, 50, seed=1, flags=flags))
    same = np.array_equal(out[1], base[1])
    print(f"flags {flags:3d}: rays {out[2].rays}  identical to brute force: {same}")
    assert same or flags & rt.FLAG_V4_MODEL
n = 300
rows = np.zeros((n + 1, 9))
rows[0] = [0.0, -1000.5, 0.0, 1000.0, 0, 0.5, 0.5, 0.5, 0.0]
for k in range(n):
    rows[k + 1] = [0.0, 0.0, -2.0 - 0.35 * k, 0.15, 2, 1.0, 1.0, 1.0, 1.5]
bcam = rt.Camera(lookfrom=(0.0, 0.0, 0.0), lookat=(0.0, 0.0, -1.0), vfov=8.0, aspect_ratio=32 / 18, aperture=0.0, focus_dist=1.0)
out = rt.render_image(rt.Scene(rows), bcam, rt.Config(32, 18, 2, 50, seed=8))
print("bead row (list overflow): rays", out[2].rays)
big = np.concatenate([cover] + [cover[1:] + np.array([0, 0, 30.0 * k, 0, 0, 0, 0, 0, 0]) for k in (1, 2)])
out = rt.render_image(rt.Scene(big), cam, rt.Config(W, H, 2, 50, seed=2))
outb = rt.render_image(rt.Scene(big), cam, rt.Config(W, H, 2, 50, seed=2, flags=rt.FLAG_BVH))
outs = rt.render_image(rt.Scene(big), cam, rt.Config(W, H, 2, 50, seed=2, flags=rt.FLAG_SIMT_FILTER))
print(f"{len(big)} spheres: rays {out[2].rays}, BVH identical: {np.array_equal(out[1], outb[1])}, "
      f"FFMA2 loop identical: {np.array_equal(out[1], outs[1])}")
a = rt.primary_hits_static(sc, cam, W, H, True)
b = rt.primary_hits(sc, cam, W, H, True)
print("static-scratch primary gate identical:", all(np.array_equal(x, y) for x, y in zip(a, b)))
fr = rt.Frame(W, H)
cfg = rt.Config(W, H, 2, 50, seed=3)
fr.render(sc, cam, cfg); fr.reduce_local(cfg)
fr.finalize_to_file(cfg, "/tmp/sanitize.ppm", rt.PNM_P6)
print("sanitize cases done")
