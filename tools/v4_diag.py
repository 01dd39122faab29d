"""Diagnostic: tile-wise mean signed difference between RT_FLAG_V4_MODEL and the reference CUDA program."""
import os, subprocess, sys
import numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from conftest import load_pkg
from test_v4_model import _meta, HARNESS
rt = load_pkg()
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 200
prefix = "/tmp/refdiag"
subprocess.check_call([HARNESS, prefix, str(spp), "7"])
m = _meta(prefix + ".meta.txt")
W, H, n = int(m["W"]), int(m["H"]), int(m["N"])
rows = np.fromfile(prefix + ".scene.f64", np.float64).reshape(n, 9)
ref = np.fromfile(prefix + ".linear.f32", np.float32).reshape(H, W, 3).astype(np.float64) / spp
aperture = 2.0 * m["focus_dist"] * np.tan(np.radians(m["defocus_angle"] / 2.0))
cam = rt.Camera(lookfrom=tuple(m["lookfrom"]), lookat=tuple(m["lookat"]), vup=tuple(m["vup"]),
                vfov=m["vfov"], aspect_ratio=W / H, aperture=float(aperture), focus_dist=m["focus_dist"])
scene = rt.Scene(rows)
np.set_printoptions(linewidth=250, precision=1, suppress=True)
for name, flags in (("v4", rt.FLAG_V4_MODEL), ("v3", 0)):
    for seed in (1, 2):
        lin = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=seed, flags=flags))[1] / spp
        T = 120
        d = (lin - ref)
        tiles = d.reshape(H // T, T, W // T, T, 3).mean(axis=(1, 3))
        print(name, "seed", seed, "bias", d.mean(axis=(0, 1)), "tile max", np.abs(tiles).max())
        print((tiles.mean(axis=2) * 1e4))
    if name == "v4":
        a = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=1, flags=flags))[1] / spp
        b = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=2, flags=flags))[1] / spp
        tiles = (a - b).reshape(H // T, T, W // T, T, 3).mean(axis=(1, 3))
        print("ours seed1 - ours seed2: bias", (a - b).mean(axis=(0, 1)), "tile max", np.abs(tiles).max())
# material census
print("kinds", np.bincount(rows[:, 4].astype(int)))
