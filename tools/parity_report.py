import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np
from conftest import load_pkg, golden, cam_for
import oracle_lib as orc
rt = load_pkg()
for name, W, H, spp in [("three", 96, 54, 32), ("cover", 64, 36, 16), ("cover", 160, 90, 8)]:
    rows = golden(f"{name}_small.npz")["spheres"]
    scene = rt.Scene(rows)
    rgb, lin, st = rt.render_image(scene, cam_for(rt, name, W, H), rt.Config(W, H, spp, 50, seed=7))
    olin, samples, rays = orc.render(rows, orc.scene_camera(name, W, H), W, H, spp, mode=orc.RNG_PHILOX, seed=7)
    err = np.abs(lin / spp - olin / spp)
    oq = orc.write_color(olin, spp)
    print(name, W, H, spp, "median", np.median(err), "mean", err.mean(), "p99", np.percentile(err, 99), "rays", st.rays, rays, "8bit>1:", (np.abs(rgb.astype(int) - oq.astype(int)) > 1).mean(), "8bit!=:", (rgb != oq).mean())
name, W, H = "three", 128, 72
rows = golden(f"{name}_small.npz")["spheres"]; scene = rt.Scene(rows)
for off in (0, 5, 123456):
    cfg = rt.Config(W, H, 1, 50, seed=3, sample_offset=off)
    rgb, lin, st = rt.render_image(scene, cam_for(rt, name, W, H), cfg)
    olin, _, _ = orc.render(rows, orc.scene_camera(name, W, H), W, H, 1, mode=orc.RNG_PHILOX, seed=3, sample_offset=off)
    print("single", off, (np.abs(lin - olin).max(axis=2) < 1e-4).mean())
