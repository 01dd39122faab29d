"""Find samples where the BVH render differs from brute force and ask the f64 oracle who is right."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import _load_pkg
import oracle_lib as orc
rt = _load_pkg()
rows = np.load(os.path.join(ROOT, "tests", "golden", "cover_small.npz"))["spheres"]
W, H = 1200, 675
cam = rt.Camera(aspect_ratio=W / H)
ocam = orc.scene_camera("cover", W, H)
scene = rt.Scene(rows)
found = 0
for off in range(0, 64):
    a = rt.render_image(scene, cam, rt.Config(W, H, 1, 50, seed=0, sample_offset=off))
    b = rt.render_image(scene, cam, rt.Config(W, H, 1, 50, seed=0, sample_offset=off, flags=rt.FLAG_BVH))
    d = np.argwhere(np.any(a[1] != b[1], axis=2))
    if len(d):
        print(f"sample {off}: {len(d)} differing pixels, rays {a[2].rays} vs {b[2].rays}")
    for (y, x) in d[:3]:
        olin, _, _ = orc.render(rows, ocam, W, H, 1, mode=orc.RNG_PHILOX, seed=0, sample_offset=off, rows=(int(y), int(y) + 1))
        print(f"   px ({x},{y}) brute {a[1][y, x]} bvh {b[1][y, x]} oracle {olin[0, x]}")
        found += 1
    if found >= 6:
        break
print("done, found", found)
