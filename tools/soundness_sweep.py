"""Randomised soundness sweep (run on a GPU box): 60 random sphere soups (33..1400 spheres, extents 0.5..200, radii over
up to 4 decades, scene offsets up to 2e4 from the origin); the tensor-core filter and the FFMA2 loop must each render the
bytes of the unfiltered f64 scan (RT_FLAG_NO_FILTER).  Last run (r02): failures: 0."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import _load_pkg
rt = _load_pkg()
from test_filter_soundness import soup, identical
rng = np.random.default_rng(123)
bad = 0
for k in range(60):
    n = int(rng.integers(33, 1400)); spread = float(np.exp(rng.uniform(np.log(0.5), np.log(200.0))))
    rmin = spread * float(np.exp(rng.uniform(np.log(2e-4), np.log(0.05)))); rmax = rmin * float(np.exp(rng.uniform(0, np.log(100.0))))
    off = tuple(rng.uniform(-1, 1, 3) * float(rng.choice([0.0, 10.0, 1000.0, 20000.0])))
    rows = soup(1000 + k, n, spread, rmin, rmax, off)
    offa = np.asarray(off)
    W, H = 64, 36
    cam = rt.Camera(lookfrom=tuple(offa + np.array([2.5 * spread, 0.6 * spread, 1.5 * spread])), lookat=tuple(offa),
                    vfov=35.0, aspect_ratio=W / H, aperture=0.02 * spread, focus_dist=3.0 * spread)
    try:
        a = identical(rt, rows, cam, W, H, 4, seed=k)
        s = identical(rt, rows, cam, W, H, 4, seed=k, flags=rt.FLAG_SIMT_FILTER)
        assert np.array_equal(a[1], s[1])
        print(k, n, round(spread, 2), "ok rays/sample", round(a[2].rays / a[2].samples, 2), flush=True)
    except AssertionError as e:
        bad += 1
        print(k, n, spread, rmin, rmax, off, "FAIL", e, flush=True)
print("failures:", bad)
