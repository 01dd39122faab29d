"""Soundness of the conservative f32 line-sphere filter: with
RT_FLAG_NO_FILTER the pre-test is switched off and every ray resolves EVERY sphere in f64, like
hittable_list::hit as written (src/cpu/hittable_list.h:28-43, sphere.h:24-51).  If the filter ever
dropped a true hit, some path would change; the image and the ray count must therefore be
BYTE-IDENTICAL with and without the flag.  Scenes: the reference's two, a row of 300 beads
(candidate overflow), the 2.5 k-sphere stress scene (rows fed through L1), and views from far
away (|o| ~ 1e4, where the |o|^2 slack term dominates)."""
import numpy as np
import pytest

from conftest import cam_for, golden
from test_bvh import stress_rows

pytestmark = pytest.mark.gpu


def identical(rt, rows, cam, W, H, spp, seed=9, flags=0):
    scene = rt.Scene(rows)
    a = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=seed, flags=flags))
    b = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=seed, flags=flags | rt.FLAG_NO_FILTER))
    assert a[2].rays == b[2].rays and a[2].samples == b[2].samples
    assert np.array_equal(a[1], b[1]), f"{(a[1] != b[1]).any(axis=2).sum()} pixels differ in linear sums"
    assert np.array_equal(a[0], b[0])
    assert a[1].max() > 0
    return a


@pytest.mark.parametrize("name,W,H,spp", [("cover", 240, 135, 16), ("three", 200, 112, 32)])
def test_reference_scenes(rt, name, W, H, spp):
    rows = golden(f"{name}_small.npz")["spheres"]
    identical(rt, rows, cam_for(rt, name, W, H), W, H, spp)
    identical(rt, rows, cam_for(rt, name, W, H), W, H, spp // 2, seed=10, flags=rt.FLAG_V4_MODEL)


def test_bead_row_overflow(rt):
    n = 300
    rows = np.zeros((n + 1, 9))
    rows[0] = [0.0, -1000.5, 0.0, 1000.0, 0, 0.5, 0.5, 0.5, 0.0]
    for k in range(n):
        rows[k + 1] = [0.0, 0.0, -2.0 - 0.35 * k, 0.15, 2, 1.0, 1.0, 1.0, 1.5]
    cam = rt.Camera(lookfrom=(0.0, 0.0, 0.0), lookat=(0.0, 0.0, -1.0), vfov=8.0, aspect_ratio=32 / 18,
                    aperture=0.0, focus_dist=1.0)
    identical(rt, rows, cam, 32, 18, 4)


def test_stress_scene_2500(rt, tmp_path):
    rows = stress_rows(25, tmp_path)
    assert len(rows) > 2000
    identical(rt, rows, cam_for(rt, "cover", 96, 54), 96, 54, 4)


@pytest.mark.parametrize("dist", [1.0e3, 1.0e4])
def test_far_camera(rt, dist):
    """The cover scene seen through a narrow lens from |o| ~ dist: the threshold's slack term
    32 * 2^-24 * |o|^2 is ~200 at 1e4, far above r^2 = 0.04 of the small spheres -- the filter
    passes almost everything there, and must still never drop a true hit."""
    rows = golden("cover_small.npz")["spheres"]
    d = np.array([13.0, 2.0, 3.0]) / np.linalg.norm([13.0, 2.0, 3.0])
    cam = rt.Camera(lookfrom=tuple(d * dist), lookat=(0.0, 0.5, 0.0), vfov=20.0 * 13.0 / dist,
                    aspect_ratio=96 / 54, aperture=0.0, focus_dist=dist)
    a = identical(rt, rows, cam, 96, 54, 8)
    assert a[2].rays > 1.5 * a[2].samples  # the spheres are really in view


def soup(seed, n, spread, rmin, rmax, offset=(0.0, 0.0, 0.0)):
    """random spheres of all three materials around `offset` (some hollow, some overlapping) on a big ground"""
    rng = np.random.default_rng(seed)
    rows = np.zeros((n + 1, 9))
    off = np.asarray(offset)
    rows[0] = [off[0], off[1] - 1000.0 - spread, off[2], 1000.0, 0, 0.5, 0.5, 0.5, 0.0]
    c = rng.uniform(-spread, spread, (n, 3)) + off
    r = np.exp(rng.uniform(np.log(rmin), np.log(rmax), n))
    kind = rng.integers(0, 3, n)
    r = np.where((kind == 2) & (rng.random(n) < 0.3), -0.9 * r, r)  # hollow bubbles
    rows[1:, :3] = c; rows[1:, 3] = r; rows[1:, 4] = kind
    rows[1:, 5:8] = rng.uniform(0.2, 1.0, (n, 3))
    rows[1:, 8] = np.where(kind == 2, 1.5, rng.uniform(0.0, 0.5, n))
    return rows


@pytest.mark.parametrize("seed,n,spread,rmin,rmax,offset", [
    (1, 200, 4.0, 0.05, 1.0, (0, 0, 0)),          # dense soup, many overlaps, rays start inside spheres
    (2, 700, 30.0, 0.01, 3.0, (0, 0, 0)),         # radii over 2.5 decades
    (3, 500, 5.0, 0.002, 0.05, (0, 0, 0)),        # tiny spheres: r^2 down to 4e-6 against |c|^2 ~ 50
    (4, 300, 5.0, 0.1, 1.0, (3000.0, -2000.0, 1500.0)),  # everything far from the origin: |c|^2 ~ 1.5e7
    (5, 1500, 20.0, 0.05, 1.5, (0, 0, 0)),        # beyond 960 spheres: streamed operand
])
def test_random_soups_both_filters(rt, seed, n, spread, rmin, rmax, offset):
    """Random geometry the reference scenes never show.  The tensor-core filter (default) and the FFMA2 loop
    must each render the bytes of the unfiltered f64 scan."""
    rows = soup(seed, n, spread, rmin, rmax, offset)
    off = np.asarray(offset, float)
    W, H = 80, 45
    cam = rt.Camera(lookfrom=tuple(off + np.array([2.5 * spread, 0.6 * spread, 1.5 * spread])), lookat=tuple(off),
                    vfov=35.0, aspect_ratio=W / H, aperture=0.02 * spread, focus_dist=3.0 * spread)
    a = identical(rt, rows, cam, W, H, 6, seed=seed)
    assert a[2].filter_unit == rt.FILTER_TENSOR
    s = identical(rt, rows, cam, W, H, 6, seed=seed, flags=rt.FLAG_SIMT_FILTER)
    assert np.array_equal(a[1], s[1]) and a[2].rays == s[2].rays
    assert a[2].rays > 1.2 * a[2].samples
