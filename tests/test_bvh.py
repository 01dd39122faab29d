"""Optional BVH mode (north-star: "brute-force over the reference's sphere order ... plus an
optional BVH built on the host").  The traversal must return exactly the sphere the reference's
linear scan returns (smallest root, later sphere on ties), so with the same Philox keys the BVH
image is BIT-IDENTICAL to the brute-force image — on the reference's scenes and on the larger
synthetic scenes of BASELINE.json configs[3] built by the façade's stress_scene()."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, cam_for, golden

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "examples", "rt_cover")


def stress_rows(half, tmp_path):
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "examples"), "rt_cover"], stdout=subprocess.DEVNULL)
    p = tmp_path / f"stress{half}.f64"
    subprocess.check_call([BIN, "--scene", "stress", "--half", str(half), "--dump-scene", str(p)])
    return np.fromfile(p, np.float64).reshape(-1, 9)


def both(rt, rows, cam, W, H, spp, seed=3):
    scene = rt.Scene(rows)
    a = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=seed))
    b = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=seed, flags=rt.FLAG_BVH))
    return a, b


@pytest.mark.parametrize("name,W,H,spp", [("cover", 160, 90, 32), ("three", 160, 90, 64)])
def test_bvh_image_identical_on_reference_scenes(rt, name, W, H, spp):
    rows = golden(f"{name}_small.npz")["spheres"]
    (rgb0, lin0, st0), (rgb1, lin1, st1) = both(rt, rows, cam_for(rt, name, W, H), W, H, spp)
    assert st0.rays == st1.rays and st0.samples == st1.samples
    assert np.array_equal(lin0, lin1) and np.array_equal(rgb0, rgb1)


@pytest.mark.parametrize("half,W,H,spp", [(25, 128, 72, 8), (158, 64, 36, 2)])
def test_bvh_image_identical_on_stress_scenes(rt, tmp_path, half, W, H, spp):
    """~2.5 k spheres (brute force reads its filter rows through L1, beyond the constant bank) and
    the ~100 k-sphere scene of configs[3]."""
    rows = stress_rows(half, tmp_path)
    assert len(rows) > (2000 if half == 25 else 99000)
    (rgb0, lin0, st0), (rgb1, lin1, st1) = both(rt, rows, cam_for(rt, "cover", W, H), W, H, spp)
    assert st0.rays == st1.rays
    assert np.array_equal(lin0, lin1) and np.array_equal(rgb0, rgb1)
    assert lin0.max() > 0


def test_brute_force_beyond_constant_bank_vs_oracle(rt, orc, tmp_path):
    """Scenes of more than ~1000 rows take the L1-fed variant of the filter loop: check it against
    the f64 oracle on the same Philox stream (same tolerances as test_same_stream_vs_oracle)."""
    rows = stress_rows(18, tmp_path)  # ~1300 spheres
    assert 1016 < len(rows) < 2000
    W, H, spp = 64, 36, 8
    scene = rt.Scene(rows)
    rgb, lin, st = rt.render_image(scene, cam_for(rt, "cover", W, H), rt.Config(W, H, spp, 50, seed=7))
    olin, samples, rays = orc.render(rows, orc.scene_camera("cover", W, H), W, H, spp,
                                     mode=orc.RNG_PHILOX, seed=7)
    err = np.abs(lin / spp - olin / spp)
    assert st.samples == samples
    assert np.median(err) < 1e-5 and err.mean() < 2e-3
    assert abs(st.rays - rays) / rays < 5e-3


def test_bvh_handout_order_does_not_change_the_image(rt):
    """BVH frames order their next render differently from brute-force frames (heavy pixels first,
    every other pixel in scan order: csrc/rt_kernels.cu, cost_keys_kernel) and a frame object
    falls back to scan order when the mode changes between renders.  All of that is scheduling:
    every render of the same scene on a re-used frame object is byte-identical to a fresh
    brute-force render."""
    W, H, spp = 160, 90, 48
    scene = rt.Scene(golden("cover_small.npz")["spheres"])
    cam = cam_for(rt, "cover", W, H)

    def render(frame, flags):
        cfg = rt.Config(W, H, spp, 50, seed=21, flags=flags)
        frame.render(scene, cam, cfg)
        frame.reduce_local(cfg)
        rgb, lin, st = frame.finalize(cfg, want_linear=True)
        return rgb.copy(), lin.copy(), st.rays

    fresh = render(rt.Frame(W, H), 0)
    fr = rt.Frame(W, H)
    for flags in (rt.FLAG_BVH, rt.FLAG_BVH, 0, 0, rt.FLAG_BVH):  # scan, heavy-first, scan, cost order, scan
        rgb, lin, rays = render(fr, flags)
        assert rays == fresh[2]
        assert np.array_equal(rgb, fresh[0]) and np.array_equal(lin, fresh[1])
