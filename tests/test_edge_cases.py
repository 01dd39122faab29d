"""Edge cases of the hot path through the C ABI (GPU): empty world, ragged sample counts (canonical
slices with zero samples), smallest and odd image sizes, depth 0/1, candidate-list overflow, and
size-independent properties at BASELINE.json's full frame size (progressive accumulation,
determinism, sample/ray accounting)."""
import numpy as np
import pytest

from conftest import cam_for, golden

pytestmark = pytest.mark.gpu


def test_empty_world_is_the_sky_gradient(rt, orc):
    W, H, spp = 64, 36, 8
    rows = np.zeros((0, 9))
    scene = rt.Scene(rows)
    cam = cam_for(rt, "cover", W, H)
    rgb, lin, st = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=1))
    olin, samples, rays = orc.render(rows, orc.scene_camera("cover", W, H), W, H, spp, seed=1)
    assert st.samples == samples == W * H * spp and st.rays == rays == samples  # one segment per sample
    assert np.abs(lin - olin).max() / spp < 1e-5
    assert np.all(lin[..., 2] / spp > 0.99)  # blue channel of (1-t)*1 + t*1.0


@pytest.mark.parametrize("spp", [1, 2, 7, 13, 63, 65])
def test_ragged_sample_counts(rt, orc, spp):
    """spp below / not a multiple of the 64 canonical slices: some slices hold no sample."""
    W, H = 48, 27
    rows = golden("three_small.npz")["spheres"]
    rgb, lin, st = rt.render_image(rt.Scene(rows), cam_for(rt, "three", W, H), rt.Config(W, H, spp, 50, seed=4))
    olin, samples, rays = orc.render(rows, orc.scene_camera("three", W, H), W, H, spp, seed=4)
    assert st.samples == samples == W * H * spp
    assert abs(st.rays - rays) <= max(3, 2e-3 * rays)
    err = np.abs(lin - olin) / spp
    assert np.median(err) < 1e-6 and err.mean() < 2e-4


@pytest.mark.parametrize("W,H", [(2, 2), (3, 5), (101, 37), (257, 3)])
def test_small_and_odd_image_sizes(rt, orc, W, H):
    rows = golden("three_small.npz")["spheres"]
    spp = 8
    cam = rt.Camera(lookfrom=(3.0, 3.0, 2.0), lookat=(0.0, 0.0, -1.0), vfov=20.0, aspect_ratio=W / H,
                    aperture=0.0, focus_dist=1.0)
    ocam = orc.camera((3.0, 3.0, 2.0), (0.0, 0.0, -1.0), (0.0, 1.0, 0.0), 20.0, W / H, 0.0, 1.0)
    rgb, lin, st = rt.render_image(rt.Scene(rows), cam, rt.Config(W, H, spp, 50, seed=6))
    olin, samples, rays = orc.render(rows, ocam, W, H, spp, seed=6)
    assert rgb.shape == (H, W, 3) and st.samples == samples
    assert (np.abs(lin - olin).max(axis=2) / spp < 1e-4).mean() >= 0.99


def test_image_size_limits_are_checked(rt):
    rows = golden("three_small.npz")["spheres"]
    with pytest.raises(rt.RtError):
        rt.render_image(rt.Scene(rows), cam_for(rt, "three", 1, 1), rt.Config(1, 1, 1, 50))
    with pytest.raises(rt.RtError):  # 2^26 pixels: the item encoding holds 26 pixel bits
        rt.Frame(8192, 8192)


@pytest.mark.parametrize("depth", [0, 1, 2])
def test_bounce_limit(rt, orc, depth):
    """depth 0 is black (main.cc:16-17); depth 1 sees the sky only where the primary ray misses."""
    W, H, spp = 64, 36, 8
    rows = golden("three_small.npz")["spheres"]
    rgb, lin, st = rt.render_image(rt.Scene(rows), cam_for(rt, "three", W, H), rt.Config(W, H, spp, depth, seed=2))
    olin, samples, rays = orc.render(rows, orc.scene_camera("three", W, H), W, H, spp, depth=depth, seed=2)
    if depth == 0:
        assert not lin.any() and not rgb.any() and not olin.any()
    else:
        assert np.abs(lin - olin).mean() / spp < 1e-4
        assert abs(st.rays - rays) <= 2e-3 * rays + 2


def test_candidate_overflow_falls_back_to_the_full_scan(rt, orc):
    """A ray down a row of 300 glass beads passes the f32 filter for every one of them: 75 groups,
    more than the 64 candidate entries a slot holds, so the shade pass must test every sphere
    exactly."""
    W, H, spp = 32, 18, 4
    n = 300
    rows = np.zeros((n + 1, 9))
    rows[0] = [0.0, -1000.5, 0.0, 1000.0, 0, 0.5, 0.5, 0.5, 0.0]
    for k in range(n):
        rows[k + 1] = [0.0, 0.0, -2.0 - 0.35 * k, 0.15, 2, 1.0, 1.0, 1.0, 1.5]
    cam = rt.Camera(lookfrom=(0.0, 0.0, 0.0), lookat=(0.0, 0.0, -1.0), vfov=8.0, aspect_ratio=W / H,
                    aperture=0.0, focus_dist=1.0)
    ocam = orc.camera((0.0, 0.0, 0.0), (0.0, 0.0, -1.0), (0.0, 1.0, 0.0), 8.0, W / H, 0.0, 1.0)
    rgb, lin, st = rt.render_image(rt.Scene(rows), cam, rt.Config(W, H, spp, 50, seed=8))
    olin, samples, rays = orc.render(rows, ocam, W, H, spp, seed=8)
    assert rays / samples > 4  # the beads are really traversed
    assert abs(st.rays - rays) / rays < 0.02
    assert np.median(np.abs(lin - olin)) / spp < 1e-5
    blin = rt.render_image(rt.Scene(rows), cam, rt.Config(W, H, spp, 50, seed=8, flags=rt.FLAG_BVH))[1]
    assert np.array_equal(blin, lin)


# ----------------------------------------------------------------- full-size properties
def test_full_size_progressive_and_accounting(rt):
    """BASELINE.json configs[1] frame size (1200x675, 486 spheres): (i) two progressive halves
    (sample_offset) add up to the one-shot frame — the same Philox keys, only the f32 summation
    order differs; (ii) rendering twice is byte-identical; (iii) sample, ray and test counters
    obey samples = W*H*spp, tests = rays * N, 1 <= rays/sample <= 50."""
    W, H, spp = 1200, 675, 32
    rows = golden("cover_small.npz")["spheres"]
    scene = rt.Scene(rows)
    cam = cam_for(rt, "cover", W, H)
    rgb, lin, st = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=3))
    rgb2, lin2, st2 = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=3))
    assert np.array_equal(rgb, rgb2) and np.array_equal(lin, lin2) and st.rays == st2.rays
    a = rt.render_image(scene, cam, rt.Config(W, H, spp // 2, 50, seed=3))
    b = rt.render_image(scene, cam, rt.Config(W, H, spp // 2, 50, seed=3, sample_offset=spp // 2))
    assert a[2].rays + b[2].rays == st.rays
    assert np.abs((a[1] + b[1]) - lin).max() <= 1e-5 * spp
    assert st.samples == W * H * spp and st.sphere_tests == st.rays * len(rows)
    assert 1.0 <= st.rays / st.samples <= 50.0
    assert lin.min() >= 0.0 and lin.max() <= spp * (1 + 1e-6)  # radiance per sample is in [0, 1]
