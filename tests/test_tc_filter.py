"""The brute-force filter on the tensor cores (csrc/rt_tc.cu, fused_tc_kernel in csrc/rt_wave.cu):
scenes of 32 to 960 spheres run `every ray against every sphere` (src/cpu/hittable_list.h:28-43) as
one 3xTF32 contraction with tcgen05.mma.  It only has to be CONSERVATIVE (never drop a line-sphere
pass); the f64 resolve behind it decides like the reference.  So: (1) images, linear sums and ray
counts are byte-identical to the FFMA2 filter loop (RT_FLAG_SIMT_FILTER) in every model; (2) the
probe's pass masks contain every pass of the exact line test in f64 on real rays, and the
contraction's error stays under the kappa bound the slack is built from."""
import numpy as np
import pytest

from conftest import cam_for, golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,W,H,spp", [("cover", 240, 135, 32), ("three", 200, 112, 64), ("cover", 1200, 675, 2)])
@pytest.mark.parametrize("flags", [0, 16, 16 | 32])  # src/cpu model, v4 model, v4 + f32 hit
def test_tensor_core_and_simt_filters_render_the_same_bytes(rt, name, W, H, spp, flags):
    rows = golden(f"{name}_small.npz")["spheres"]
    scene, cam = rt.Scene(rows), cam_for(rt, name, W, H)
    a = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=3, flags=flags))
    b = rt.render_image(scene, cam, rt.Config(W, H, spp, 50, seed=3, flags=flags | rt.FLAG_SIMT_FILTER))
    assert a[2].rays == b[2].rays and a[2].samples == b[2].samples
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[0], b[0])
    assert a[1].max() > 0
    # rt_stats says which unit ran the filter (bench.py's roofline line quotes it)
    assert a[2].filter_unit == (rt.FILTER_TENSOR if len(rows) >= 32 else rt.FILTER_SIMT)
    assert b[2].filter_unit == rt.FILTER_SIMT


@pytest.mark.parametrize("copies", [3, 6])
def test_scenes_beyond_the_operand_limit_stream_it_through_shared_memory(rt, copies):
    """more than 960 spheres: the sphere operand no longer fits a CTA's shared memory and streams through
    two chunk buffers (fused_tcs_kernel); same bytes as the FFMA2 loop and as the BVH."""
    cover = golden("cover_small.npz")["spheres"]
    big = np.concatenate([cover] + [cover[1:] + np.array([0, 0, 30.0 * k, 0, 0, 0, 0, 0, 0]) for k in range(1, copies)])
    assert len(big) > 960
    W, H = 96, 54
    cam = cam_for(rt, "cover", W, H)
    a = rt.render_image(rt.Scene(big), cam, rt.Config(W, H, 4, 50, seed=2))
    s = rt.render_image(rt.Scene(big), cam, rt.Config(W, H, 4, 50, seed=2, flags=rt.FLAG_SIMT_FILTER))
    b = rt.render_image(rt.Scene(big), cam, rt.Config(W, H, 4, 50, seed=2, flags=rt.FLAG_BVH))
    assert np.array_equal(a[1], s[1]) and np.array_equal(a[1], b[1]) and np.array_equal(a[0], b[0])
    assert a[2].rays == s[2].rays == b[2].rays
    assert a[2].filter_unit == rt.FILTER_TENSOR and s[2].filter_unit == rt.FILTER_SIMT and b[2].filter_unit == rt.FILTER_NONE


def probe_rays(rows, n, seed):
    rng = np.random.default_rng(seed)
    k = rng.integers(0, len(rows), n)
    nrm = rng.standard_normal((n, 3)); nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    o = rows[k, :3] + np.abs(rows[k, 3:4]) * nrm * rng.choice([1.0, 1.0, 3.0], (n, 1))
    o[: n // 4] = np.array([13.0, 2.0, 3.0])  # camera position
    d = rng.standard_normal((n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    ro = np.concatenate([o, np.ones((n, 1))], axis=1).astype(np.float32)
    rd = np.concatenate([d, np.zeros((n, 1))], axis=1).astype(np.float32)
    rd[:, :3] /= np.linalg.norm(rd[:, :3].astype(np.float64), axis=1, keepdims=True).astype(np.float32)
    return ro, rd


@pytest.mark.parametrize("name", ["cover", "three"])
def test_probe_masks_contain_every_exact_line_pass(rt, name):
    rows = golden(f"{name}_small.npz")["spheres"]
    n = 4096
    ro, rd = probe_rays(rows, n, seed=5)
    bits, vals, ms = rt.tc_filter_probe(rt.Scene(rows), ro, rd, n_values=n)
    ns = len(rows)
    o, d = ro[:, :3].astype(np.float64), rd[:, :3].astype(np.float64)
    c, r = rows[:, :3], rows[:, 3]
    oc = c[None, :, :] - o[:, None, :]
    b = np.einsum("rsk,rk->rs", oc, d)
    e = b * b - ((oc * oc).sum(axis=2) - r[None, :] ** 2) * (d * d).sum(axis=1)[:, None]
    exact = e >= 0
    assert not (exact & ~bits[:, :ns]).any(), "the tensor-core filter dropped a true line-sphere pass"
    assert not bits[:, ns:].any()  # padding rows never pass
    # selectivity: not more than 25 % above the exact line test on these rays
    assert bits[:, :ns].sum() <= 1.25 * exact.sum() + n
    # error of the contraction against f64, in units of the bound the slack is built from
    kappa = 2.0 ** -16
    c2, o2 = (c * c).sum(axis=1), (o * o).sum(axis=1)
    bb = (c @ d.T - (o * d).sum(axis=1)[None, :]).T
    e64 = bb * bb + 2 * (o @ c.T) + (r * r - c2)[None, :] - o2[:, None]
    slack = kappa * (2 * c2 + np.abs(r * r - c2))[None, :] + ((2 * kappa + 4e-7) * o2)[:, None]
    err = np.abs(vals[:, :ns] - (e64 + slack)) / (kappa * ((2 * c2 + np.abs(r * r - c2))[None, :] + (2 * o2)[:, None]))
    assert err.max() < 0.25, err.max()  # measured 0.04


@pytest.mark.parametrize("n", [32, 33, 63, 64, 65, 128, 129, 448, 959, 960, 961, 1024, 1025, 1344, 1345, 1537])
def test_sphere_counts_around_tile_and_chunk_boundaries(rt, n):
    """the cover scene tiled / truncated to n spheres: whole and ragged last tiles, the resident / streamed limit
    (960), whole and ragged last chunks (6 tiles = 384 spheres).  Tensor cores and FFMA2 loop: same bytes."""
    cover = golden("cover_small.npz")["spheres"]
    reps = -(-n // len(cover))
    rows = np.concatenate([cover] + [cover[1:] + np.array([0, 0, 30.0 * k, 0, 0, 0, 0, 0, 0]) for k in range(1, reps + 1)])[:n]
    W, H = 80, 45
    cam = cam_for(rt, "cover", W, H)
    a = rt.render_image(rt.Scene(rows), cam, rt.Config(W, H, 4, 50, seed=n))
    s = rt.render_image(rt.Scene(rows), cam, rt.Config(W, H, 4, 50, seed=n, flags=rt.FLAG_SIMT_FILTER))
    assert a[2].filter_unit == rt.FILTER_TENSOR and s[2].filter_unit == rt.FILTER_SIMT
    assert a[2].rays == s[2].rays
    assert np.array_equal(a[1], s[1]) and np.array_equal(a[0], s[0])
