"""The C-ABI surface around the hot path (GPU): stream-ordered (asynchronous) calls, in-place
scene update, the per-thread frame cache of rt_render_image, the persistent multi-GPU object on
one device, the banded finalise / PNM writers, and the BVH flag on an empty world."""
import numpy as np
import pytest

from conftest import cam_for, golden

pytestmark = pytest.mark.gpu


def test_async_calls_on_a_caller_stream(rt):
    """With a caller-owned stream rt_render / reduce_local / finalize only enqueue: outputs (pinned
    host memory) and stats are valid after the caller's synchronisation and equal the synchronous
    path byte for byte."""
    import torch
    W, H, spp = 160, 90, 32
    rows = golden("cover_small.npz")["spheres"]
    scene, cam, cfg = rt.Scene(rows), cam_for(rt, "cover", W, H), rt.Config(W, H, spp, 50, seed=5)
    ref_rgb, ref_lin, ref_st = rt.render_image(scene, cam, cfg)
    stream = torch.cuda.Stream()
    fr = rt.Frame(W, H)
    fr.set_stream(stream.cuda_stream)
    rgb = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory()
    lin = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
    st = fr.render(scene, cam, cfg)
    assert st.samples == W * H * spp and st.render_ms == 0.0  # host-known fields only
    fr.reduce_local(cfg)
    fr.finalize(cfg, rgb_out=rgb.numpy(), linear_out=lin.numpy())
    stream.synchronize()
    full = fr.stats()
    assert np.array_equal(rgb.numpy(), ref_rgb) and np.array_equal(lin.numpy(), ref_lin)
    assert full.rays == ref_st.rays and full.samples == ref_st.samples
    assert full.render_ms > 0 and full.frame_ms >= full.render_ms
    fr.set_stream(None)  # back to the frame's own stream: synchronous again
    st2 = fr.render(scene, cam, cfg)
    assert st2.render_ms > 0


def test_scene_update_in_place(rt):
    """rt_scene_update re-uses the allocation: the image equals a freshly created scene's, also
    with fewer spheres than the capacity, also in BVH mode (hierarchy rebuilt), and a scene that
    does not fit is refused."""
    W, H, spp = 128, 72, 16
    cover = golden("cover_small.npz")["spheres"]
    cam, cfg = cam_for(rt, "cover", W, H), rt.Config(W, H, spp, 50, seed=2)
    moved = cover.copy()
    moved[1:, 0] += 0.25
    fewer = cover[:200]
    sc = rt.Scene(cover)
    base = rt.render_image(sc, cam, cfg)[1]
    for rows in (moved, fewer, cover):
        sc.update(rows)
        for flags in (0, rt.FLAG_BVH):
            c = rt.Config(W, H, spp, 50, seed=2, flags=flags)
            assert np.array_equal(rt.render_image(sc, cam, c)[1], rt.render_image(rt.Scene(rows), cam, c)[1])
    assert np.array_equal(rt.render_image(sc, cam, cfg)[1], base)
    with pytest.raises(rt.RtError, match="capacity"):
        sc.update(np.concatenate([cover, cover]))


def test_frame_cache_of_render_image(rt):
    """Two sizes alternate through the one cached frame per thread; results do not depend on what
    the cached frame rendered before; release is idempotent."""
    rows = golden("three_small.npz")["spheres"]
    sc = rt.Scene(rows)
    out = {}
    for W, H in ((96, 54), (64, 36), (96, 54), (64, 36)):
        lin = rt.render_image(sc, cam_for(rt, "three", W, H), rt.Config(W, H, 8, 50, seed=1))[1]
        assert (W, H) not in out or np.array_equal(out[(W, H)], lin)
        out[(W, H)] = lin
    rt.release_cached_frames()
    rt.release_cached_frames()
    lin = rt.render_image(sc, cam_for(rt, "three", 96, 54), rt.Config(96, 54, 8, 50, seed=1))[1]
    assert np.array_equal(out[(96, 54)], lin)


def test_multi_object_on_one_device(rt):
    """rt_multi with a single device is the single-GPU image; it renders several frames and
    scenes without re-creating anything."""
    W, H, spp = 128, 72, 16
    cover, three = golden("cover_small.npz")["spheres"], golden("three_small.npz")["spheres"]
    m = rt.Multi([0], W, H)
    for rows, name in ((cover, "cover"), (three, "three"), (cover, "cover")):
        m.set_scene(rows)
        cam, cfg = cam_for(rt, name, W, H), rt.Config(W, H, spp, 50, seed=4)
        rgb, lin, st = m.render(cam, cfg, want_linear=True)
        ref = rt.render_image(rt.Scene(rows), cam, cfg)
        assert np.array_equal(rgb, ref[0]) and np.array_equal(lin, ref[1])
        assert st.rays == ref[2].rays and st.samples == ref[2].samples and st.render_ms > 0
    with pytest.raises(rt.RtError, match="does not match"):
        m.render(cam_for(rt, "cover", 64, 36), rt.Config(64, 36, 4))
    m.close()


@pytest.mark.parametrize("W,H", [(160, 90), (101, 37)])
def test_banded_finalise_and_pnm_writers(rt, tmp_path, W, H):
    """rt_finalize_to_file (bands of 64 rows, copy of band k+1 under the formatting of band k)
    writes the same bytes as rt_write_ppm of the whole image; P6 holds the raw rows; the band
    callback sees every row once, top to bottom."""
    rows = golden("three_small.npz")["spheres"]
    sc, cam, cfg = rt.Scene(rows), cam_for(rt, "three", W, H), rt.Config(W, H, 8, 50, seed=6)
    fr = rt.Frame(W, H)
    fr.render(sc, cam, cfg)
    fr.reduce_local(cfg)
    rgb, _, st = fr.finalize(cfg)
    p3, p3b, p6 = tmp_path / "a.ppm", tmp_path / "b.ppm", tmp_path / "c.ppm"
    rt.write_ppm(str(p3), rgb)
    st2 = fr.finalize_to_file(cfg, str(p3b))
    assert p3.read_bytes() == p3b.read_bytes()
    assert st2.rays == st.rays
    fr.finalize_to_file(cfg, str(p6), rt.PNM_P6)
    raw = p6.read_bytes()
    head = f"P6\n{W} {H}\n255\n".encode()
    assert raw.startswith(head) and raw[len(head):] == rgb.tobytes()
    seen = []
    fr.finalize_bands(cfg, 16, lambda band, r0, r1: seen.append((r0, r1, band)))
    assert [s[0] for s in seen] == list(range(0, H, 16)) and seen[-1][1] == H
    assert np.array_equal(np.concatenate([s[2] for s in seen]), rgb)
    with pytest.raises(rt.RtError):
        fr.finalize_to_file(cfg, str(tmp_path / "no" / "dir.ppm"))


def test_empty_world_with_bvh_flag(rt):
    """an empty world has no hierarchy to walk; the BVH flag must fall back to
    the brute-force kernel instead of traversing a degenerate root."""
    W, H, spp = 64, 36, 4
    sc = rt.Scene(np.zeros((0, 9)))
    cam = cam_for(rt, "cover", W, H)
    a = rt.render_image(sc, cam, rt.Config(W, H, spp, 50, seed=1))
    b = rt.render_image(sc, cam, rt.Config(W, H, spp, 50, seed=1, flags=rt.FLAG_BVH))
    assert np.array_equal(a[1], b[1]) and b[2].rays == W * H * spp


def test_flag_combinations_are_checked(rt):
    rows = golden("three_small.npz")["spheres"]
    sc, cam = rt.Scene(rows), cam_for(rt, "three", 32, 18)
    for flags in (rt.FLAG_NO_FILTER | rt.FLAG_BVH, rt.FLAG_NO_FILTER | rt.FLAG_F32_HIT,
                  rt.FLAG_F32_HIT | rt.FLAG_BVH, 4, 1 << 20):
        with pytest.raises(rt.RtError):
            rt.render_image(sc, cam, rt.Config(32, 18, 2, 50, flags=flags))
