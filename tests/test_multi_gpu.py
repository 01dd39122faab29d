"""One process, several GPUs through the C ABI (rt_render_image_multi): north-star gate 3 on real
devices — the N-GPU image is byte-identical to the 1-GPU image.  Needs >= 2 GPUs (skipped on the
single-GPU test box; run with `gpurun --gpus 2`)."""
import ctypes as C

import numpy as np
import pytest

from conftest import cam_for, golden

pytestmark = pytest.mark.gpu


def device_count(rt):
    n = C.c_int()
    assert rt.lib().rt_device_count(C.byref(n)) == 0
    return n.value


@pytest.mark.parametrize("name", ["cover", "three"])
def test_multi_gpu_image_is_byte_identical(rt, name):
    ndev = device_count(rt)
    if ndev < 2:
        pytest.skip("needs at least two GPUs")
    W, H, spp = 160, 90, 64
    rows = golden(f"{name}_small.npz")["spheres"]
    cam = cam_for(rt, name, W, H)
    cfg = rt.Config(W, H, spp, 50, seed=11)
    base_rgb, base_lin, base_st = rt.render_image(rt.Scene(rows), cam, cfg)
    for G in (2, 4, 8):
        if G > ndev:
            break
        rgb, lin, st = rt.render_image_multi(rows, cam, cfg, list(range(G)))
        assert st.samples == base_st.samples and st.rays == base_st.rays
        assert np.array_equal(rgb, base_rgb), f"G={G}: 8-bit image differs"
        assert np.array_equal(lin, base_lin), f"G={G}: linear sums differ"


def test_persistent_multi_object_is_byte_identical(rt):
    """rt_multi_*: scenes, frames and peer access set up once; device 0's finalise kernel reads the
    other GPUs' partials through peer pointers.  Several frames (different seeds, then another
    scene) through the same object equal the single-GPU images byte for byte."""
    ndev = device_count(rt)
    if ndev < 2:
        pytest.skip("needs at least two GPUs")
    W, H, spp = 160, 90, 64
    for G in (2, 4, 8):
        if G > ndev:
            break
        m = rt.Multi(list(range(G)), W, H)
        for name, seed in (("cover", 1), ("cover", 2), ("three", 3)):
            rows = golden(f"{name}_small.npz")["spheres"]
            cam, cfg = cam_for(rt, name, W, H), rt.Config(W, H, spp, 50, seed=seed)
            m.set_scene(rows)
            rgb, lin, st = m.render(cam, cfg, want_linear=True)
            base_rgb, base_lin, base_st = rt.render_image(rt.Scene(rows), cam, cfg)
            assert st.rays == base_st.rays and st.samples == base_st.samples
            assert np.array_equal(rgb, base_rgb) and np.array_equal(lin, base_lin), f"G={G} {name}"
        m.close()


def test_multi_gpu_argument_errors(rt):
    rows = golden("three_small.npz")["spheres"]
    cam = cam_for(rt, "three", 32, 18)
    with pytest.raises(rt.RtError, match="ngpu must be"):
        rt.render_image_multi(rows, cam, rt.Config(32, 18, 4), [0, 0, 0])
    with pytest.raises(rt.RtError, match="listed twice"):
        rt.render_image_multi(rows, cam, rt.Config(32, 18, 4), [0, 0])
