"""rt_primary_hits_static: the primary-hit table without any device allocation (static __device__
scratch, include/rt_b200.h).  No counterpart in the reference.
The check is the existing entry point: rt_primary_hits on the same inputs must give the
same four arrays bit for bit (same kernels, only the staging memory differs) — on two calls in a row
with DIFFERENT frames, so that contents left in the shared array by the first call cannot pass for
results of the second.  Collected last on purpose (pytest -x)."""
import numpy as np
import pytest

from conftest import cam_for, golden

pytestmark = pytest.mark.gpu


def _same(a, b):
    for x, y, what in zip(a, b, ("ids", "t", "normal", "front_face")):
        assert x.dtype == y.dtype and x.shape == y.shape, what
        assert np.array_equal(x.view(np.uint8), y.view(np.uint8)), what  # bit-exact, NaN-safe


@pytest.mark.parametrize("use_f64", [True, False])
def test_static_matches_allocating_entry_point_twice(rt, use_f64):
    cover = rt.Scene(golden("cover_small.npz")["spheres"])
    three = rt.Scene(golden("three_small.npz")["spheres"])
    # call 1: the largest frame the static array takes (512 x 256 = RT_STATIC_MAX_PIXELS)
    W, H = 512, 256
    assert W * H == rt.STATIC_MAX_PIXELS
    got1 = rt.primary_hits_static(cover, cam_for(rt, "cover", W, H), W, H, use_f64)
    ref1 = rt.primary_hits(cover, cam_for(rt, "cover", W, H), W, H, use_f64)
    _same(got1, ref1)
    assert (ref1[0] >= 0).any() and (ref1[0] < 0).any()  # hits and sky both present
    # call 2, straight after: another scene, another (smaller, ragged) frame in the same array
    W, H = 203, 117
    got2 = rt.primary_hits_static(three, cam_for(rt, "three", W, H), W, H, use_f64)
    ref2 = rt.primary_hits(three, cam_for(rt, "three", W, H), W, H, use_f64)
    _same(got2, ref2)
    # and the first frame again: nothing of call 2 survives in it
    _same(rt.primary_hits_static(cover, cam_for(rt, "cover", 512, 256), 512, 256, use_f64), ref1)


def test_static_against_golden(rt):
    """the reference-generated table (tests/golden) through the no-allocation entry point"""
    g = golden("cover_primary_200x112.npz")
    W, H = int(g["W"]), int(g["H"])
    scene = rt.Scene(golden("cover_small.npz")["spheres"])
    ids, t, nrm, ff = rt.primary_hits_static(scene, cam_for(rt, "cover", W, H), W, H, True)
    assert np.array_equal(ids, g["ids"])
    assert np.array_equal(ff, g["front_face"])


def test_static_on_a_caller_stream(rt):
    torch = pytest.importorskip("torch")
    st = torch.cuda.Stream()
    W, H = 160, 90
    scene = rt.Scene(golden("three_small.npz")["spheres"])
    cam = cam_for(rt, "three", W, H)
    _same(rt.primary_hits_static(scene, cam, W, H, True, stream=st.cuda_stream), rt.primary_hits(scene, cam, W, H, True))


def test_static_rejects_frames_beyond_the_maximum(rt):
    scene = rt.Scene(golden("three_small.npz")["spheres"])
    W, H = 512, 257  # one row more than RT_STATIC_MAX_PIXELS
    with pytest.raises(rt.RtError, match="RT_STATIC_MAX_PIXELS"):
        rt.primary_hits_static(scene, cam_for(rt, "three", W, H), W, H, True)
    # the allocating entry point takes the same frame
    ids, *_ = rt.primary_hits(scene, cam_for(rt, "three", W, H), W, H, True)
    assert ids.shape == (H, W)
