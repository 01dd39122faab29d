"""Pin the CPU oracle (oracle/rt_oracle.c) to the reference's own compiled code.

Golden files were produced by oracle/_ref/ref_harness (reference TU included unmodified,
tests/golden/make_golden.py).  Everything here is bit-exact: integer ids, IEEE doubles.
"""
import os
import subprocess
import tempfile

import numpy as np
import pytest

from conftest import golden


def test_philox_known_answers(orc):
    # Random123 kat_vectors, philox4x32-10
    assert orc.philox([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    f = 0xffffffff
    assert orc.philox([f, f, f, f], [f, f]) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert orc.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_mt19937_known_answer(orc):
    # C++ standard [rand.predef]: the 10000th invocation of a default mt19937 is 4123659995.
    # random_double() consumes two draws; the 5000th call used draws 9999 and 10000:
    # value = (g9999 + g10000 * 2^32) / 2^64, so floor(value * 2^32) == g10000.
    ctx = orc.Ctx()
    x = 0.0
    for _ in range(5000):
        x = ctx.random_double()
    assert int(x * 4294967296.0) == 4123659995


@pytest.mark.parametrize("name", ["cover", "three"])
def test_scene_dump_bit_exact(orc, name):
    g = golden(f"{name}_small.npz")
    rows = orc.cover_scene(orc.Ctx()) if name == "cover" else orc.three_scene()
    assert rows.shape == g["spheres"].shape
    assert np.array_equal(rows, g["spheres"])
    if name == "cover":  # 486 spheres = 382 lambertian + 77 metal + 27 glass
        kinds = rows[:, 4].astype(int)
        assert len(rows) == 486 and (np.bincount(kinds) == [382, 77, 27]).all()


@pytest.mark.parametrize("name", ["cover", "three"])
def test_sequential_render_bit_exact(orc, name):
    """Same mt19937 stream, same op order -> identical f64 sums and identical 8-bit PPM."""
    g = golden(f"{name}_small.npz")
    W, H, spp = int(g["W"]), int(g["H"]), int(g["spp"])
    ctx = orc.Ctx()
    rows = orc.cover_scene(ctx) if name == "cover" else orc.three_scene()
    cam = orc.scene_camera(name, W, H)
    lin, samples, rays = orc.render(rows, cam, W, H, spp, mode=orc.RNG_MT19937, ctx=ctx)
    assert samples == W * H * spp and rays > samples
    assert np.array_equal(lin, g["linear"])
    assert np.array_equal(orc.write_color(lin, spp), g["ppm"])


@pytest.mark.parametrize("name", ["cover", "three"])
def test_primary_hits_bit_exact(orc, name):
    g = golden(f"{name}_small.npz")
    W, H = int(g["W"]), int(g["H"])
    ids, t, nrm, ff = orc.primary_hits(g["spheres"], orc.scene_camera(name, W, H), W, H)
    assert np.array_equal(ids, g["ids"]) and np.array_equal(ff, g["front_face"])
    assert np.array_equal(t, g["t"]) and np.array_equal(nrm, g["normal"])
    g2 = golden(f"{name}_primary_200x112.npz")
    W, H = int(g2["W"]), int(g2["H"])
    ids, t, nrm, ff = orc.primary_hits(g["spheres"], orc.scene_camera(name, W, H), W, H)
    assert np.array_equal(ids, g2["ids"]) and np.array_equal(ff, g2["front_face"])
    assert np.array_equal(t, g2["t"]) and np.array_equal(nrm, g2["normal"])
    assert (ids >= 0).mean() > 0.3  # the table is not trivially empty


def test_write_color_analytic(orc):
    # color.h:14-22: 0 -> 0; spp (i.e. mean 1.0) -> int(256*0.999) = 255; mean .25 -> 128
    spp = 7
    sums = np.array([[0.0, 1.0 * spp, 0.25 * spp]])
    assert orc.write_color(sums, spp).tolist() == [[0, 255, 128]]


@pytest.mark.skipif(not os.path.exists("/root/reference/src/cpu/main.cc"),
                    reason="live reference only exists in the build container")
def test_live_reference_other_size(orc):
    """A size that is not in the golden set, straight against the reference binary."""
    subprocess.check_call(["make", "-C", os.path.join(os.path.dirname(orc.LIB)), "ref"],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    W, H, spp = 33, 21, 3
    with tempfile.TemporaryDirectory() as tmp:
        lin_p = os.path.join(tmp, "lin")
        subprocess.check_call([orc.REF_HARNESS, "--scene", "cover", "--w", str(W), "--h", str(H),
                               "--spp", str(spp), "--linear", lin_p])
        ref = np.fromfile(lin_p, np.float64).reshape(H, W, 3)
    ctx = orc.Ctx()
    rows = orc.cover_scene(ctx)
    lin, _, _ = orc.render(rows, orc.scene_camera("cover", W, H), W, H, spp,
                           mode=orc.RNG_MT19937, ctx=ctx)
    assert np.array_equal(lin, ref)


@pytest.mark.skipif(not os.path.exists("/root/reference/src/cpu-multi-threading/main.cc"),
                    reason="live reference only exists in the build container")
def test_mt_program_on_one_thread_is_the_oracle(orc):
    """The CPU baseline bench.py also times is the reference's cpu_multi_threading_ray_tracer, built
    by oracle/Makefile with its three workload literals read from the environment.  On ONE thread
    it visits pixels and samples in src/cpu/main.cc's order, so its P3 output must be the oracle's
    image byte for byte -- the generated copy changes the workload and nothing else."""
    odir = os.path.dirname(orc.LIB)
    subprocess.check_call(["make", "-C", odir, "mt_ref"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    W, spp = 48, 2
    H = int(W / (16.0 / 9.0))
    env = dict(os.environ, RT_MT_WIDTH=str(W), RT_MT_SPP=str(spp), RT_MT_THREADS="1")
    r = subprocess.run([os.path.join(odir, "_ref", "cpu_mt_tracer")], env=env, capture_output=True,
                       text=True, check=True)
    tok = r.stdout.split()
    assert tok[:4] == ["P3", str(W), str(H), "255"]
    img = np.array(tok[4:], dtype=np.int64).reshape(H, W, 3)
    assert any(l.startswith("Time Cost (ms)") for l in r.stderr.splitlines())
    ctx = orc.Ctx()
    rows = orc.cover_scene(ctx)
    lin, _, _ = orc.render(rows, orc.scene_camera("cover", W, H), W, H, spp, mode=orc.RNG_MT19937, ctx=ctx)
    assert np.array_equal(img, orc.write_color(lin.reshape(-1, 3), spp).reshape(H, W, 3))
