"""Book-step scene pack (include/rt/book_scenes.hpp, gpu_book_scenes.hpp):
the scenes, cameras and — for the diffuse-only steps — material RULES of the reference's
archive/listing34…70 through the rt:: facade and of archive-gpu/image7…23 through rt::gpu.

CPU: every step's world and camera equal what tests/golden/make_book_golden.py parsed out of
the reference's own listings (tests/golden/book_steps.json); the last two steps equal the
scene dumps of the reference's compiled code (cover/three goldens).
GPU: every step renders through the CUDA path and matches the f64 oracle on the same Philox
stream — negative-radius bubbles, touching spheres, pinhole and 90-degree cameras included."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, golden

EX = os.path.join(ROOT, "examples")
BIN = os.path.join(EX, "rt_book")
STEPS = json.load(open(os.path.join(GOLDEN, "book_steps.json")))
ALL = list(STEPS) + ["defocus_blur", "final_cover", "gpu_final_cover"]


@pytest.fixture(scope="module")
def rt_book(rt):
    rt.lib()
    subprocess.check_call(["make", "-C", EX, "rt_book"], stdout=subprocess.DEVNULL)
    return BIN


def dump(binary, name, tmp_path):
    p = tmp_path / f"{name}.f64"
    subprocess.check_call([binary, name, "--dump", str(p)])
    a = np.fromfile(p, np.float64)
    rows, cam = a[:-12].reshape(-1, 9), a[-12:]
    return rows, dict(lookfrom=cam[0:3].tolist(), lookat=cam[3:6].tolist(), vfov=cam[6], aspect=cam[7],
                      aperture=cam[8], focus_dist=cam[9], flags=int(cam[10]), max_depth=int(cam[11]))


def test_list_names_every_step(rt_book):
    out = subprocess.check_output([rt_book, "--list"], text=True).splitlines()
    names = [l.split("\t")[0] for l in out]
    assert names[:14] == ["diffuse_first", "diffuse_gamma", "diffuse_lambertian", "diffuse_hemisphere",
                          "shiny_metal", "fuzzed_metal", "glass_pair", "glass_single",
                          "hollow_glass", "wide_angle", "distant_view", "zoomed_view",
                          "defocus_blur", "final_cover"]
    assert set(STEPS) <= set(names) and names[-1] == "gpu_final_cover" and len(names) == 28


@pytest.mark.parametrize("name", list(STEPS))
def test_step_equals_reference_listing(rt_book, name, tmp_path):
    rows, cam = dump(rt_book, name, tmp_path)
    g = STEPS[name]
    assert np.array_equal(rows, np.array(g["rows"], np.float64)), g["listing"]
    flags, depth = cam.pop("flags"), cam.pop("max_depth")
    assert cam == g["camera"]
    assert flags == g["flags"] and depth == g["max_depth"]
    line = [l for l in subprocess.check_output([rt_book, "--list"], text=True).splitlines() if l.startswith(name + "\t")][0]
    assert g["listing"] in line
    assert f"\t{g['image_width']} px\t{g['samples_per_pixel']} spp\tdepth {g['max_depth']}" in line


def test_last_steps_equal_reference_dumps(rt_book, tmp_path):
    rows, cam = dump(rt_book, "defocus_blur", tmp_path)
    assert np.array_equal(rows, golden("three_small.npz")["spheres"])
    assert cam["lookfrom"] == [3, 3, 2] and cam["aperture"] == 2.0 and cam["focus_dist"] == np.sqrt(27.0)
    assert cam["flags"] == 0 and cam["max_depth"] == 50
    rows, cam = dump(rt_book, "final_cover", tmp_path)
    assert np.array_equal(rows, golden("cover_small.npz")["spheres"])
    assert cam["lookfrom"] == [13, 2, 3] and (cam["aperture"], cam["focus_dist"], cam["aspect"]) == (0.1, 10.0, 1.5)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ALL)
def test_step_renders_like_the_oracle(rt, orc, rt_book, name, tmp_path):
    rows, c = dump(rt_book, name, tmp_path)
    W, H, spp = (64, 36, 8) if name == "final_cover" else (96, 54, 16)
    cam = rt.Camera(lookfrom=tuple(c["lookfrom"]), lookat=tuple(c["lookat"]), vfov=c["vfov"], aspect_ratio=W / H,
                    aperture=c["aperture"], focus_dist=c["focus_dist"])
    ocam = orc.camera(c["lookfrom"], c["lookat"], (0.0, 1.0, 0.0), c["vfov"], W / H, c["aperture"], c["focus_dist"])
    flags, depth = c["flags"], c["max_depth"]
    v4 = bool(flags & rt.FLAG_V4_MODEL)
    rgb, lin, st = rt.render_image(rt.Scene(rows), cam, rt.Config(W, H, spp, depth, seed=9, flags=flags))
    olin, samples, rays = orc.render(rows, ocam, W, H, spp, depth=depth, seed=9,
                                     model=orc.MODEL_GPU if v4 else orc.MODEL_CPU)
    err = np.abs(lin / spp - olin / spp)
    assert st.samples == samples
    assert np.median(err) < 1e-6 and err.mean() < 1e-4, (np.median(err), err.mean())
    assert abs(st.rays - rays) / rays < 3e-3
    if flags & rt.FLAG_LINEAR_OUTPUT:  # write_color before the book adds gamma (archive/listing30/color.h)
        oq = (256 * np.clip(olin / spp, 0.0, 0.999)).astype(np.uint8)
    else:
        oq = orc.write_color(olin, spp)
    assert (np.abs(rgb.astype(int) - oq.astype(int)) > 1).mean() < 3e-3
    if name.startswith(("diffuse", "gpu_hemisphere", "gpu_lambertian")):
        assert st.rays / st.samples > 1.5 and set(rows[:, 4]) <= {0.0, 3.0, 4.0}


@pytest.mark.gpu
def test_driver_writes_the_listing_sized_image(rt_book, tmp_path):
    out = tmp_path / "hollow.ppm"
    r = subprocess.run([rt_book, "hollow_glass", "--spp", "4", "--out", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert out.read_text().split()[:4] == ["P3", "400", "225", "255"]  # archive/listing61/main.cc:43-45
    out = tmp_path / "g.ppm"   # a step of the CUDA program's history, at its own 1920 x 1080
    r = subprocess.run([rt_book, "gpu_hemisphere_depth10", "--spp", "2", "--out", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert out.read_text().split()[:4] == ["P3", "1920", "1080", "255"]  # archive-gpu/image7/camera.h:43
