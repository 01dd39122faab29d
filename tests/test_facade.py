"""The C++ façade (include/rt/rtweekend.hpp: the reference's class names over the C ABI) and the
example driver shaped like the reference's main() (examples/cover.cc)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, cam_for, golden

EX = os.path.join(ROOT, "examples")
BIN = os.path.join(EX, "rt_cover")


@pytest.fixture(scope="module")
def rt_cover(rt):
    rt.lib()
    subprocess.check_call(["make", "-C", EX, "rt_cover"], stdout=subprocess.DEVNULL)
    return BIN


@pytest.mark.parametrize("name", ["cover", "three"])
def test_facade_scene_equals_reference_dump(rt_cover, name, tmp_path):
    """random_scene() / three_sphere_scene() built through hittable_list::add(make_shared<sphere>)
    flatten to exactly the scene the reference's own random_scene() produced (golden dump)."""
    p = tmp_path / "scene.f64"
    subprocess.check_call([rt_cover, "--scene", name, "--dump-scene", str(p)])
    rows = np.fromfile(p, np.float64).reshape(-1, 9)
    assert np.array_equal(rows, golden(f"{name}_small.npz")["spheres"])


def test_facade_fails_loudly_without_gpu(rt_cover):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([rt_cover, "--width", "32", "--spp", "1"], capture_output=True, text=True)
    assert r.returncode == 99 and "rt_b200 error -2" in r.stderr and r.stdout == ""


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cover", "three"])
def test_example_ppm_equals_binding_render(rt, rt_cover, name, tmp_path):
    """Same kernels, same seed -> the C++ driver's P3 text is byte-identical to the image the
    ctypes binding renders, and has the reference's header/row format."""
    W, H, spp = 96, 54, 8
    out = tmp_path / "img.ppm"
    r = subprocess.run([rt_cover, "--scene", name, "--width", str(W), "--aspect", "16:9", "--spp",
                        str(spp), "--seed", "5", "--out", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Time Cost (ms)" in r.stderr
    tok = out.read_text().split()
    assert tok[:4] == ["P3", str(W), str(H), "255"]
    img = np.array(tok[4:], np.uint8).reshape(H, W, 3)
    scene = rt.Scene(golden(f"{name}_small.npz")["spheres"])
    rgb, _, _ = rt.render_image(scene, cam_for(rt, name, W, H), rt.Config(W, H, spp, 50, seed=5))
    # the driver's aspect is 16/9 exactly, the binding's W/H = 96/54 = 16/9 as well
    assert np.array_equal(img, rgb)
