"""The C++ façade (include/rt/rtweekend.hpp: the reference's class names over the C ABI) and the
example driver shaped like the reference's main() (examples/cover.cc)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, cam_for, golden

EX = os.path.join(ROOT, "examples")
BIN = os.path.join(EX, "rt_cover")
GPU_BIN = os.path.join(EX, "rt_gpu_cover")


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


# ------------------------------------------------------------------ rt::gpu facade (src/gpu API)
@pytest.fixture(scope="module")
def rt_gpu_cover(rt):
    rt.lib()
    subprocess.check_call(["make", "-C", EX, "rt_gpu_cover"], stdout=subprocess.DEVNULL)
    return GPU_BIN


def gpu_world(binary, seed, tmp_path):
    p = tmp_path / f"world{seed}.f64"
    subprocess.check_call([binary, "--seed", str(seed), "--dump-scene", str(p)])
    return np.fromfile(p, np.float64).reshape(-1, 9)


def test_gpu_facade_new_world(rt_gpu_cover, tmp_path):
    """new_world(world, materials, seed) follows src/gpu/main.cu:18-75: ground first, the three
    big spheres last, a 22 x 22 grid of r = 0.2 spheres kept when farther than 0.9 from
    (4, 0.2, 0), ~80 % lambertian / 15 % metal (fuzz < 0.5, albedo in [0.5, 1)) / 5 % glass 1.5."""
    rows = gpu_world(rt_gpu_cover, 7, tmp_path)
    assert np.array_equal(rows, gpu_world(rt_gpu_cover, 7, tmp_path))
    assert not np.array_equal(rows[:50], gpu_world(rt_gpu_cover, 8, tmp_path)[:50])
    assert 470 <= len(rows) <= 488
    assert rows[0].tolist() == [0, -1000, 0, 1000, 0, 0.5, 0.5, 0.5, 0]
    assert rows[-3, :5].tolist() == [0, 1, 0, 1, 2] and rows[-3, 8] == 1.5
    assert rows[-2, :5].tolist() == [-4, 1, 0, 1, 0]
    assert rows[-1, :5].tolist() == [4, 1, 0, 1, 1] and rows[-1, 8] == 0
    small = rows[1:-3]
    assert np.all(small[:, 3] == np.float32(0.2)) and np.all(small[:, 1] == np.float32(0.2))
    assert np.all(np.hypot(small[:, 0] - 4.0, small[:, 2]) > 0.9)
    kinds = np.bincount(small[:, 4].astype(int), minlength=3) / len(small)
    assert 0.72 < kinds[0] < 0.88 and 0.09 < kinds[1] < 0.21 and 0.01 < kinds[2] < 0.10
    metal = small[small[:, 4] == 1]
    assert np.all((metal[:, 5:8] >= 0.5) & (metal[:, 5:8] < 1.0)) and np.all((metal[:, 8] >= 0) & (metal[:, 8] < 0.5))
    assert np.all(small[small[:, 4] == 2][:, 8] == 1.5)


def test_gpu_facade_fails_loudly_without_gpu(rt_gpu_cover):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([rt_gpu_cover, "--width", "32", "--spp", "1"], capture_output=True, text=True)
    # checkCudaErrors convention of the reference: message on stderr, exit(99) (cuda_utility.h:8-18)
    assert r.returncode == 99 and "rt_b200 error = -2" in r.stderr and r.stdout == ""


@pytest.mark.gpu
def test_gpu_facade_image_equals_binding_render(rt, rt_gpu_cover, tmp_path):
    """render(frame_buffer, cam, world, seed) + output_image == the binding's RT_FLAG_V4_MODEL
    render of the dumped world with the camera new_camera describes (defocus_angle 0.6 deg ->
    aperture 2 * 10 * tan(0.3 deg), aspect = W / H)."""
    W, spp, seed = 96, 8, 7
    H = int(W / (np.float32(16.0) / np.float32(9.0)))
    out = tmp_path / "img.ppm"
    r = subprocess.run([rt_gpu_cover, "--width", str(W), "--spp", str(spp), "--seed", str(seed), "--out", str(out)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert f"Image Size = {W} x {H}" in r.stderr and "Time Cost (ms)" in r.stderr
    tok = out.read_text().split()
    assert tok[:4] == ["P3", str(W), str(H), "255"]
    img = np.array(tok[4:], np.uint8).reshape(H, W, 3)
    rows = gpu_world(rt_gpu_cover, seed, tmp_path)
    cam = rt.Camera(aspect_ratio=W / H, aperture=2.0 * 10.0 * np.tan(np.radians(np.float32(0.6)) / 2.0), focus_dist=10.0)
    rgb, _, _ = rt.render_image(rt.Scene(rows), cam, rt.Config(W, H, spp, 50, seed=seed, flags=rt.FLAG_V4_MODEL))
    assert (img != rgb).mean() < 1e-3  # aperture differs in the last bits (f32 angle vs f64)
