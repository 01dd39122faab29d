"""The C-ABI boundary on a machine WITHOUT a GPU: the library loads, exports exactly what
include/rt_b200.h declares, validates arguments, and fails loudly (never falls back to a CPU
path) when no CUDA device exists."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden

HEADER = os.path.join(ROOT, "include", "rt_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree(rt):
    assert declared_functions() == sorted(rt.SYMBOLS)


def test_library_exports_every_declared_symbol(rt):
    L = rt.lib()
    for name in declared_functions():
        assert hasattr(L, name), f"librt_b200.so does not export {name}"
    out = subprocess.run(["nm", "-D", "--defined-only", rt.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (rt_[a-z0-9_]+)", out))
    assert set(declared_functions()) <= exported
    assert L.rt_version().decode().endswith("(sm_100a)")


def test_struct_layouts_match_header(rt):
    """sizeof of the ctypes mirrors == what a C compiler makes of the header."""
    import tempfile
    prog = ('#include "rt_b200.h"\n#include <stdio.h>\nint main(void){printf("%zu %zu %zu %zu\\n",'
            "sizeof(rt_sphere),sizeof(rt_camera),sizeof(rt_config),sizeof(rt_stats));return 0;}\n")
    with tempfile.TemporaryDirectory() as tmp:
        c = os.path.join(tmp, "s.c")
        open(c, "w").write(prog)
        subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                               "-o", os.path.join(tmp, "s"), c])
        sizes = [int(x) for x in subprocess.check_output([os.path.join(tmp, "s")]).split()]
    assert sizes == [C.sizeof(rt._Sphere), C.sizeof(rt._Camera), C.sizeof(rt._Config), C.sizeof(rt._Stats)]


def test_argument_validation_needs_no_device(rt):
    L = rt.lib()
    h = C.c_void_p()
    assert L.rt_scene_create(None, C.c_uint32(3), C.c_int32(0), C.byref(h)) == -1  # RT_E_INVALID
    assert b"NULL" in L.rt_last_error()
    arr, n = rt.spheres_from_rows(golden("three_small.npz")["spheres"])
    arr[1].kind = 7
    assert L.rt_scene_create(arr, C.c_uint32(n), C.c_int32(0), C.byref(h)) == -1
    assert b"material kind" in L.rt_last_error()
    arr[1].kind = 0
    arr[2].radius = 0.0
    assert L.rt_scene_create(arr, C.c_uint32(n), C.c_int32(0), C.byref(h)) == -1
    assert L.rt_frame_create(C.c_uint32(1), C.c_uint32(1), C.c_int32(0), C.byref(h)) == -1


def test_no_cpu_fallback(rt):
    """Without a CUDA device every compute entry point returns RT_E_CUDA."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(rt.RtError, match="error -2"):
        rt.Scene(golden("three_small.npz")["spheres"])
    with pytest.raises(rt.RtError, match="error -2"):
        rt.Frame(16, 16)


def test_write_ppm_matches_reference_format(rt, tmp_path):
    """P3 header + 'r g b\\n' per pixel, rows top->bottom (src/cpu/main.cc:109, color.h:20-22):
    the golden 8-bit values written through rt_write_ppm reproduce the reference's own text."""
    g = golden("three_small.npz")
    ppm = g["ppm"].astype(np.uint8)
    H, W, _ = ppm.shape
    p = tmp_path / "o.ppm"
    rt.write_ppm(str(p), ppm)
    expect = f"P3\n{W} {H}\n255\n" + "".join(f"{r} {g_} {b}\n" for r, g_, b in ppm.reshape(-1, 3))
    assert p.read_text() == expect
