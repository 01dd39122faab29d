"""ctypes wrapper of oracle/liboracle.so (the CPU parity oracle).  TEST INFRASTRUCTURE:
imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")
REF_HARNESS = os.path.join(ORACLE_DIR, "_ref", "ref_harness")

RNG_MT19937, RNG_PHILOX = 0, 1


class OrcCamera(C.Structure):
    _fields_ = [("origin", C.c_double * 3), ("lower_left", C.c_double * 3),
                ("horizontal", C.c_double * 3), ("vertical", C.c_double * 3),
                ("u", C.c_double * 3), ("v", C.c_double * 3), ("w", C.c_double * 3),
                ("lens_radius", C.c_double)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"])
        L = C.CDLL(LIB)
        L.orc_create.restype = C.c_void_p
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_random_double.restype = C.c_double
        L.orc_random_double.argtypes = [C.c_void_p]
        L.orc_camera_init.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double,
                                      C.c_double, C.c_double, C.c_double]
        L.orc_cover_scene.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.orc_three_scene.argtypes = [C.c_void_p, C.c_int]
        L.orc_primary_hits.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_primary_hits_rows.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                            C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_render.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                 C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_uint32, C.c_int,
                                 C.c_int, C.c_void_p, C.c_void_p]
        L.orc_write_color.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_closest_hit.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_double,
                                      C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


class Ctx:
    """mt19937 stream shared by scene build and rendering, like the reference."""

    def __init__(self):
        self.h = C.c_void_p(lib().orc_create())

    def random_double(self) -> float:
        return lib().orc_random_double(self.h)

    def __del__(self):
        try:
            lib().orc_destroy(self.h)
        except Exception:
            pass


def camera(lookfrom, lookat, vup, vfov, aspect, aperture, focus) -> OrcCamera:
    cam = OrcCamera()
    d3 = lambda a: (C.c_double * 3)(*a)
    lib().orc_camera_init(C.byref(cam), d3(lookfrom), d3(lookat), d3(vup), vfov, aspect, aperture,
                          focus)
    return cam


def cover_scene(ctx: Ctx) -> np.ndarray:
    sp = np.zeros((1024, 9))
    n = lib().orc_cover_scene(ctx.h, sp.ctypes.data, 1024)
    assert n > 0
    return sp[:n].copy()


def three_scene() -> np.ndarray:
    sp = np.zeros((16, 9))
    n = lib().orc_three_scene(sp.ctypes.data, 16)
    return sp[:n].copy()


COVER_CAM = dict(lookfrom=(13.0, 2.0, 3.0), lookat=(0.0, 0.0, 0.0), vup=(0.0, 1.0, 0.0), vfov=20.0,
                 aperture=0.1, focus=10.0)                      # src/cpu/main.cc:93-99
THREE_CAM = dict(lookfrom=(3.0, 3.0, 2.0), lookat=(0.0, 0.0, -1.0), vup=(0.0, 1.0, 0.0), vfov=20.0,
                 aperture=2.0, focus=float(np.sqrt(27.0)))       # archive/listing69/main.cc:64-71


def scene_camera(name: str, W: int, H: int) -> OrcCamera:
    p = COVER_CAM if name == "cover" else THREE_CAM
    return camera(p["lookfrom"], p["lookat"], p["vup"], p["vfov"], W / H, p["aperture"], p["focus"])


def primary_hits(spheres: np.ndarray, cam: OrcCamera, W: int, H: int):
    spheres = np.ascontiguousarray(spheres, np.float64)
    ids = np.zeros((H, W), np.int32)
    t = np.zeros((H, W))
    nrm = np.zeros((H, W, 3))
    ff = np.zeros((H, W), np.uint8)
    lib().orc_primary_hits(spheres.ctypes.data, len(spheres), C.byref(cam), W, H, ids.ctypes.data,
                           t.ctypes.data, nrm.ctypes.data, ff.ctypes.data)
    return ids, t, nrm, ff


def _bands(H: int, workers: int):
    edges = np.linspace(0, H, min(H, workers * 4) + 1).astype(int)
    return [(int(a), int(b)) for a, b in zip(edges[:-1], edges[1:]) if b > a]


def primary_hits_parallel(spheres: np.ndarray, cam: OrcCamera, W: int, H: int, workers: int | None = None):
    """primary_hits over row bands on a thread pool (ctypes releases the GIL): full-size tables
    (1200x675, 3840x2160) in seconds instead of minutes."""
    from concurrent.futures import ThreadPoolExecutor
    spheres = np.ascontiguousarray(spheres, np.float64)
    workers = workers or os.cpu_count() or 1
    ids = np.zeros((H, W), np.int32)
    t = np.zeros((H, W))
    nrm = np.zeros((H, W, 3))
    ff = np.zeros((H, W), np.uint8)

    def one(band):
        a, b = band
        lib().orc_primary_hits_rows(spheres.ctypes.data, len(spheres), C.byref(cam), W, H, a, b,
                                    ids[a:b].ctypes.data, t[a:b].ctypes.data, nrm[a:b].ctypes.data,
                                    ff[a:b].ctypes.data)
    with ThreadPoolExecutor(workers) as ex:
        list(ex.map(one, _bands(H, workers)))
    return ids, t, nrm, ff


def render_parallel(spheres: np.ndarray, cam: OrcCamera, W: int, H: int, spp: int, depth: int = 50,
                    seed: int = 0, sample_offset: int = 0, model: int = 0, workers: int | None = None):
    """render(mode=RNG_PHILOX) over row bands on a thread pool -> (linear sums [H][W][3], samples,
    rays).  The Philox stream is keyed by (pixel, sample), so bands are independent."""
    from concurrent.futures import ThreadPoolExecutor
    spheres = np.ascontiguousarray(spheres, np.float64)
    workers = workers or os.cpu_count() or 1
    lib().orc_set_model(model)
    lin = np.zeros((H, W, 3))
    tot = [0, 0]

    def one(band):
        a, b = band
        st = (C.c_uint64 * 2)()
        lib().orc_render(None, spheres.ctypes.data, len(spheres), C.byref(cam), W, H, spp, depth,
                         RNG_PHILOX, seed, sample_offset, a, b, lin[a:b].ctypes.data, st)
        return int(st[0]), int(st[1])
    with ThreadPoolExecutor(workers) as ex:
        for s, r in ex.map(one, _bands(H, workers)):
            tot[0] += s; tot[1] += r
    lib().orc_set_model(MODEL_CPU)
    return lin, tot[0], tot[1]


MODEL_CPU, MODEL_GPU = 0, 1  # src/cpu (spec) | src/gpu (v4 book) semantics


def render(spheres: np.ndarray, cam: OrcCamera, W: int, H: int, spp: int, depth: int = 50,
           mode: int = RNG_PHILOX, seed: int = 0, sample_offset: int = 0, ctx: Ctx | None = None,
           rows=None, model: int = MODEL_CPU):
    """-> (linear f64 sums [rows][W][3], samples, rays)"""
    spheres = np.ascontiguousarray(spheres, np.float64)
    lib().orc_set_model(model)
    r0, r1 = rows if rows is not None else (0, H)
    lin = np.zeros((r1 - r0, W, 3))
    st = (C.c_uint64 * 2)()
    lib().orc_render(ctx.h if ctx is not None else None, spheres.ctypes.data, len(spheres),
                     C.byref(cam), W, H, spp, depth, mode, seed, sample_offset, r0, r1,
                     lin.ctypes.data, st)
    lib().orc_set_model(MODEL_CPU)
    return lin, int(st[0]), int(st[1])


def write_color(lin_sums: np.ndarray, spp: int) -> np.ndarray:
    """color.h:8-23 applied per pixel -> uint8 [..., 3]"""
    flat = np.ascontiguousarray(lin_sums, np.float64).reshape(-1, 3)
    out = np.zeros((len(flat), 3), np.int32)
    tmp = (C.c_int * 3)()
    for k in range(len(flat)):
        lib().orc_write_color(flat[k].ctypes.data, spp, tmp)
        out[k] = tmp[:]
    return out.reshape(lin_sums.shape).astype(np.uint8)


def philox(ctr, key):
    c = (C.c_uint32 * 4)(*ctr)
    k = (C.c_uint32 * 2)(*key)
    o = (C.c_uint32 * 4)()
    lib().orc_philox4x32_10(c, k, o)
    return list(o)


def closest_hit(spheres: np.ndarray, org, dirv, t_min=0.001, t_max=float("inf"), model: int = MODEL_CPU):
    """hittable_list::hit over the rows in insertion order -> (sphere index or -1, t)"""
    spheres = np.ascontiguousarray(spheres, np.float64)
    t = C.c_double(0.0)
    d3 = lambda a: (C.c_double * 3)(*a)
    lib().orc_set_model(model)
    idx = lib().orc_closest_hit(spheres.ctypes.data, len(spheres), d3(org), d3(dirv), t_min, t_max,
                                C.byref(t), None, None, None)
    lib().orc_set_model(MODEL_CPU)
    return idx, t.value
