"""rt_b200 — ctypes binding of the C ABI in include/rt_b200.h (librt_b200.so).

This is test/bench plumbing around the product library, shaped like the reference's own
vocabulary (spheres with lambertian/metal/dielectric materials, a camera built from
lookfrom/lookat/vup/vfov/aspect/aperture/focus_dist, src/cpu/main.cc:93-99).  There is no
CPU fallback: if librt_b200.so is missing, importing the binding raises.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# RT_B200_LIB selects an experimental build of the same library (kernel A/B runs)
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "librt_b200.so")

LAMBERTIAN, METAL, DIELECTRIC = 0, 1, 2
DIFFUSE_BALL, DIFFUSE_HEMI = 3, 4  # the book's earlier diffuse rules (archive/listing33, listing40)
SLICES = 64
FLAG_COUNT_RAYS = 1
FLAG_BVH = 8
FLAG_V4_MODEL = 16  # src/gpu (v4 book) semantics, see include/rt_b200.h
FLAG_F32_HIT = 32   # f32 sphere::hit like src/gpu/sphere.h (compatibility/diagnostic)
FLAG_NO_FILTER = 64  # soundness switch: no f32 pre-test, every ray resolves every sphere in f64
FLAG_LINEAR_OUTPUT = 128  # write_color without the gamma sqrt (the book's first diffuse steps)
FLAG_SIMT_FILTER = 256  # FFMA2 filter loop instead of the tensor-core contraction (same image)
PNM_P3, PNM_P6 = 3, 6


class RtError(RuntimeError):
    pass


class _Sphere(C.Structure):
    _fields_ = [("center", C.c_double * 3), ("radius", C.c_double), ("kind", C.c_uint32),
                ("_pad", C.c_uint32), ("albedo", C.c_double * 3), ("fuzz_or_ir", C.c_double)]


class _Camera(C.Structure):
    _fields_ = [("lookfrom", C.c_double * 3), ("lookat", C.c_double * 3), ("vup", C.c_double * 3),
                ("vfov_deg", C.c_double), ("aspect_ratio", C.c_double), ("aperture", C.c_double),
                ("focus_dist", C.c_double)]


class _Config(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("samples_per_pixel", C.c_uint32),
                ("max_depth", C.c_uint32), ("seed", C.c_uint64), ("sample_offset", C.c_uint32),
                ("slice_begin", C.c_uint32), ("slice_end", C.c_uint32), ("device", C.c_int32),
                ("flags", C.c_uint32)]


class _Stats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("rays", C.c_uint64), ("sphere_tests", C.c_uint64),
                ("render_ms", C.c_float), ("finalize_ms", C.c_float), ("launches", C.c_uint32),
                ("frame_ms", C.c_float), ("filter_unit", C.c_uint32)]


# every symbol include/rt_b200.h declares
SYMBOLS = [
    "rt_last_error", "rt_version", "rt_device_count", "rt_scene_create", "rt_scene_destroy",
    "rt_scene_count", "rt_frame_create", "rt_frame_destroy", "rt_frame_partial_ptr", "rt_render",
    "rt_frame_reduce_local", "rt_finalize", "rt_render_image", "rt_primary_hits", "rt_write_ppm",
    "rt_frame_set_stream", "rt_measure_fp32_peak", "rt_render_image_multi",
    "rt_scene_update", "rt_frame_stats", "rt_release_cached_frames", "rt_write_pnm",
    "rt_finalize_bands", "rt_finalize_to_file", "rt_multi_create", "rt_multi_set_scene",
    "rt_multi_render", "rt_multi_destroy", "rt_tc_filter_probe", "rt_primary_hits_static",
]

BAND_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_uint8), C.c_uint32, C.c_uint32)

_lib = None


def lib() -> C.CDLL:
    """Load librt_b200.so (fails loudly when it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; "
                          "g.build()'` (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.rt_last_error.restype = C.c_char_p
        L.rt_version.restype = C.c_char_p
        for name in SYMBOLS[2:]:
            getattr(L, name).restype = C.c_int
        _lib = L
    return _lib


def _check(rc: int) -> None:
    if rc != 0:
        raise RtError(f"rt_b200 error {rc}: {lib().rt_last_error().decode()}")


@dataclass
class Camera:
    """camera(lookfrom, lookat, vup, vfov, aspect_ratio, aperture, focus_dist), camera.h:8-10"""
    lookfrom: tuple = (13.0, 2.0, 3.0)
    lookat: tuple = (0.0, 0.0, 0.0)
    vup: tuple = (0.0, 1.0, 0.0)
    vfov: float = 20.0
    aspect_ratio: float = 16.0 / 9.0
    aperture: float = 0.1
    focus_dist: float = 10.0

    def c(self) -> _Camera:
        return _Camera((C.c_double * 3)(*self.lookfrom), (C.c_double * 3)(*self.lookat),
                       (C.c_double * 3)(*self.vup), self.vfov, self.aspect_ratio, self.aperture,
                       self.focus_dist)


@dataclass
class Config:
    width: int
    height: int
    samples_per_pixel: int
    max_depth: int = 50
    seed: int = 0
    sample_offset: int = 0
    slice_begin: int = 0
    slice_end: int = SLICES
    device: int = 0
    flags: int = 0

    def c(self) -> _Config:
        return _Config(self.width, self.height, self.samples_per_pixel, self.max_depth, self.seed,
                       self.sample_offset, self.slice_begin, self.slice_end, self.device, self.flags)


@dataclass
class Stats:
    samples: int = 0
    rays: int = 0
    sphere_tests: int = 0
    render_ms: float = 0.0
    finalize_ms: float = 0.0
    launches: int = 0
    frame_ms: float = 0.0
    filter_unit: int = 0  # FILTER_NONE / FILTER_SIMT / FILTER_TENSOR

    @staticmethod
    def of(s: _Stats) -> "Stats":
        return Stats(s.samples, s.rays, s.sphere_tests, s.render_ms, s.finalize_ms, s.launches,
                     s.frame_ms, s.filter_unit)


def spheres_from_rows(rows: np.ndarray):
    """[n][9] f64 rows (cx cy cz radius kind r g b fuzz|ir — the oracle's scene dump layout,
    reference insertion order) -> ctypes rt_sphere array."""
    rows = np.ascontiguousarray(rows, dtype=np.float64).reshape(-1, 9)
    arr = (_Sphere * max(len(rows), 1))()
    for k, r in enumerate(rows):
        arr[k].center[:] = r[0:3]
        arr[k].radius = r[3]
        arr[k].kind = int(r[4])
        arr[k].albedo[:] = r[5:8]
        arr[k].fuzz_or_ir = r[8]
    return arr, len(rows)


class Scene:
    """Device-resident flattened hittable_list (src/cpu/hittable_list.h)."""

    def __init__(self, rows: np.ndarray, device: int = 0):
        arr, n = spheres_from_rows(rows)
        self.n, self.device = n, device
        self._h = C.c_void_p()
        _check(lib().rt_scene_create(arr, C.c_uint32(n), C.c_int32(device), C.byref(self._h)))

    def update(self, rows: np.ndarray, cuda_stream: int | None = None) -> None:
        """hittable_list::clear + add without re-allocating (rt_scene_update): one asynchronous
        host->device copy from the scene's pinned staging buffer on `cuda_stream`."""
        arr, n = spheres_from_rows(rows)
        _check(lib().rt_scene_update(self._h, arr, C.c_uint32(n), C.c_void_p(cuda_stream or 0)))
        self.n = n

    def update_raw(self, arr, n: int, cuda_stream: int | None = None) -> None:
        """rt_scene_update from a prepared ctypes rt_sphere array (no Python-side conversion)."""
        _check(lib().rt_scene_update(self._h, arr, C.c_uint32(n), C.c_void_p(cuda_stream or 0)))
        self.n = n

    def blob_bytes(self) -> int:
        """bytes one upload of this scene moves host->device (the four SoA row arrays)"""
        cap = max((self.n + 7) // 8 * 8, 8)
        return cap * 32 + (cap + 8) * 16 + cap * 16 + cap * 8

    def close(self):
        if self._h:
            lib().rt_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Frame:
    """Per-GPU accumulation buffers of one frame."""

    def __init__(self, width: int, height: int, device: int = 0):
        self.width, self.height, self.device = width, height, device
        self._h = C.c_void_p()
        _check(lib().rt_frame_create(C.c_uint32(width), C.c_uint32(height), C.c_int32(device),
                                     C.byref(self._h)))

    def set_stream(self, cuda_stream: int | None) -> None:
        _check(lib().rt_frame_set_stream(self._h, C.c_void_p(cuda_stream or 0)))

    def partial_ptr(self):
        p, nbytes = C.c_void_p(), C.c_size_t()
        _check(lib().rt_frame_partial_ptr(self._h, C.byref(p), C.byref(nbytes)))
        return p.value, nbytes.value

    def render(self, scene: Scene, cam: Camera, cfg: Config) -> Stats:
        st, cc, cf = _Stats(), cam.c(), cfg.c()
        _check(lib().rt_render(scene._h, C.byref(cc), C.byref(cf), self._h, C.byref(st)))
        return Stats.of(st)

    def reduce_local(self, cfg: Config) -> None:
        cf = cfg.c()
        _check(lib().rt_frame_reduce_local(self._h, C.byref(cf)))

    def stats(self) -> Stats:
        """Timings / ray count of the last render + finalise (after the stream was synchronised)."""
        st = _Stats()
        _check(lib().rt_frame_stats(self._h, C.byref(st)))
        return Stats.of(st)

    def finalize_to_file(self, cfg: Config, path: str, fmt: int = PNM_P3, partial_ptrs=None) -> Stats:
        """Finalise straight into a PNM file, band by band (rt_finalize_to_file)."""
        if partial_ptrs is None:
            partial_ptrs = [self.partial_ptr()[0]]
        n = len(partial_ptrs)
        ptrs = (C.c_void_p * n)(*partial_ptrs)
        st, cf = _Stats(), cfg.c()
        _check(lib().rt_finalize_to_file(self._h, C.byref(cf), ptrs, C.c_uint32(n), path.encode(),
                                         C.c_int(fmt), C.byref(st)))
        return Stats.of(st)

    def finalize_bands(self, cfg: Config, band_rows: int, fn, partial_ptrs=None) -> Stats:
        """fn(rows uint8 [r1-r0][W][3], r0, r1) is called per band of rows, top to bottom."""
        if partial_ptrs is None:
            partial_ptrs = [self.partial_ptr()[0]]
        n = len(partial_ptrs)
        ptrs = (C.c_void_p * n)(*partial_ptrs)
        W = self.width

        def tramp(_user, p, r0, r1):
            fn(np.ctypeslib.as_array(p, shape=(r1 - r0, W, 3)).copy(), r0, r1)
            return 0
        cb = BAND_FN(tramp)
        st, cf = _Stats(), cfg.c()
        _check(lib().rt_finalize_bands(self._h, C.byref(cf), ptrs, C.c_uint32(n), C.c_uint32(band_rows),
                                       cb, None, C.byref(st)))
        return Stats.of(st)

    def finalize(self, cfg: Config, partial_ptrs=None, want_linear=False, rgb_out=None,
                 linear_out=None):
        """-> (rgb8 [H][W][3] uint8, linear [H][W][3] float32 | None, Stats)"""
        if partial_ptrs is None:
            partial_ptrs = [self.partial_ptr()[0]]
        n = len(partial_ptrs)
        ptrs = (C.c_void_p * n)(*partial_ptrs)
        rgb = rgb_out if rgb_out is not None else np.empty((self.height, self.width, 3), np.uint8)
        lin = linear_out
        if lin is None and want_linear:
            lin = np.empty((self.height, self.width, 3), np.float32)
        st, cf = _Stats(), cfg.c()
        _check(lib().rt_finalize(self._h, C.byref(cf), ptrs, C.c_uint32(n),
                                 C.c_void_p(rgb.ctypes.data),
                                 C.c_void_p(lin.ctypes.data) if lin is not None else None,
                                 C.byref(st)))
        return rgb, lin, Stats.of(st)

    def close(self):
        if self._h:
            lib().rt_frame_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def render_image(scene: Scene, cam: Camera, cfg: Config, want_linear=True):
    """One-call render on one GPU -> (rgb8, linear sums, Stats)."""
    rgb = np.empty((cfg.height, cfg.width, 3), np.uint8)
    lin = np.empty((cfg.height, cfg.width, 3), np.float32) if want_linear else None
    st, cc, cf = _Stats(), cam.c(), cfg.c()
    _check(lib().rt_render_image(scene._h, C.byref(cc), C.byref(cf), C.c_void_p(rgb.ctypes.data),
                                 C.c_void_p(lin.ctypes.data) if lin is not None else None,
                                 C.byref(st)))
    return rgb, lin, Stats.of(st)


def render_image_multi(rows: np.ndarray, cam: Camera, cfg: Config, devices, want_linear=True):
    """One process, several GPUs (rt_render_image_multi) -> (rgb8, linear sums, Stats)."""
    arr, n = spheres_from_rows(rows)
    rgb = np.empty((cfg.height, cfg.width, 3), np.uint8)
    lin = np.empty((cfg.height, cfg.width, 3), np.float32) if want_linear else None
    st, cc, cf = _Stats(), cam.c(), cfg.c()
    devs = (C.c_int32 * len(devices))(*devices)
    _check(lib().rt_render_image_multi(arr, C.c_uint32(n), C.byref(cc), C.byref(cf), devs,
                                       C.c_uint32(len(devices)), C.c_void_p(rgb.ctypes.data),
                                       C.c_void_p(lin.ctypes.data) if lin is not None else None,
                                       C.byref(st)))
    return rgb, lin, Stats.of(st)


class Multi:
    """One process, several GPUs (rt_multi_*): scenes, frames and peer access set up once."""

    def __init__(self, devices, width: int, height: int):
        self.devices, self.width, self.height = list(devices), width, height
        self._h = C.c_void_p()
        devs = (C.c_int32 * len(self.devices))(*self.devices)
        _check(lib().rt_multi_create(devs, C.c_uint32(len(self.devices)), C.c_uint32(width),
                                     C.c_uint32(height), C.byref(self._h)))

    def set_scene(self, rows: np.ndarray) -> None:
        arr, n = spheres_from_rows(rows)
        _check(lib().rt_multi_set_scene(self._h, arr, C.c_uint32(n)))

    def render(self, cam: Camera, cfg: Config, want_linear=False, rgb_out=None):
        rgb = rgb_out if rgb_out is not None else np.empty((self.height, self.width, 3), np.uint8)
        lin = np.empty((self.height, self.width, 3), np.float32) if want_linear else None
        st, cc, cf = _Stats(), cam.c(), cfg.c()
        _check(lib().rt_multi_render(self._h, C.byref(cc), C.byref(cf), C.c_void_p(rgb.ctypes.data),
                                     C.c_void_p(lin.ctypes.data) if lin is not None else None,
                                     C.byref(st)))
        return rgb, lin, Stats.of(st)

    def close(self):
        if self._h:
            lib().rt_multi_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def tc_filter_probe(scene: Scene, ray_o: np.ndarray, ray_d: np.ndarray, n_values: int = 0, repeats: int = 1):
    """The filter as a tf32 contraction on the tensor cores (rt_tc_filter_probe).
    -> (pass mask bool [n][tiles * T], raw values f32 [n_values][tiles * T] | None, ms); T = spheres per tile"""
    ray_o = np.ascontiguousarray(ray_o, np.float32).reshape(-1, 4)
    ray_d = np.ascontiguousarray(ray_d, np.float32).reshape(-1, 4)
    n = len(ray_o)
    tiles_max = 32
    masks = np.zeros((n, tiles_max), np.uint64)
    vals = np.zeros((n_values, tiles_max * 64), np.float32) if n_values else None
    tiles, tsz, ms = C.c_uint32(), C.c_uint32(), C.c_float()
    _check(lib().rt_tc_filter_probe(scene._h, C.c_void_p(ray_o.ctypes.data), C.c_void_p(ray_d.ctypes.data),
                                    C.c_uint32(n), C.c_void_p(masks.ctypes.data), C.byref(tiles), C.byref(tsz),
                                    C.c_void_p(vals.ctypes.data) if vals is not None else None,
                                    C.c_uint32(n_values), C.c_uint32(repeats), C.byref(ms)))
    t, T = tiles.value, tsz.value
    m = masks.reshape(-1)[: n * t].reshape(n, t)
    bits = np.unpackbits(m.view(np.uint8).reshape(n, t, 8), axis=2, bitorder="little")[:, :, :T].reshape(n, t * T).astype(bool)
    v = vals.reshape(-1)[: n_values * t * T].reshape(n_values, t * T) if vals is not None else None
    return bits, v, ms.value


def release_cached_frames() -> None:
    _check(lib().rt_release_cached_frames())


def primary_hits(scene: Scene, cam: Camera, width: int, height: int, use_f64: bool, device=0):
    """-> ids int32 [H][W], t f64 [H][W], normal f64 [H][W][3], front_face u8 [H][W]"""
    ids = np.empty((height, width), np.int32)
    t = np.empty((height, width), np.float64)
    nrm = np.empty((height, width, 3), np.float64)
    ff = np.empty((height, width), np.uint8)
    cc = cam.c()
    _check(lib().rt_primary_hits(scene._h, C.byref(cc), C.c_uint32(width), C.c_uint32(height),
                                 C.c_int32(device), C.c_int(1 if use_f64 else 0),
                                 C.c_void_p(ids.ctypes.data), C.c_void_p(t.ctypes.data),
                                 C.c_void_p(nrm.ctypes.data), C.c_void_p(ff.ctypes.data)))
    return ids, t, nrm, ff


STATIC_MAX_PIXELS = 131072  # RT_STATIC_MAX_PIXELS
FILTER_NONE, FILTER_SIMT, FILTER_TENSOR = 0, 1, 2  # rt_stats.filter_unit


def primary_hits_static(scene: Scene, cam: Camera, width: int, height: int, use_f64: bool, device=0, stream=0):
    """primary_hits without any device allocation (rt_primary_hits_static: static scratch of
    STATIC_MAX_PIXELS pixels per GPU; larger frames raise).  No counterpart in the reference."""
    ids = np.empty((height, width), np.int32)
    t = np.empty((height, width), np.float64)
    nrm = np.empty((height, width, 3), np.float64)
    ff = np.empty((height, width), np.uint8)
    cc = cam.c()
    _check(lib().rt_primary_hits_static(scene._h, C.byref(cc), C.c_uint32(width), C.c_uint32(height),
                                        C.c_int32(device), C.c_int(1 if use_f64 else 0), C.c_void_p(stream),
                                        C.c_void_p(ids.ctypes.data), C.c_void_p(t.ctypes.data),
                                        C.c_void_p(nrm.ctypes.data), C.c_void_p(ff.ctypes.data)))
    return ids, t, nrm, ff


def measure_fp32_peak(device: int = 0):
    """-> (TFLOP/s of back-to-back FFMA, estimated SM clock in MHz)"""
    tf, mhz = C.c_float(), C.c_float()
    _check(lib().rt_measure_fp32_peak(C.c_int32(device), C.byref(tf), C.byref(mhz)))
    return tf.value, mhz.value


def write_ppm(path: str, rgb8: np.ndarray, fmt: int = PNM_P3) -> None:
    h, w, _ = rgb8.shape
    rgb8 = np.ascontiguousarray(rgb8, np.uint8)
    _check(lib().rt_write_pnm(path.encode(), C.c_void_p(rgb8.ctypes.data), C.c_uint32(w),
                              C.c_uint32(h), C.c_int(fmt)))
