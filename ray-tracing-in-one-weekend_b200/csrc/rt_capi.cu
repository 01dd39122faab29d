// csrc/rt_capi.cu — the C ABI of include/rt_b200.h: host-side flattening of the reference's
// scene/camera objects, buffer management, kernel launches, error reporting.
// No CPU fallback: every compute entry point needs a CUDA device.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rt_b200.h"
#include "rt_kernels.h"

using namespace rtb;

static thread_local char g_err[512] = "";

static int fail(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}

#define CU(expr)                                                                         \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess)                                                              \
      return fail(RT_E_CUDA, "CUDA error %d (%s) at %s:%d '%s'", (int)e__,               \
                  cudaGetErrorString(e__), __FILE__, __LINE__, #expr);                   \
  } while (0)

// CU for the create functions: the half-built object is destroyed before the error is returned
#define CU_OR(cleanup, expr)                                                             \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess) {                                                            \
      cleanup;                                                                           \
      return fail(RT_E_CUDA, "CUDA error %d (%s) at %s:%d '%s'", (int)e__,               \
                  cudaGetErrorString(e__), __FILE__, __LINE__, #expr);                   \
    }                                                                                    \
  } while (0)

struct rt_scene {
  int device;
  uint32_t n;
  float4 *fast;
  double4 *exact;
  float4 *mat;
  float2 *aux;
  uint32_t n_pad;
  std::vector<float4> host_fast;  // host copy of the filter rows (kernel-parameter path)
  BvhNode *bvh_nodes;             // optional hierarchy, built on the host at scene creation
  uint32_t *bvh_prim;
  uint32_t bvh_node_count;
};

struct rt_frame {
  int device;
  uint32_t w, h, npix;
  float4 *accum;    // [RT_SLICES][npix]
  float4 *partial;  // [npix]
  uint32_t *rgb_words;
  float *linear;
  unsigned long long *counters;  // [0] work queue head, [1] ray total
  cudaStream_t stream;      // stream in use (own_stream unless rt_frame_set_stream)
  cudaStream_t own_stream;
  cudaEvent_t ev[4];
  int sm_count;
  size_t smem_optin;
  uint32_t last_n;
  void *pool;               // wavefront path-slot pool (allocated on first use)
  uint32_t pool_slots;
  unsigned int *live;       // device counter
  unsigned int *host_live;  // pinned
  // hand-out order of the next frame: pixels by descending cost of the frame rendered before
  uint32_t *order, *sort_keys, *sort_keys_out, *sort_iota, *heavy;
  void *sort_temp;
  size_t sort_temp_bytes;
  bool order_valid;
  bool order_heavy_only;    // the order was made for BVH frames (heavy pixels first, the rest in scan order)
};

extern "C" const char *rt_last_error(void) { return g_err; }
extern "C" const char *rt_version(void) { return "rt_b200 0.1 (sm_100a)"; }

extern "C" int rt_device_count(int *count) {
  if (!count) return fail(RT_E_INVALID, "count is NULL");
  CU(cudaGetDeviceCount(count));
  return RT_OK;
}

// ---------------------------------------------------------------- scene
extern "C" int rt_scene_create(const rt_sphere *sp, uint32_t n, int32_t device, rt_scene **out) {
  if (!out) return fail(RT_E_INVALID, "out is NULL");
  *out = nullptr;
  if (n && !sp) return fail(RT_E_INVALID, "spheres is NULL");
  if (n > (1u << 24)) return fail(RT_E_INVALID, "too many spheres (%u)", n);
  // rows are padded to a multiple of 8 (the filter loop takes two groups of 4 per trip); a padding row has
  // w = -inf so that b*b + q = -inf never reaches the threshold
  const uint32_t npad = n ? (n + kRowPad - 1) / kRowPad * kRowPad : kRowPad;
  std::vector<float4> fast(npad + kRowTail, make_float4(0.f, 0.f, 0.f, -INFINITY));  // + prefetch group
  std::vector<double4> exact(npad, make_double4(0, 0, 0, 1));
  std::vector<float4> mat(npad, make_float4(0.f, 0.f, 0.f, 0.f));
  std::vector<float2> aux(npad, make_float2(1.f, 0.f));
  for (uint32_t k = 0; k < n; ++k) {
    const rt_sphere &s = sp[k];
    if (s.kind > RT_DIELECTRIC) return fail(RT_E_INVALID, "sphere %u: bad material kind %u", k, s.kind);
    if (!(s.radius != 0.0) || !std::isfinite(s.radius))
      return fail(RT_E_INVALID, "sphere %u: radius must be finite and non-zero", k);
    double cx = s.center[0], cy = s.center[1], cz = s.center[2], r = s.radius;
    exact[k] = make_double4(cx, cy, cz, r);
    // Filter row: the line o + t d (|d| = 1) meets the sphere iff
    //   (c.d - o.d)^2 + 2 c.o + (r^2 - |c|^2) >= |o|^2.
    // The f32 evaluation error is bounded by SLACK*(|c|^2 + r^2) + SLACK*|o|^2;
    // the sphere part is folded into w here (rounded up), the ray part in the kernel.
    float fx = (float)cx, fy = (float)cy, fz = (float)cz;
    double c2 = cx * cx + cy * cy + cz * cz, r2 = r * r;
    double w = (r2 - c2) + (double)RTB_FILTER_SLACK * (c2 + r2);
    float wf = (float)w;
    if ((double)wf < w) wf = std::nextafter(wf, INFINITY);
    fast[k] = make_float4(fx, fy, fz, wf);
    double param = s.fuzz_or_ir;
    if (s.kind == RT_METAL) param = param < 1 ? param : 1;  // src/cpu/material.h:38
    if (s.kind == RT_DIELECTRIC)
      mat[k] = make_float4(1.f, 1.f, 1.f, (float)param);
    else
      mat[k] = make_float4((float)s.albedo[0], (float)s.albedo[1], (float)s.albedo[2],
                           s.kind == RT_METAL ? (float)param : 0.f);
    uint32_t kind_bits = s.kind;
    float kind_f;
    memcpy(&kind_f, &kind_bits, 4);
    aux[k] = make_float2((float)(1.0 / r), kind_f);
  }
  CU(cudaSetDevice(device));
  rt_scene *sc = new (std::nothrow) rt_scene();
  if (!sc) return fail(RT_E_NOMEM, "out of host memory");
  sc->device = device; sc->n = n; sc->n_pad = npad;
  size_t m = npad;
  cudaError_t e;
  if ((e = cudaMalloc(&sc->fast, (m + kRowTail) * sizeof(float4))) != cudaSuccess ||
      (e = cudaMalloc(&sc->exact, m * sizeof(double4))) != cudaSuccess ||
      (e = cudaMalloc(&sc->mat, m * sizeof(float4))) != cudaSuccess ||
      (e = cudaMalloc(&sc->aux, m * sizeof(float2))) != cudaSuccess) {
    rt_scene_destroy(sc);
    return fail(RT_E_CUDA, "cudaMalloc failed: %s", cudaGetErrorString(e));
  }
  CU_OR(rt_scene_destroy(sc), cudaMemcpy(sc->fast, fast.data(), (m + kRowTail) * sizeof(float4), cudaMemcpyHostToDevice));
  CU_OR(rt_scene_destroy(sc), cudaMemcpy(sc->exact, exact.data(), m * sizeof(double4), cudaMemcpyHostToDevice));
  CU_OR(rt_scene_destroy(sc), cudaMemcpy(sc->mat, mat.data(), m * sizeof(float4), cudaMemcpyHostToDevice));
  CU_OR(rt_scene_destroy(sc), cudaMemcpy(sc->aux, aux.data(), m * sizeof(float2), cudaMemcpyHostToDevice));
  sc->host_fast = fast;
  {
    // the north-star's optional BVH: built here on the host, used when a render asks for it
    std::vector<BvhNode> nodes;
    std::vector<uint32_t> prim;
    build_bvh(exact.data(), n, nodes, prim);
    if (prim.empty()) prim.push_back(0);
    sc->bvh_node_count = (uint32_t)nodes.size();
    if ((e = cudaMalloc(&sc->bvh_nodes, nodes.size() * sizeof(BvhNode))) != cudaSuccess ||
        (e = cudaMalloc(&sc->bvh_prim, prim.size() * sizeof(uint32_t))) != cudaSuccess) {
      rt_scene_destroy(sc);
      return fail(RT_E_CUDA, "cudaMalloc failed: %s", cudaGetErrorString(e));
    }
    CU_OR(rt_scene_destroy(sc), cudaMemcpy(sc->bvh_nodes, nodes.data(), nodes.size() * sizeof(BvhNode), cudaMemcpyHostToDevice));
    CU_OR(rt_scene_destroy(sc), cudaMemcpy(sc->bvh_prim, prim.data(), prim.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
  }
  *out = sc;
  return RT_OK;
}

extern "C" int rt_scene_destroy(rt_scene *sc) {
  if (!sc) return RT_OK;
  cudaSetDevice(sc->device);
  cudaFree(sc->fast); cudaFree(sc->exact); cudaFree(sc->mat); cudaFree(sc->aux);
  cudaFree(sc->bvh_nodes); cudaFree(sc->bvh_prim);
  delete sc;
  return RT_OK;
}

extern "C" int rt_scene_count(const rt_scene *sc, uint32_t *n) {
  if (!sc || !n) return fail(RT_E_INVALID, "NULL argument");
  *n = sc->n;
  return RT_OK;
}

static SceneView view_of(const rt_scene *sc) {
  SceneView v;
  v.fast = sc->fast; v.exact = sc->exact; v.mat = sc->mat; v.aux = sc->aux; v.n = sc->n; v.n_pad = sc->n_pad;
  v.bvh_nodes = sc->bvh_nodes; v.bvh_prim = sc->bvh_prim;
  return v;
}

// ---------------------------------------------------------------- camera
// camera::camera, src/cpu/camera.h:8-26, same operation order, in double.
struct CamFrame {
  double origin[3], llc[3], hor[3], ver[3], u[3], v[3], w[3], lens_radius;
};

static void unit3(const double a[3], double out[3]) {
  double inv = 1 / std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
  for (int k = 0; k < 3; ++k) out[k] = inv * a[k];
}
static void cross3(const double a[3], const double b[3], double out[3]) {
  out[0] = a[1] * b[2] - a[2] * b[1];
  out[1] = a[2] * b[0] - a[0] * b[2];
  out[2] = a[0] * b[1] - a[1] * b[0];
}

static void derive_camera(const rt_camera *c, CamFrame *f) {
  const double pi = 3.1415926535897932385;
  double theta = c->vfov_deg * pi / 180.0;
  double h = std::tan(theta / 2);
  double vh = 2.0 * h, vw = c->aspect_ratio * vh;
  double d[3] = {c->lookfrom[0] - c->lookat[0], c->lookfrom[1] - c->lookat[1],
                 c->lookfrom[2] - c->lookat[2]};
  unit3(d, f->w);
  double cr[3];
  cross3(c->vup, f->w, cr);
  unit3(cr, f->u);
  cross3(f->w, f->u, f->v);
  double kh = c->focus_dist * vw, kv = c->focus_dist * vh;
  for (int k = 0; k < 3; ++k) {
    f->origin[k] = c->lookfrom[k];
    f->hor[k] = kh * f->u[k];
    f->ver[k] = kv * f->v[k];
    f->llc[k] = f->origin[k] - 0.5 * f->hor[k] - 0.5 * f->ver[k] - c->focus_dist * f->w[k];
  }
  f->lens_radius = c->aperture / 2;
}

static CameraF32 narrow_camera(const CamFrame &f, uint32_t W, uint32_t H, bool pixel_square = false) {
  CameraF32 c;
  for (int k = 0; k < 3; ++k) {
    c.origin[k] = (float)f.origin[k];
    c.llc_minus_origin[k] = (float)(f.llc[k] - f.origin[k]);
    c.horizontal[k] = (float)f.hor[k]; c.vertical[k] = (float)f.ver[k];
    c.u[k] = (float)f.u[k]; c.v[k] = (float)f.v[k];
  }
  c.lens_radius = (float)f.lens_radius;
  // src/cpu/main.cc:116-117 divides by W-1 / H-1; the v4 camera (src/gpu/camera.h:92-103,140-157)
  // tiles the viewport with W x H pixel footprints: pixel00 + i du + j dv + (xi - 1/2)(du + dv)
  // = upper_left + (i + xi) viewport_u / W + (j + xi) viewport_v / H
  c.inv_wm1 = (float)(1.0 / (double)(pixel_square ? W : W - 1));
  c.inv_hm1 = (float)(1.0 / (double)(pixel_square ? H : H - 1));
  return c;
}

// ---------------------------------------------------------------- frame
extern "C" int rt_frame_create(uint32_t width, uint32_t height, int32_t device, rt_frame **out) {
  if (!out) return fail(RT_E_INVALID, "out is NULL");
  *out = nullptr;
  if (width < 2 || height < 2 || (uint64_t)width * height >= (1ull << kItemShift))
    return fail(RT_E_INVALID, "image size %ux%u out of range (min 2x2)", width, height);
  CU(cudaSetDevice(device));
  rt_frame *f = new (std::nothrow) rt_frame();
  if (!f) return fail(RT_E_NOMEM, "out of host memory");
  memset(f, 0, sizeof *f);
  f->device = device; f->w = width; f->h = height; f->npix = width * height;
  int sm_count = 0, smem_optin = 0;
  CU_OR(delete f, cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device));
  CU_OR(delete f, cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
  f->sm_count = sm_count;
  f->smem_optin = (size_t)smem_optin;
  size_t quads = (f->npix + 3) / 4;
  cudaError_t e;
  if ((e = cudaMalloc(&f->accum, (size_t)RT_SLICES * f->npix * sizeof(float4))) != cudaSuccess ||
      (e = cudaMalloc(&f->partial, (size_t)f->npix * sizeof(float4))) != cudaSuccess ||
      (e = cudaMalloc(&f->rgb_words, quads * 12)) != cudaSuccess ||
      (e = cudaMalloc(&f->linear, (size_t)f->npix * 3 * sizeof(float))) != cudaSuccess ||
      (e = cudaMalloc(&f->counters, 2 * sizeof(unsigned long long))) != cudaSuccess ||
      (e = cudaStreamCreateWithFlags(&f->own_stream, cudaStreamNonBlocking)) != cudaSuccess) {
    rt_frame_destroy(f);
    return fail(RT_E_CUDA, "frame allocation failed: %s", cudaGetErrorString(e));
  }
  f->stream = f->own_stream;
  for (int k = 0; k < 4; ++k) CU_OR(rt_frame_destroy(f), cudaEventCreate(&f->ev[k]));
  *out = f;
  return RT_OK;
}

extern "C" int rt_frame_destroy(rt_frame *f) {
  if (!f) return RT_OK;
  cudaSetDevice(f->device);
  cudaFree(f->accum); cudaFree(f->partial); cudaFree(f->rgb_words); cudaFree(f->linear);
  cudaFree(f->counters);
  cudaFree(f->pool); cudaFree(f->live);
  cudaFree(f->order); cudaFree(f->sort_keys); cudaFree(f->sort_keys_out); cudaFree(f->sort_iota);
  cudaFree(f->sort_temp); cudaFree(f->heavy);
  if (f->host_live) cudaFreeHost(f->host_live);
  for (int k = 0; k < 4; ++k)
    if (f->ev[k]) cudaEventDestroy(f->ev[k]);
  if (f->own_stream) cudaStreamDestroy(f->own_stream);
  delete f;
  return RT_OK;
}

extern "C" int rt_frame_set_stream(rt_frame *f, void *cuda_stream) {
  if (!f) return fail(RT_E_INVALID, "frame is NULL");
  f->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : f->own_stream;
  return RT_OK;
}

extern "C" int rt_frame_partial_ptr(rt_frame *f, void **dev_ptr, size_t *bytes) {
  if (!f) return fail(RT_E_INVALID, "frame is NULL");
  if (dev_ptr) *dev_ptr = f->partial;
  if (bytes) *bytes = (size_t)f->npix * sizeof(float4);
  return RT_OK;
}

static int check_cfg(const rt_config *cfg, const rt_frame *f) {
  if (!cfg || !f) return fail(RT_E_INVALID, "NULL argument");
  if (cfg->width != f->w || cfg->height != f->h)
    return fail(RT_E_INVALID, "config %ux%u does not match frame %ux%u", cfg->width, cfg->height,
                f->w, f->h);
  if (cfg->device != f->device) return fail(RT_E_INVALID, "config device != frame device");
  uint32_t cnt = cfg->slice_end - cfg->slice_begin;
  if (cfg->slice_begin >= cfg->slice_end || cfg->slice_end > RT_SLICES ||
      (cnt & (cnt - 1)) != 0 || (cfg->slice_begin % cnt) != 0)
    return fail(RT_E_INVALID, "slices [%u,%u) must be an aligned power-of-two block of %u",
                cfg->slice_begin, cfg->slice_end, RT_SLICES);
  if (cfg->samples_per_pixel == 0) return fail(RT_E_INVALID, "samples_per_pixel must be > 0");
  if ((cfg->flags & RT_FLAG_F32_HIT) && (cfg->flags & (RT_FLAG_BVH | RT_FLAG_WAVEFRONT)))
    return fail(RT_E_INVALID, "RT_FLAG_F32_HIT works with the brute-force fused kernel only");
  return RT_OK;
}

// ---------------------------------------------------------------- the hot path
// 0 = fused persistent kernel (default), 2 = two-kernel wavefront (A/B variant of the same code)
static int render_mode(const rt_config *cfg) {
  if (cfg->flags & RT_FLAG_F32_HIT) return 0;
  if (const char *e = getenv("RT_B200_MODE")) return strcmp(e, "wave") == 0 ? 2 : 0;
  return (cfg->flags & RT_FLAG_WAVEFRONT) ? 2 : 0;
}

extern "C" int rt_render(const rt_scene *sc, const rt_camera *cam, const rt_config *cfg,
                         rt_frame *f, rt_stats *stats) {
  if (!sc || !cam) return fail(RT_E_INVALID, "NULL argument");
  int rc = check_cfg(cfg, f);
  if (rc) return rc;
  if (sc->device != f->device) return fail(RT_E_INVALID, "scene and frame live on different devices");
  CU(cudaSetDevice(f->device));
  CamFrame cf;
  derive_camera(cam, &cf);
  WaveParams w;
  w.scene = view_of(sc);
  w.model_v4 = (cfg->flags & RT_FLAG_V4_MODEL) ? 1u : 0u;
  w.cam = narrow_camera(cf, f->w, f->h, w.model_v4 != 0u);
  w.width = f->w; w.height = f->h;
  w.spp = cfg->samples_per_pixel; w.max_depth = cfg->max_depth;
  w.seed = cfg->seed; w.sample_offset = cfg->sample_offset;
  w.slice_begin = cfg->slice_begin; w.slice_end = cfg->slice_end;
  w.accum = f->accum; w.counter = f->counters;
  const bool bvh_frame = (cfg->flags & RT_FLAG_BVH) != 0;
  w.order = f->order_valid && f->order_heavy_only == bvh_frame && !getenv("RT_B200_NO_COST_ORDER") ? f->order : nullptr;
  w.heavy = f->heavy;
  f->last_n = sc->n;
  // hand-out order = largest slices first; trailing slices without samples are not queued
  const uint32_t n_own = cfg->slice_end - cfg->slice_begin;
  w.n_orders = 0;
  for (uint32_t ord = 0; ord < n_own; ++ord) {
    uint32_t a, b;
    slice_range(slice_in_order(ord, cfg->slice_begin, cfg->slice_end), cfg->samples_per_pixel, a, b);
    if (b > a) w.n_orders = ord + 1;
  }

  uint32_t launches = 0;
  CU(cudaMemsetAsync(f->counters, 0, sizeof(unsigned long long), f->stream));
  CU(cudaEventRecord(f->ev[0], f->stream));
  if (cfg->max_depth == 0) {
    // ray_color returns black at depth <= 0 (src/cpu/main.cc:16-17)
    CU(cudaMemsetAsync(f->accum + (size_t)cfg->slice_begin * f->npix, 0,
                       (size_t)(cfg->slice_end - cfg->slice_begin) * f->npix * sizeof(float4),
                       f->stream));
  } else {
    for (uint32_t ord = w.n_orders; ord < n_own; ++ord)  // planes of the slices that are not queued
      CU(cudaMemsetAsync(f->accum + (size_t)slice_in_order(ord, cfg->slice_begin, cfg->slice_end) * f->npix,
                         0, (size_t)f->npix * sizeof(float4), f->stream));
    // path slots live in a pool in device memory (allocated on first use)
    if (!f->pool) {
      uint32_t slots = wave_default_slots(f->sm_count);
      if (const char *e = getenv("RT_B200_WAVE_SLOTS")) slots = (uint32_t)atol(e) / 1024u * 1024u;
      if (slots < 1024u) return fail(RT_E_INVALID, "RT_B200_WAVE_SLOTS too small");
      CU(cudaMalloc(&f->pool, wave_pool_bytes(slots)));
      CU(cudaMalloc(&f->live, sizeof(unsigned int)));
      CU(cudaMallocHost(&f->host_live, sizeof(unsigned int)));
      f->pool_slots = slots;
    }
    w.live = f->live;
    WaveRunInfo info;
    if (render_mode(cfg) == 2) {
      wave_carve(w, f->pool, f->pool_slots);
      CU(run_wavefront(w, sc->host_fast.data(), f->sm_count, f->smem_optin, f->stream, f->host_live, &info));
    } else {
      CU(run_fused(w, f->pool, f->pool_slots, sc->host_fast.data(), (cfg->flags & RT_FLAG_BVH) != 0,
                   (cfg->flags & RT_FLAG_F32_HIT) != 0, f->sm_count, f->smem_optin, f->stream, &info));
    }
    launches = info.launches;
  }
  CU(cudaEventRecord(f->ev[1], f->stream));
  CU(cudaStreamSynchronize(f->stream));
  if (stats) {
    memset(stats, 0, sizeof *stats);
    uint64_t s0 = 0;
    for (uint32_t k = cfg->slice_begin; k < cfg->slice_end; ++k) {
      uint32_t a, b;
      slice_range(k, cfg->samples_per_pixel, a, b);
      s0 += b - a;
    }
    stats->samples = s0 * f->npix;
    CU(cudaEventElapsedTime(&stats->render_ms, f->ev[0], f->ev[1]));
    stats->launches = launches;
  }
  return RT_OK;
}

extern "C" int rt_frame_reduce_local(rt_frame *f, const rt_config *cfg) {
  int rc = check_cfg(cfg, f);
  if (rc) return rc;
  CU(cudaSetDevice(f->device));
  CU(launch_reduce_local(f->accum, f->npix, cfg->slice_begin, cfg->slice_end, f->partial,
                         f->sm_count, f->stream));
  // the ray counts of this frame order the work queue of the next one (heaviest pixels first)
  if (!f->order) {
    f->sort_temp_bytes = cost_order_temp_bytes(f->npix);
    CU(cudaMalloc(&f->order, (size_t)f->npix * 4));
    CU(cudaMalloc(&f->sort_keys, (size_t)f->npix * 4));
    CU(cudaMalloc(&f->sort_keys_out, (size_t)f->npix * 4));
    CU(cudaMalloc(&f->sort_iota, (size_t)f->npix * 4));
    CU(cudaMalloc(&f->sort_temp, f->sort_temp_bytes ? f->sort_temp_bytes : 16));
    CU(cudaMalloc(&f->heavy, 4));
  }
  uint32_t own = 0;
  for (uint32_t k = cfg->slice_begin; k < cfg->slice_end; ++k) {
    uint32_t a, b;
    slice_range(k, cfg->samples_per_pixel, a, b);
    own += b - a;
  }
  if (own) {
    CU(launch_cost_order(f->partial, f->npix, 1.0f / (float)own, f->sort_keys, f->sort_keys_out,
                         f->sort_iota, f->order, f->heavy, (cfg->flags & RT_FLAG_BVH) != 0, f->sort_temp,
                         f->sort_temp_bytes, f->sm_count, f->stream));
    f->order_valid = true;
    f->order_heavy_only = (cfg->flags & RT_FLAG_BVH) != 0;
  }
  CU(cudaStreamSynchronize(f->stream));
  return RT_OK;
}

extern "C" int rt_finalize(rt_frame *f, const rt_config *cfg, const void *const *partials,
                           uint32_t count, uint8_t *rgb8_out, float *linear_out, rt_stats *stats) {
  if (!f || !cfg || !partials) return fail(RT_E_INVALID, "NULL argument");
  if (cfg->width != f->w || cfg->height != f->h) return fail(RT_E_INVALID, "config/frame size mismatch");
  if (!(count == 1 || count == 2 || count == 4 || count == 8))
    return fail(RT_E_INVALID, "partial count must be 1, 2, 4 or 8 (got %u)", count);
  if (cfg->samples_per_pixel == 0) return fail(RT_E_INVALID, "samples_per_pixel must be > 0");
  CU(cudaSetDevice(f->device));
  const float4 *ptrs[8] = {};
  for (uint32_t k = 0; k < count; ++k) {
    if (!partials[k]) return fail(RT_E_INVALID, "partials[%u] is NULL", k);
    ptrs[k] = static_cast<const float4 *>(partials[k]);
  }
  CU(cudaMemsetAsync(f->counters + 1, 0, sizeof(unsigned long long), f->stream));
  CU(cudaEventRecord(f->ev[2], f->stream));
  float scale = 1.0f / (float)cfg->samples_per_pixel;
  CU(launch_finalize(ptrs, count, f->npix, scale, f->rgb_words, linear_out ? f->linear : nullptr,
                     f->counters + 1, f->sm_count, f->stream));
  CU(cudaEventRecord(f->ev[3], f->stream));
  if (rgb8_out)
    CU(cudaMemcpyAsync(rgb8_out, f->rgb_words, (size_t)f->npix * 3, cudaMemcpyDeviceToHost, f->stream));
  if (linear_out)
    CU(cudaMemcpyAsync(linear_out, f->linear, (size_t)f->npix * 3 * sizeof(float),
                       cudaMemcpyDeviceToHost, f->stream));
  unsigned long long rays = 0;
  CU(cudaMemcpyAsync(&rays, f->counters + 1, sizeof rays, cudaMemcpyDeviceToHost, f->stream));
  CU(cudaStreamSynchronize(f->stream));
  if (stats) {
    stats->rays = rays;
    stats->sphere_tests = rays * (uint64_t)f->last_n;
    CU(cudaEventElapsedTime(&stats->finalize_ms, f->ev[2], f->ev[3]));
    if (cudaEventElapsedTime(&stats->frame_ms, f->ev[0], f->ev[3]) != cudaSuccess) {
      stats->frame_ms = 0.f;  // rt_render has not run on this frame
      (void)cudaGetLastError();
    }
    stats->launches += 1;
  }
  return RT_OK;
}

extern "C" int rt_render_image(const rt_scene *sc, const rt_camera *cam, const rt_config *cfg_in,
                               uint8_t *rgb8_out, float *linear_out, rt_stats *stats) {
  if (!sc || !cam || !cfg_in) return fail(RT_E_INVALID, "NULL argument");
  rt_config cfg = *cfg_in;
  cfg.slice_begin = 0; cfg.slice_end = RT_SLICES;
  rt_frame *f = nullptr;
  int rc = rt_frame_create(cfg.width, cfg.height, cfg.device, &f);
  if (rc) return rc;
  rt_stats st;
  memset(&st, 0, sizeof st);
  rc = rt_render(sc, cam, &cfg, f, &st);
  if (!rc) rc = rt_frame_reduce_local(f, &cfg);
  if (!rc) {
    st.launches += 1;
    const void *parts[1] = {f->partial};
    rc = rt_finalize(f, &cfg, parts, 1, rgb8_out, linear_out, &st);
  }
  rt_frame_destroy(f);
  if (!rc && stats) *stats = st;
  return rc;
}

// ---------------------------------------------------------------- one process, several GPUs
extern "C" int rt_render_image_multi(const rt_sphere *spheres, uint32_t n, const rt_camera *cam,
                                     const rt_config *cfg_in, const int32_t *devices, uint32_t ngpu,
                                     uint8_t *rgb8_out, float *linear_out, rt_stats *stats) {
  if (!cam || !cfg_in || !devices) return fail(RT_E_INVALID, "NULL argument");
  if (!(ngpu == 1 || ngpu == 2 || ngpu == 4 || ngpu == 8))
    return fail(RT_E_INVALID, "ngpu must be 1, 2, 4 or 8 (got %u)", ngpu);
  for (uint32_t a = 0; a < ngpu; ++a)
    for (uint32_t b = a + 1; b < ngpu; ++b)
      if (devices[a] == devices[b]) return fail(RT_E_INVALID, "device %d listed twice", devices[a]);
  std::vector<rt_scene *> scenes(ngpu, nullptr);
  std::vector<rt_frame *> frames(ngpu, nullptr);
  std::vector<rt_stats> st(ngpu);
  std::vector<int> rcs(ngpu, RT_OK);
  std::vector<std::string> errs(ngpu);
  const uint32_t per = RT_SLICES / ngpu;
  auto work = [&](uint32_t g) {
    rt_config c = *cfg_in;
    c.device = devices[g];
    c.slice_begin = g * per; c.slice_end = (g + 1) * per;
    int rc = rt_scene_create(spheres, n, c.device, &scenes[g]);
    if (!rc) rc = rt_frame_create(c.width, c.height, c.device, &frames[g]);
    memset(&st[g], 0, sizeof(rt_stats));
    if (!rc) rc = rt_render(scenes[g], cam, &c, frames[g], &st[g]);
    if (!rc) rc = rt_frame_reduce_local(frames[g], &c);
    rcs[g] = rc;
    if (rc) errs[g] = rt_last_error();  // the message is thread-local
  };
  {
    std::vector<std::thread> th;
    for (uint32_t g = 1; g < ngpu; ++g) th.emplace_back(work, g);
    work(0);
    for (auto &t : th) t.join();
  }
  int rc = RT_OK;
  for (uint32_t g = 0; g < ngpu && !rc; ++g)
    if (rcs[g]) rc = fail(rcs[g], "GPU %d: %s", devices[g], errs[g].c_str());
  std::vector<float4 *> staged(ngpu, nullptr);
  if (!rc) {
    // partial sums of the other GPUs -> devices[0], bytes only (peer copy over NVLink when the
    // topology allows it, through the host otherwise)
    rt_frame *f0 = frames[0];
    cudaError_t e = cudaSetDevice(devices[0]);
    const size_t bytes = (size_t)f0->npix * sizeof(float4);
    std::vector<const void *> parts(ngpu);
    parts[0] = f0->partial;
    for (uint32_t g = 1; g < ngpu && e == cudaSuccess; ++g) {
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, devices[0], devices[g]) == cudaSuccess && can) {
        cudaError_t pe = cudaDeviceEnablePeerAccess(devices[g], 0);
        if (pe != cudaSuccess) (void)cudaGetLastError();  // already enabled
      }
      e = cudaMalloc(&staged[g], bytes);
      if (e == cudaSuccess) e = cudaMemcpyPeer(staged[g], devices[0], frames[g]->partial, devices[g], bytes);
      parts[g] = staged[g];
    }
    if (e != cudaSuccess) rc = fail(RT_E_CUDA, "gathering partial sums failed: %s", cudaGetErrorString(e));
    rt_config call = *cfg_in;
    call.device = devices[0]; call.slice_begin = 0; call.slice_end = RT_SLICES;
    rt_stats fin;
    memset(&fin, 0, sizeof fin);
    if (!rc) rc = rt_finalize(f0, &call, parts.data(), ngpu, rgb8_out, linear_out, &fin);
    if (!rc && stats) {
      *stats = fin;
      stats->samples = 0; stats->render_ms = 0.f;
      for (uint32_t g = 0; g < ngpu; ++g) {
        stats->samples += st[g].samples;
        stats->launches += st[g].launches + 1;
        if (st[g].render_ms > stats->render_ms) stats->render_ms = st[g].render_ms;
      }
      stats->frame_ms = stats->render_ms + fin.finalize_ms;
    }
  }
  for (uint32_t g = 0; g < ngpu; ++g) {
    if (staged[g]) { cudaSetDevice(devices[0]); cudaFree(staged[g]); }
    rt_frame_destroy(frames[g]);
    rt_scene_destroy(scenes[g]);
  }
  return rc;
}

// ---------------------------------------------------------------- accuracy gate
extern "C" int rt_primary_hits(const rt_scene *sc, const rt_camera *cam, uint32_t W, uint32_t H,
                               int32_t device, int use_f64, int32_t *ids, double *t, double *normal,
                               uint8_t *front_face) {
  if (!sc || !cam || !ids || !t || !normal || !front_face) return fail(RT_E_INVALID, "NULL argument");
  if (W < 2 || H < 2) return fail(RT_E_INVALID, "image size %ux%u out of range (min 2x2)", W, H);
  if (device != sc->device) return fail(RT_E_INVALID, "scene lives on another device");
  CU(cudaSetDevice(device));
  size_t npix = (size_t)W * H;
  int32_t *d_ids = nullptr; double *d_t = nullptr, *d_n = nullptr; uint8_t *d_ff = nullptr;
  CU(cudaMalloc(&d_ids, npix * 4));
  CU(cudaMalloc(&d_t, npix * 8));
  CU(cudaMalloc(&d_n, npix * 24));
  CU(cudaMalloc(&d_ff, npix));
  CamFrame cf;
  derive_camera(cam, &cf);
  cudaError_t e;
  if (use_f64) {
    CameraF64 c64;
    for (int k = 0; k < 3; ++k) {
      c64.origin[k] = cf.origin[k]; c64.lower_left[k] = cf.llc[k];
      c64.horizontal[k] = cf.hor[k]; c64.vertical[k] = cf.ver[k];
    }
    e = launch_primary_f64(sc->exact, sc->n, c64, W, H, d_ids, d_t, d_n, d_ff, 0);
  } else {
    e = launch_primary_f32(view_of(sc), narrow_camera(cf, W, H), W, H, d_ids, d_t, d_n, d_ff, 0);
  }
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e == cudaSuccess) e = cudaMemcpy(ids, d_ids, npix * 4, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(t, d_t, npix * 8, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(normal, d_n, npix * 24, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(front_face, d_ff, npix, cudaMemcpyDeviceToHost);
  cudaFree(d_ids); cudaFree(d_t); cudaFree(d_n); cudaFree(d_ff);
  if (e != cudaSuccess) return fail(RT_E_CUDA, "primary-hit pass failed: %s", cudaGetErrorString(e));
  return RT_OK;
}

// ---------------------------------------------------------------- measurement aid
extern "C" int rt_measure_fp32_peak(int32_t device, float *tflops, float *sm_clock_mhz_est) {
  if (!tflops) return fail(RT_E_INVALID, "NULL argument");
  CU(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  const int blocks = prop.multiProcessorCount * 8, iters = 20000;
  float *out = nullptr;
  long long *cyc = nullptr, hcyc = 0;
  CU(cudaMalloc(&out, (size_t)blocks * 256 * sizeof(float)));
  CU(cudaMalloc(&cyc, sizeof(long long)));
  cudaEvent_t a, b;
  CU(cudaEventCreate(&a));
  CU(cudaEventCreate(&b));
  float best = 0.f, best_ms = 0.f;
  for (int rep = 0; rep < 4; ++rep) {  // first rep is the warm-up
    CU(cudaEventRecord(a, 0));
    CU(launch_ffma_peak(out, blocks, iters, cyc, 0));
    CU(cudaEventRecord(b, 0));
    CU(cudaEventSynchronize(b));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, a, b));
    double flops = 2.0 * 64.0 * (double)iters * 256.0 * blocks;
    float tf = (float)(flops / (ms * 1e-3) / 1e12);
    if (rep > 0 && tf > best) { best = tf; best_ms = ms; }
  }
  CU(cudaMemcpy(&hcyc, cyc, sizeof hcyc, cudaMemcpyDeviceToHost));
  cudaFree(out); cudaFree(cyc); cudaEventDestroy(a); cudaEventDestroy(b);
  *tflops = best;
  // block 0's cycle count over the kernel's wall time ~ SM clock (block 0 runs the whole time
  // only if there is a single wave: 8 blocks/SM x 256 threads fits, so it does)
  if (sm_clock_mhz_est) *sm_clock_mhz_est = best_ms > 0 ? (float)(hcyc / (best_ms * 1e3)) : 0.f;
  return RT_OK;
}

// ---------------------------------------------------------------- output
extern "C" int rt_write_ppm(const char *path, const uint8_t *rgb8, uint32_t W, uint32_t H) {
  if (!path || !rgb8) return fail(RT_E_INVALID, "NULL argument");
  FILE *fp = strcmp(path, "-") == 0 ? stdout : fopen(path, "w");
  if (!fp) return fail(RT_E_IO, "cannot open %s", path);
  fprintf(fp, "P3\n%u %u\n255\n", W, H);  // src/cpu/main.cc:109
  // "r g b\n" per pixel (src/cpu/color.h:20-22), formatted through a small digit table
  static char lut[256][4];
  static int lut_len[256];
  if (!lut_len[255])
    for (int v = 0; v < 256; ++v) lut_len[v] = snprintf(lut[v], 4, "%d", v);
  std::vector<char> line((size_t)W * 12 + 1);
  for (uint32_t y = 0; y < H; ++y) {
    char *o = line.data();
    const uint8_t *p = rgb8 + (size_t)y * W * 3;
    for (uint32_t x = 0; x < W; ++x, p += 3) {
      for (int c = 0; c < 3; ++c) {
        memcpy(o, lut[p[c]], lut_len[p[c]]);
        o += lut_len[p[c]];
        *o++ = c == 2 ? '\n' : ' ';
      }
    }
    if (fwrite(line.data(), 1, (size_t)(o - line.data()), fp) != (size_t)(o - line.data())) {
      if (fp != stdout) fclose(fp);
      return fail(RT_E_IO, "short write to %s", path);
    }
  }
  if (fp != stdout) fclose(fp);
  return RT_OK;
}
