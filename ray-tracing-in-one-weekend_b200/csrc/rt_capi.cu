// csrc/rt_capi.cu — the C ABI of include/rt_b200.h: host-side flattening of the reference's
// scene/camera objects, buffer management, kernel launches, error reporting.
// No CPU fallback: every compute entry point needs a CUDA device.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"
#include "rt_capi_internal.h"

using namespace rtb;

static thread_local char g_err[512] = "";

int rt_fail(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}
#define fail rt_fail

extern "C" const char *rt_last_error(void) { return g_err; }
extern "C" const char *rt_version(void) { return "rt_b200 0.2 (sm_100a)"; }

extern "C" int rt_device_count(int *count) {
  if (!count) return fail(RT_E_INVALID, "count is NULL");
  CU(cudaGetDeviceCount(count));
  return RT_OK;
}

// RT_B200_* experiment switches: read once per process (C++11 thread-safe static)
namespace rtb {
const EnvCfg &env_cfg() {
  static const EnvCfg cfg = [] {
    EnvCfg c;
    const char *e;
    c.no_tc = getenv("RT_B200_NO_TC") != nullptr;
    c.no_tc_stream = getenv("RT_B200_NO_TC_STREAM") != nullptr;
    c.no_const_rows = getenv("RT_B200_NO_CONST_ROWS") != nullptr;
    c.no_cost_order = getenv("RT_B200_NO_COST_ORDER") != nullptr;
    e = getenv("RT_B200_HEAVY_QUARTER_RAYS");
    c.heavy_quarter_rays = e ? atoi(e) : -1;
    return c;
  }();
  return cfg;
}
}  // namespace rtb

// ---------------------------------------------------------------- scene
// One device allocation ("blob") and one pinned staging buffer of the same layout:
//   exact [cap] double4 | fast [cap + kRowTail] float4 | mat [cap] float4 | aux [cap] float2
static size_t scene_blob_bytes(uint32_t cap) {
  return (size_t)cap * 32 + ((size_t)cap + kRowTail) * 16 + (size_t)cap * 16 + (size_t)cap * 8;
}

static void scene_views(unsigned char *b, uint32_t cap, double4 *&exact, float4 *&fast, float4 *&mat,
                        float2 *&aux) {
  exact = reinterpret_cast<double4 *>(b); b += (size_t)cap * 32;
  fast = reinterpret_cast<float4 *>(b); b += ((size_t)cap + kRowTail) * 16;
  mat = reinterpret_cast<float4 *>(b); b += (size_t)cap * 16;
  aux = reinterpret_cast<float2 *>(b);
}

static int check_spheres(const rt_sphere *sp, uint32_t n) {
  if (n && !sp) return fail(RT_E_INVALID, "spheres is NULL");
  if (n > (1u << 24)) return fail(RT_E_INVALID, "too many spheres (%u)", n);
  for (uint32_t k = 0; k < n; ++k) {
    if (sp[k].kind > RT_KIND_MAX) return fail(RT_E_INVALID, "sphere %u: bad material kind %u", k, sp[k].kind);
    if (!(sp[k].radius != 0.0) || !std::isfinite(sp[k].radius))
      return fail(RT_E_INVALID, "sphere %u: radius must be finite and non-zero", k);
  }
  return RT_OK;
}

// hittable_list (insertion order) -> the four SoA row arrays, in the staging buffer
static void flatten_scene(const rt_sphere *sp, uint32_t n, uint32_t cap, unsigned char *stage) {
  double4 *exact; float4 *fast, *mat; float2 *aux;
  scene_views(stage, cap, exact, fast, mat, aux);
  // a padding row has w = -inf so that b*b + q = -inf never reaches the threshold
  for (uint32_t k = 0; k < cap + kRowTail; ++k) fast[k] = make_float4(0.f, 0.f, 0.f, -INFINITY);
  for (uint32_t k = 0; k < cap; ++k) {
    exact[k] = make_double4(0, 0, 0, 1);
    mat[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    aux[k] = make_float2(1.f, 0.f);
  }
  for (uint32_t k = 0; k < n; ++k) {
    const rt_sphere &s = sp[k];
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
}

static uint32_t padded_rows(uint32_t n) { return n ? (n + kRowPad - 1) / kRowPad * kRowPad : kRowPad; }

extern "C" int rt_scene_create(const rt_sphere *sp, uint32_t n, int32_t device, rt_scene **out) {
  if (!out) return fail(RT_E_INVALID, "out is NULL");
  *out = nullptr;
  int rc = check_spheres(sp, n);
  if (rc) return rc;
  CU(cudaSetDevice(device));
  rt_scene *sc = new (std::nothrow) rt_scene();
  if (!sc) return fail(RT_E_NOMEM, "out of host memory");
  sc->device = device; sc->n = n;
  sc->n_pad = sc->cap = padded_rows(n);
  sc->blob_bytes = scene_blob_bytes(sc->cap);
  CU_OR(rt_scene_destroy(sc), cudaMalloc(&sc->blob, sc->blob_bytes));
  CU_OR(rt_scene_destroy(sc), cudaMallocHost(&sc->stage, sc->blob_bytes));
  CU_OR(rt_scene_destroy(sc), cudaEventCreateWithFlags(&sc->staged, cudaEventDisableTiming));
  scene_views(sc->blob, sc->cap, sc->exact, sc->fast, sc->mat, sc->aux);
  flatten_scene(sp, n, sc->cap, sc->stage);
  CU_OR(rt_scene_destroy(sc), cudaMemcpy(sc->blob, sc->stage, sc->blob_bytes, cudaMemcpyHostToDevice));
  {
    double4 *e; float4 *f, *m; float2 *a;
    scene_views(sc->stage, sc->cap, e, f, m, a);
    sc->host_fast.assign(f, f + sc->cap + kRowTail);
  }
  *out = sc;
  return RT_OK;
}

extern "C" int rt_scene_update(rt_scene *sc, const rt_sphere *sp, uint32_t n, void *cuda_stream) {
  if (!sc) return fail(RT_E_INVALID, "scene is NULL");
  int rc = check_spheres(sp, n);
  if (rc) return rc;
  if (padded_rows(n) > sc->cap)
    return fail(RT_E_INVALID, "rt_scene_update: %u spheres do not fit the scene's capacity of %u rows", n, sc->cap);
  CU(cudaSetDevice(sc->device));
  // the staging buffer may still feed the previous update's copy
  if (sc->staged_pending) CU(cudaEventSynchronize(sc->staged));
  sc->staged_pending = false;
  flatten_scene(sp, n, sc->cap, sc->stage);
  sc->n = n; sc->n_pad = padded_rows(n);
  {
    double4 *e; float4 *f, *m; float2 *a;
    scene_views(sc->stage, sc->cap, e, f, m, a);
    sc->host_fast.assign(f, f + sc->cap + kRowTail);
  }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  CU(cudaMemcpyAsync(sc->blob, sc->stage, sc->blob_bytes, cudaMemcpyHostToDevice, st));
  CU(cudaEventRecord(sc->staged, st));
  sc->staged_pending = true;
  std::lock_guard<std::mutex> lk(sc->bvh_mu);
  sc->bvh_built = false;  // rebuilt by the next RT_FLAG_BVH render
  sc->tc_built = false;   // ... and the tensor-core operand by the next render that uses it
  return RT_OK;
}

extern "C" int rt_scene_destroy(rt_scene *sc) {
  if (!sc) return RT_OK;
  cudaSetDevice(sc->device);
  if (sc->staged_pending) cudaEventSynchronize(sc->staged);
  cudaFree(sc->blob);
  if (sc->stage) cudaFreeHost(sc->stage);
  if (sc->staged) cudaEventDestroy(sc->staged);
  cudaFree(sc->bvh_nodes); cudaFree(sc->bvh_prim); cudaFree(sc->tc_op);
  delete sc;
  return RT_OK;
}

extern "C" int rt_scene_count(const rt_scene *sc, uint32_t *n) {
  if (!sc || !n) return fail(RT_E_INVALID, "NULL argument");
  *n = sc->n;
  return RT_OK;
}

// The north-star's optional BVH: built on the host (rt_bvh.cu) at the first render that asks
// for it, from the exact rows in the staging buffer; uploaded once.
int rt_scene_ensure_bvh(const rt_scene *csc) {
  rt_scene *sc = const_cast<rt_scene *>(csc);
  std::lock_guard<std::mutex> lk(sc->bvh_mu);
  if (sc->bvh_built) return RT_OK;
  if (sc->staged_pending) CU(cudaEventSynchronize(sc->staged));
  double4 *e; float4 *f, *m; float2 *a;
  scene_views(sc->stage, sc->cap, e, f, m, a);
  std::vector<BvhNode> nodes;
  std::vector<uint32_t> prim;
  build_bvh(e, sc->n, nodes, prim);
  if (prim.empty()) prim.push_back(0);
  BvhNode *dn = nullptr;
  uint32_t *dp = nullptr;
  CU(cudaMalloc(&dn, nodes.size() * sizeof(BvhNode)));
  CU_OR(cudaFree(dn), cudaMalloc(&dp, prim.size() * sizeof(uint32_t)));
  CU_OR((cudaFree(dn), cudaFree(dp)), cudaMemcpy(dn, nodes.data(), nodes.size() * sizeof(BvhNode), cudaMemcpyHostToDevice));
  CU_OR((cudaFree(dn), cudaFree(dp)), cudaMemcpy(dp, prim.data(), prim.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
  cudaFree(sc->bvh_nodes); cudaFree(sc->bvh_prim);
  sc->bvh_nodes = dn; sc->bvh_prim = dp;
  sc->bvh_node_count = (uint32_t)nodes.size();
  sc->bvh_built = true;
  return RT_OK;
}

// Sphere operand of the tensor-core filter (rt_tc.cu), from the exact rows in the staging buffer.
int rt_scene_ensure_tc(const rt_scene *csc) {
  rt_scene *sc = const_cast<rt_scene *>(csc);
  std::lock_guard<std::mutex> lk(sc->bvh_mu);
  if (sc->tc_built) return RT_OK;
  if (sc->staged_pending) CU(cudaEventSynchronize(sc->staged));
  std::vector<float> op;
  uint32_t tiles = 0;
  tc_build_sphere_operand(reinterpret_cast<const double4 *>(sc->stage), sc->n, op, tiles);
  float *d = nullptr;
  CU(cudaMalloc(&d, op.size() * sizeof(float)));
  CU_OR(cudaFree(d), cudaMemcpy(d, op.data(), op.size() * sizeof(float), cudaMemcpyHostToDevice));
  cudaFree(sc->tc_op);
  sc->tc_op = d; sc->tc_tiles = tiles; sc->tc_built = true;
  return RT_OK;
}

SceneView rt_scene_view(const rt_scene *sc) {
  SceneView v;
  v.fast = sc->fast; v.exact = sc->exact; v.mat = sc->mat; v.aux = sc->aux; v.n = sc->n; v.n_pad = sc->n_pad;
  v.bvh_nodes = sc->bvh_built ? sc->bvh_nodes : nullptr;
  v.bvh_prim = sc->bvh_built ? sc->bvh_prim : nullptr;
  v.tc_op = sc->tc_built ? sc->tc_op : nullptr;
  v.tc_tiles = sc->tc_built ? sc->tc_tiles : 0u;
  return v;
}

// ---------------------------------------------------------------- camera
// camera::camera, src/cpu/camera.h:8-26, same operation order, in double.
static void unit3(const double a[3], double out[3]) {
  double inv = 1 / std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
  for (int k = 0; k < 3; ++k) out[k] = inv * a[k];
}
static void cross3(const double a[3], const double b[3], double out[3]) {
  out[0] = a[1] * b[2] - a[2] * b[1];
  out[1] = a[2] * b[0] - a[0] * b[2];
  out[2] = a[0] * b[1] - a[1] * b[0];
}

void rt_derive_camera(const rt_camera *c, CamFrame *f) {
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

CameraF32 rt_narrow_camera(const CamFrame &f, uint32_t W, uint32_t H, bool pixel_square) {
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
  if (width < 2 || height < 2 || width > 65535u || height > 65535u ||
      (uint64_t)width * height >= (1ull << kItemShift))
    return fail(RT_E_INVALID, "image size %ux%u out of range (2..65535 per side, fewer than 2^%u pixels)",
                width, height, kItemShift);
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
      (e = cudaMalloc(&f->counters, 4 * sizeof(unsigned long long))) != cudaSuccess ||
      (e = cudaMallocHost(&f->host_words, 4 * sizeof(unsigned long long))) != cudaSuccess ||
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
  cudaFree(f->pool);
  cudaFree(f->order); cudaFree(f->sort_keys); cudaFree(f->sort_keys_out); cudaFree(f->sort_iota);
  cudaFree(f->sort_temp); cudaFree(f->heavy);
  if (f->host_words) cudaFreeHost(f->host_words);
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
  if ((cfg->flags & RT_FLAG_F32_HIT) && (cfg->flags & RT_FLAG_BVH))
    return fail(RT_E_INVALID, "RT_FLAG_F32_HIT works with the brute-force scan only");
  if ((cfg->flags & RT_FLAG_SIMT_FILTER) && (cfg->flags & RT_FLAG_BVH))
    return fail(RT_E_INVALID, "RT_FLAG_SIMT_FILTER selects between the two brute-force filters; not with RT_FLAG_BVH");
  if ((cfg->flags & RT_FLAG_NO_FILTER) && (cfg->flags & (RT_FLAG_BVH | RT_FLAG_F32_HIT)))
    return fail(RT_E_INVALID, "RT_FLAG_NO_FILTER works with the brute-force scan only");
  if (cfg->flags & ~RT_FLAG_ALL) return fail(RT_E_INVALID, "unknown flag bits 0x%x", cfg->flags & ~RT_FLAG_ALL);
  return RT_OK;
}

static uint64_t samples_of(const rt_config *cfg, uint32_t npix) {
  uint64_t s0 = 0;
  for (uint32_t k = cfg->slice_begin; k < cfg->slice_end; ++k) {
    uint32_t a, b;
    slice_range(k, cfg->samples_per_pixel, a, b);
    s0 += b - a;
  }
  return s0 * npix;
}

// ---------------------------------------------------------------- the hot path
// path-slot pool, allocated on first use
static int ensure_pool(rt_frame *f) {
  if (f->pool) return RT_OK;
  const uint32_t slots = wave_default_slots(f->sm_count);
  CU(cudaMalloc(&f->pool, wave_pool_bytes(slots)));
  f->pool_slots = slots;
  return RT_OK;
}

// Enqueue the path kernel of one frame's slices on the frame's stream (no host synchronisation).
int rt_render_enqueue(const rt_scene *sc, const rt_camera *cam, const rt_config *cfg, rt_frame *f) {
  if (!sc || !cam) return fail(RT_E_INVALID, "NULL argument");
  int rc = check_cfg(cfg, f);
  if (rc) return rc;
  if (sc->device != f->device) return fail(RT_E_INVALID, "scene and frame live on different devices");
  CU(cudaSetDevice(f->device));
  // an empty world has no hierarchy to walk (its root box would be degenerate): brute force
  const bool use_bvh = (cfg->flags & RT_FLAG_BVH) != 0 && sc->n > 0;
  if (use_bvh && (rc = rt_scene_ensure_bvh(sc)) != RT_OK) return rc;
  // the brute-force filter runs on the tensor cores from 32 spheres up (a handful of spheres is a single
  // trip of the FFMA2 loop: not worth a tensor-core round trip): the sphere operand resident in shared
  // memory up to kTcMaxSpheres, streamed through it in chunks beyond; RT_FLAG_SIMT_FILTER / RT_B200_NO_TC
  // keep the FFMA2 loop, RT_B200_NO_TC_STREAM keeps it for the scenes that would stream
  const bool use_tc = !use_bvh && sc->n >= 32u && (sc->n <= kTcMaxSpheres || !env_cfg().no_tc_stream) &&
                      !(cfg->flags & (RT_FLAG_NO_FILTER | RT_FLAG_SIMT_FILTER)) && !env_cfg().no_tc;
  if (use_tc && (rc = rt_scene_ensure_tc(sc)) != RT_OK) return rc;
  CamFrame cf;
  rt_derive_camera(cam, &cf);
  WaveParams w;
  w.scene = rt_scene_view(sc);
  w.model_v4 = (cfg->flags & RT_FLAG_V4_MODEL) ? 1u : 0u;
  w.no_filter = (cfg->flags & RT_FLAG_NO_FILTER) ? 1u : 0u;
  w.cam = rt_narrow_camera(cf, f->w, f->h, w.model_v4 != 0u);
  w.width = f->w; w.height = f->h;
  w.inv_width = 1.0f / (float)f->w;
  w.spp = cfg->samples_per_pixel; w.max_depth = cfg->max_depth;
  w.seed = cfg->seed; w.sample_offset = cfg->sample_offset;
  w.accum = f->accum; w.counter = f->counters;
  w.status = reinterpret_cast<unsigned int *>(f->counters + 2);
  w.order = f->order_valid && f->order_heavy_only == use_bvh && !env_cfg().no_cost_order ? f->order : nullptr;
  w.heavy = f->heavy;
  f->last_n = sc->n;
  f->last_samples = samples_of(cfg, f->npix);
  // hand-out order = largest slices first; trailing slices without samples are not queued
  const uint32_t n_own = cfg->slice_end - cfg->slice_begin;
  w.n_orders = 0;
  memset(w.order_slice, 0, sizeof w.order_slice);
  for (uint32_t ord = 0; ord < n_own; ++ord) {
    uint32_t a, b;
    const uint32_t sl = slice_in_order(ord, cfg->slice_begin, cfg->slice_end);
    w.order_slice[ord] = (uint8_t)sl;
    slice_range(sl, cfg->samples_per_pixel, a, b);
    if (b > a) w.n_orders = ord + 1;
  }
  // canonical slices are contiguous sample ranges: slice s = [bounds[s], bounds[s + 1])
  for (uint32_t sl = 0; sl < RT_SLICES; ++sl) {
    uint32_t a, b;
    slice_range(sl, cfg->samples_per_pixel, a, b);
    w.slice_bounds[sl] = a;
    w.slice_bounds[sl + 1] = b;
  }

  uint32_t launches = 0;
  f->last_tc = false; f->last_filter = RT_FILTER_NONE;
  CU(cudaMemsetAsync(f->counters, 0, sizeof(unsigned long long), f->stream));
  CU(cudaMemsetAsync(f->counters + 2, 0, sizeof(unsigned long long), f->stream));
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
    if ((rc = ensure_pool(f)) != RT_OK) return rc;
    WaveRunInfo info;
    CU(run_fused(w, f->pool, f->pool_slots, sc->host_fast.data(), use_bvh,
                 (cfg->flags & RT_FLAG_F32_HIT) != 0, use_tc, f->sm_count, f->smem_optin, f->stream, &info));
    launches = info.launches;
    f->last_tc = info.tensor_core_filter;
    f->last_filter = info.tensor_core_filter ? RT_FILTER_TENSOR
                     : (use_bvh || (cfg->flags & RT_FLAG_NO_FILTER)) ? RT_FILTER_NONE : RT_FILTER_SIMT;
  }
  CU(cudaEventRecord(f->ev[1], f->stream));
  CU(cudaMemcpyAsync(f->host_words + 2, f->counters + 2, sizeof(unsigned long long), cudaMemcpyDeviceToHost, f->stream));
  f->rendered = true;
  f->last_launches = launches;
  return RT_OK;
}

static bool caller_stream(const rt_frame *f) { return f->stream != f->own_stream; }

extern "C" int rt_render(const rt_scene *sc, const rt_camera *cam, const rt_config *cfg,
                         rt_frame *f, rt_stats *stats) {
  int rc = rt_render_enqueue(sc, cam, cfg, f);
  if (rc) return rc;
  if (stats) {
    memset(stats, 0, sizeof *stats);
    stats->samples = f->last_samples;
    stats->launches = f->last_launches;
    stats->filter_unit = f->last_filter;
  }
  if (caller_stream(f)) return RT_OK;  // asynchronous: timings through rt_frame_stats()
  CU(cudaStreamSynchronize(f->stream));
  if (f->host_words[2]) return fail(RT_E_CUDA, "a tensor-core MMA of the filter never completed (bounded wait expired)");
  if (stats) CU(cudaEventElapsedTime(&stats->render_ms, f->ev[0], f->ev[1]));
  return RT_OK;
}

// hand-out order buffers of the next frame, allocated together
static int ensure_order(rt_frame *f) {
  if (f->order) return RT_OK;
  const size_t tb = cost_order_temp_bytes(f->npix);
  void *p[6] = {};
  const size_t sz[6] = {(size_t)f->npix * 4, (size_t)f->npix * 4, (size_t)f->npix * 4, (size_t)f->npix * 4,
                        tb ? tb : 16, 4};
  for (int k = 0; k < 6; ++k) {
    cudaError_t e = cudaMalloc(&p[k], sz[k]);
    if (e != cudaSuccess) {
      for (int j = 0; j < k; ++j) cudaFree(p[j]);
      return fail(RT_E_CUDA, "CUDA error %d (%s) allocating the hand-out order", (int)e, cudaGetErrorString(e));
    }
  }
  f->order = (uint32_t *)p[0]; f->sort_keys = (uint32_t *)p[1]; f->sort_keys_out = (uint32_t *)p[2];
  f->sort_iota = (uint32_t *)p[3]; f->sort_temp = p[4]; f->heavy = (uint32_t *)p[5];
  f->sort_temp_bytes = tb;
  return RT_OK;
}

int rt_reduce_enqueue(rt_frame *f, const rt_config *cfg) {
  int rc = check_cfg(cfg, f);
  if (rc) return rc;
  CU(cudaSetDevice(f->device));
  CU(launch_reduce_local(f->accum, f->npix, cfg->slice_begin, cfg->slice_end, f->partial,
                         f->sm_count, f->stream));
  // the ray counts of this frame order the work queue of the next one (heaviest pixels first)
  if ((rc = ensure_order(f)) != RT_OK) return rc;
  const uint64_t own = samples_of(cfg, 1);
  if (own) {
    const bool bvh = (cfg->flags & RT_FLAG_BVH) != 0 && f->last_n > 0;
    CU(launch_cost_order(f->partial, f->npix, 1.0f / (float)own, f->sort_keys, f->sort_keys_out,
                         f->sort_iota, f->order, f->heavy, bvh, f->sort_temp,
                         f->sort_temp_bytes, f->sm_count, f->stream));
    f->order_valid = true;
    f->order_heavy_only = bvh;
  }
  return RT_OK;
}

extern "C" int rt_frame_reduce_local(rt_frame *f, const rt_config *cfg) {
  int rc = rt_reduce_enqueue(f, cfg);
  if (rc) return rc;
  if (!caller_stream(f)) CU(cudaStreamSynchronize(f->stream));
  return RT_OK;
}

// finalise kernel + device->host copies, all on the frame's stream
int rt_finalize_enqueue(rt_frame *f, const rt_config *cfg, const void *const *partials, uint32_t count,
                        uint8_t *rgb8_out, float *linear_out) {
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
  CU(launch_finalize(ptrs, count, f->npix, scale, !(cfg->flags & RT_FLAG_LINEAR_OUTPUT), f->rgb_words,
                     linear_out ? f->linear : nullptr,
                     f->counters + 1, f->sm_count, f->stream));
  CU(cudaEventRecord(f->ev[3], f->stream));
  if (rgb8_out)
    CU(cudaMemcpyAsync(rgb8_out, f->rgb_words, (size_t)f->npix * 3, cudaMemcpyDeviceToHost, f->stream));
  if (linear_out)
    CU(cudaMemcpyAsync(linear_out, f->linear, (size_t)f->npix * 3 * sizeof(float),
                       cudaMemcpyDeviceToHost, f->stream));
  CU(cudaMemcpyAsync(f->host_words + 1, f->counters + 1, sizeof(unsigned long long), cudaMemcpyDeviceToHost, f->stream));
  f->finalized = true;
  return RT_OK;
}

static int fill_finalize_stats(rt_frame *f, rt_stats *stats) {
  stats->rays = f->host_words[1];
  stats->sphere_tests = stats->rays * (uint64_t)f->last_n;
  CU(cudaEventElapsedTime(&stats->finalize_ms, f->ev[2], f->ev[3]));
  if (!f->rendered || cudaEventElapsedTime(&stats->frame_ms, f->ev[0], f->ev[3]) != cudaSuccess) {
    stats->frame_ms = 0.f;  // rt_render has not run on this frame (or on another stream order)
    (void)cudaGetLastError();
  }
  return RT_OK;
}

extern "C" int rt_finalize(rt_frame *f, const rt_config *cfg, const void *const *partials,
                           uint32_t count, uint8_t *rgb8_out, float *linear_out, rt_stats *stats) {
  int rc = rt_finalize_enqueue(f, cfg, partials, count, rgb8_out, linear_out);
  if (rc) return rc;
  if (stats) stats->launches += 1;
  if (caller_stream(f)) return RT_OK;  // asynchronous: outputs valid after the caller's sync
  CU(cudaStreamSynchronize(f->stream));
  if (stats) return fill_finalize_stats(f, stats);
  return RT_OK;
}

extern "C" int rt_frame_stats(rt_frame *f, rt_stats *stats) {
  if (!f || !stats) return fail(RT_E_INVALID, "NULL argument");
  CU(cudaSetDevice(f->device));
  memset(stats, 0, sizeof *stats);
  if (f->rendered) {
    CU(cudaEventSynchronize(f->ev[1]));
    CU(cudaStreamSynchronize(f->stream));
    if (f->host_words[2]) return fail(RT_E_CUDA, "a tensor-core MMA of the filter never completed (bounded wait expired)");
    stats->samples = f->last_samples;
    stats->launches = f->last_launches;
    stats->filter_unit = f->last_filter;
    CU(cudaEventElapsedTime(&stats->render_ms, f->ev[0], f->ev[1]));
  }
  if (f->finalized) {
    CU(cudaEventSynchronize(f->ev[3]));
    CU(cudaStreamSynchronize(f->stream));  // the ray-count copy follows the event
    stats->launches += 1;
    return fill_finalize_stats(f, stats);
  }
  return RT_OK;
}

// ---------------------------------------------------------------- whole frame on one GPU
// rt_render_image keeps one frame object per calling thread (the slice planes are 0.83 GB at
// 1200x675): allocated by the first call, re-used while size and device stay the same.
namespace {
struct FrameCache {
  rt_frame *f = nullptr;
  ~FrameCache() { rt_frame_destroy(f); }
};
thread_local FrameCache g_frame_cache;
}  // namespace

extern "C" int rt_release_cached_frames(void) {
  rt_frame_destroy(g_frame_cache.f);
  g_frame_cache.f = nullptr;
  return RT_OK;
}

extern "C" int rt_render_image(const rt_scene *sc, const rt_camera *cam, const rt_config *cfg_in,
                               uint8_t *rgb8_out, float *linear_out, rt_stats *stats) {
  if (!sc || !cam || !cfg_in) return fail(RT_E_INVALID, "NULL argument");
  rt_config cfg = *cfg_in;
  cfg.slice_begin = 0; cfg.slice_end = RT_SLICES;
  rt_frame *f = g_frame_cache.f;
  if (!f || f->w != cfg.width || f->h != cfg.height || f->device != cfg.device) {
    rt_release_cached_frames();
    int rc = rt_frame_create(cfg.width, cfg.height, cfg.device, &f);
    if (rc) return rc;
    g_frame_cache.f = f;
  }
  rt_stats st;
  memset(&st, 0, sizeof st);
  int rc = rt_render(sc, cam, &cfg, f, &st);
  if (!rc) rc = rt_frame_reduce_local(f, &cfg);
  if (!rc) {
    st.launches += 1;
    const void *parts[1] = {f->partial};
    rc = rt_finalize(f, &cfg, parts, 1, rgb8_out, linear_out, &st);
  }
  if (!rc && stats) *stats = st;
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
  cudaError_t e = cudaMalloc(&d_ids, npix * 4);
  if (e == cudaSuccess) e = cudaMalloc(&d_t, npix * 8);
  if (e == cudaSuccess) e = cudaMalloc(&d_n, npix * 24);
  if (e == cudaSuccess) e = cudaMalloc(&d_ff, npix);
  if (e == cudaSuccess) {
    CamFrame cf;
    rt_derive_camera(cam, &cf);
    if (use_f64) {
      CameraF64 c64;
      for (int k = 0; k < 3; ++k) {
        c64.origin[k] = cf.origin[k]; c64.lower_left[k] = cf.llc[k];
        c64.horizontal[k] = cf.hor[k]; c64.vertical[k] = cf.ver[k];
      }
      e = launch_primary_f64(sc->exact, sc->n, c64, W, H, d_ids, d_t, d_n, d_ff, 0);
    } else {
      e = launch_primary_f32(rt_scene_view(sc), rt_narrow_camera(cf, W, H, false), W, H, d_ids, d_t, d_n, d_ff, 0);
    }
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
  int sms = 0;
  CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  const int blocks = sms * 8, iters = 20000;
  float *out = nullptr;
  long long *cyc = nullptr, hcyc = 0;
  cudaEvent_t a = nullptr, b = nullptr;
  float best = 0.f, best_ms = 0.f;
  cudaError_t e = cudaMalloc(&out, (size_t)blocks * 256 * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&cyc, sizeof(long long));
  if (e == cudaSuccess) e = cudaEventCreate(&a);
  if (e == cudaSuccess) e = cudaEventCreate(&b);
  for (int rep = 0; rep < 4 && e == cudaSuccess; ++rep) {  // first rep is the warm-up
    e = cudaEventRecord(a, 0);
    if (e == cudaSuccess) e = launch_ffma_peak(out, blocks, iters, cyc, 0);
    if (e == cudaSuccess) e = cudaEventRecord(b, 0);
    if (e == cudaSuccess) e = cudaEventSynchronize(b);
    float ms = 0.f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, a, b);
    if (e != cudaSuccess) break;
    double flops = 2.0 * 64.0 * (double)iters * 256.0 * blocks;
    float tf = (float)(flops / (ms * 1e-3) / 1e12);
    if (rep > 0 && tf > best) { best = tf; best_ms = ms; }
  }
  if (e == cudaSuccess) e = cudaMemcpy(&hcyc, cyc, sizeof hcyc, cudaMemcpyDeviceToHost);
  cudaFree(out); cudaFree(cyc);
  if (a) cudaEventDestroy(a);
  if (b) cudaEventDestroy(b);
  if (e != cudaSuccess) return fail(RT_E_CUDA, "FFMA peak probe failed: %s", cudaGetErrorString(e));
  *tflops = best;
  // block 0's cycle count over the kernel's wall time ~ SM clock (block 0 runs the whole time
  // only if there is a single wave: 8 blocks/SM x 256 threads fits, so it does)
  if (sm_clock_mhz_est) *sm_clock_mhz_est = best_ms > 0 ? (float)(hcyc / (best_ms * 1e3)) : 0.f;
  return RT_OK;
}
