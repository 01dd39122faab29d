// csrc/rt_kernels.cu — hand-written sm_100a kernels of the path-tracing hot path.
//
//   path_kernel       persistent megakernel: camera ray -> brute-force closest hit -> scatter
//                     -> ... -> sky, with per-lane path regeneration.  Replaces the loops of
//                     src/cpu/main.cc:111-123 and everything they call.
//   reduce_local      canonical pairwise tree over the slices one GPU rendered.
//   finalize_kernel   upper tree levels over per-GPU partials + write_color (color.h:8-23).
//   primary_f32       production-precision primary-hit table (accuracy gate, f32 leg).
//
// Design notes (full text in DESIGN.md):
//  * Work item = (pixel, sample slice); a lane owns R items at a time and sums its samples
//    sequentially in f32, so the image is independent of scheduling and of the GPU count.
//  * The sphere loop is a conservative f32 LINE-sphere filter, 7 FFMA + 1 FSETP per
//    (ray, sphere), on SoA float4 rows staged in shared memory by a 1-D bulk (TMA) copy.
//    Survivors (a handful per ray) are pushed to a per-lane candidate list and resolved in
//    f64 with exactly the reference's sphere::hit arithmetic, so closest-hit decisions do
//    not suffer f32 cancellation (r = 1000 ground sphere).
#include "rt_device.cuh"
#include "rt_kernels.h"

namespace rtb {

constexpr int kBlock = 256;
constexpr int kCandCap = 8;  // candidate list capacity per ray before an in-loop flush

struct Hit {
  double t;
  int id;
};

// Resolve a lane's candidate list (ascending sphere index) against the exact spheres.
__device__ __noinline__ Hit resolve_candidates(const uint32_t *list, uint32_t stride,
                                               uint32_t cnt, const double4 *exact, float ox,
                                               float oy, float oz, float dx, float dy, float dz,
                                               double t_min, Hit h) {
  for (uint32_t k = 0; k < cnt; ++k) {
    uint32_t s = list[k * stride];
    double4 S = exact[s];
    if (exact_hit(S, ox, oy, oz, dx, dy, dz, t_min, h.t)) h.id = (int)s;
  }
  return h;
}

template <int R>
struct Lane {
  float ox[R], oy[R], oz[R];     // ray origin
  float dx[R], dy[R], dz[R];     // unit direction
  float dlen[R];                 // |direction| of the reference's un-normalised ray
  float tr[R], tg[R], tb[R];     // path throughput (product of attenuations)
  float ar[R], ag[R], ab[R];     // item accumulator (sum over the slice's samples)
  float nrays[R];                // ray segments traced for this item
  uint32_t pixel[R], slice[R];
  uint32_t sample[R], sample_end[R];  // next sample / end of slice (relative to frame)
  uint32_t cur[R];               // sample index of the path in flight
  uint32_t slot[R];              // Philox slot (hit count of the path)
  int depth[R];                  // remaining bounces (src/cpu/main.cc:16)
  bool need_new[R], has_item[R], dead[R];
};

template <int R, bool SMEM>
__global__ void __launch_bounds__(kBlock, (R <= 2 ? 2 : 1))
path_kernel(const __grid_constant__ PathParams P) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const uint32_t n = P.scene.n;
  const uint32_t tid = threadIdx.x;

  // ---- shared memory carve-up: [fast float4 n][exact double4 n] (SMEM only) [cand lists]
  float4 *s_fast = reinterpret_cast<float4 *>(smem_raw);
  double4 *s_exact = reinterpret_cast<double4 *>(smem_raw + (SMEM ? (size_t)n * 16 : 0));
  uint32_t *s_cand =
      reinterpret_cast<uint32_t *>(smem_raw + (SMEM ? (size_t)n * 48 : 0));
  __shared__ __align__(8) uint64_t s_bar;

  const float4 *fast = SMEM ? s_fast : P.scene.fast;
  const double4 *exact = SMEM ? s_exact : P.scene.exact;

  if (SMEM) {
    // Scene upload into shared memory: one elected thread issues 1-D bulk (TMA) copies that
    // complete on an mbarrier; everyone waits on the barrier phase.
    if (tid == 0) {
      mbar_init(&s_bar, 1);
    }
    __syncthreads();
    if (tid == 0) {
      mbar_expect_tx(&s_bar, n * 48u);
      const uint32_t kChunk = 32768u;  // keep each bulk copy modest
      for (uint32_t off = 0; off < n * 16u; off += kChunk) {
        uint32_t bytes = min(kChunk, n * 16u - off);
        bulk_g2s(reinterpret_cast<unsigned char *>(s_fast) + off,
                 reinterpret_cast<const unsigned char *>(P.scene.fast) + off, bytes, &s_bar);
      }
      for (uint32_t off = 0; off < n * 32u; off += kChunk) {
        uint32_t bytes = min(kChunk, n * 32u - off);
        bulk_g2s(reinterpret_cast<unsigned char *>(s_exact) + off,
                 reinterpret_cast<const unsigned char *>(P.scene.exact) + off, bytes, &s_bar);
      }
    }
    mbar_wait(&s_bar, 0);
  }

  const uint32_t npix = P.width * P.height;
  const uint32_t nslices = P.slice_end - P.slice_begin;
  const uint64_t total_items = (uint64_t)npix * nslices;
  uint32_t *my_cand = s_cand + tid;  // element (r, k) at my_cand[(r*kCandCap + k) * kBlock]

  Lane<R> L;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    L.need_new[r] = true; L.has_item[r] = false; L.dead[r] = false;
    L.sample[r] = 0; L.sample_end[r] = 0;
    L.ox[r] = L.oy[r] = L.oz[r] = 0.f; L.dx[r] = 0.f; L.dy[r] = 1.f; L.dz[r] = 0.f;
    L.dlen[r] = 1.f; L.tr[r] = L.tg[r] = L.tb[r] = 0.f; L.ar[r] = L.ag[r] = L.ab[r] = 0.f;
    L.nrays[r] = 0.f; L.pixel[r] = 0; L.slice[r] = 0; L.cur[r] = 0; L.slot[r] = 0;
    L.depth[r] = 0;
  }

  for (;;) {
    // ------------------------------------------------------------ (1) regenerate paths
    bool all_dead = true;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (L.need_new[r] && !L.dead[r]) {
        while (L.sample[r] == L.sample_end[r]) {  // item exhausted: flush + fetch the next
          if (L.has_item[r]) {
            P.accum[(size_t)L.slice[r] * npix + L.pixel[r]] =
                make_float4(L.ar[r], L.ag[r], L.ab[r], L.nrays[r]);
            L.has_item[r] = false;
          }
          unsigned long long it = atomicAdd(P.counter, 1ull);
          if (it >= total_items) { L.dead[r] = true; break; }
          uint32_t sl = (uint32_t)(it / npix);
          L.pixel[r] = (uint32_t)(it - (uint64_t)sl * npix);
          L.slice[r] = P.slice_begin + sl;
          // canonical slice k covers samples [floor(k*spp/8), floor((k+1)*spp/8))
          L.sample[r] = (uint32_t)(((uint64_t)L.slice[r] * P.spp) / RT_SLICES_DEV);
          L.sample_end[r] = (uint32_t)(((uint64_t)(L.slice[r] + 1) * P.spp) / RT_SLICES_DEV);
          L.ar[r] = L.ag[r] = L.ab[r] = 0.f; L.nrays[r] = 0.f;
          L.has_item[r] = true;
        }
        if (!L.dead[r]) {
          // camera::get_ray (src/cpu/camera.h:28-35) with pixel jitter (main.cc:116-117)
          L.cur[r] = L.sample[r]++;
          uint32_t px = L.pixel[r];
          uint32_t row = px / P.width, i = px - row * P.width;
          uint32_t j = P.height - 1 - row;
          float4 xi = philox_uniforms(px, P.sample_offset + L.cur[r], 0u, P.seed);
          float u = ((float)i + xi.x) * P.cam.inv_wm1;
          float v = ((float)j + xi.y) * P.cam.inv_hm1;
          float rad = P.cam.lens_radius * sqrtf(xi.z), sn, cs;
          sincospif(2.0f * xi.w, &sn, &cs);
          float rdx = rad * cs, rdy = rad * sn;
          float offx = P.cam.u[0] * rdx + P.cam.v[0] * rdy;
          float offy = P.cam.u[1] * rdx + P.cam.v[1] * rdy;
          float offz = P.cam.u[2] * rdx + P.cam.v[2] * rdy;
          L.ox[r] = P.cam.origin[0] + offx;
          L.oy[r] = P.cam.origin[1] + offy;
          L.oz[r] = P.cam.origin[2] + offz;
          float ddx = P.cam.llc_minus_origin[0] + u * P.cam.horizontal[0] +
                      v * P.cam.vertical[0] - offx;
          float ddy = P.cam.llc_minus_origin[1] + u * P.cam.horizontal[1] +
                      v * P.cam.vertical[1] - offy;
          float ddz = P.cam.llc_minus_origin[2] + u * P.cam.horizontal[2] +
                      v * P.cam.vertical[2] - offz;
          float len = sqrtf(ddx * ddx + ddy * ddy + ddz * ddz), inv = 1.0f / len;
          L.dx[r] = ddx * inv; L.dy[r] = ddy * inv; L.dz[r] = ddz * inv; L.dlen[r] = len;
          L.tr[r] = L.tg[r] = L.tb[r] = 1.0f;
          L.depth[r] = (int)P.max_depth;
          L.slot[r] = 0;
          L.need_new[r] = false;
        }
      }
      all_dead = all_dead && L.dead[r];
    }
    if (all_dead) break;

    // ------------------------------------------------------------ (2) closest hit
    // hittable_list::hit (src/cpu/hittable_list.h:28-43) over all n spheres in order.
    float nod[R], ox2[R], oy2[R], oz2[R], oot[R];
    uint32_t cnt[R];
    Hit hit[R];
    double tmin[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      ox2[r] = 2.0f * L.ox[r]; oy2[r] = 2.0f * L.oy[r]; oz2[r] = 2.0f * L.oz[r];
      nod[r] = -(L.ox[r] * L.dx[r] + L.oy[r] * L.dy[r] + L.oz[r] * L.dz[r]);
      float oo = L.ox[r] * L.ox[r] + L.oy[r] * L.oy[r] + L.oz[r] * L.oz[r];
      // dead / idle lanes can never pass the filter
      oot[r] = (L.dead[r] || L.need_new[r]) ? __int_as_float(0x7f800000)
                                            : oo * (1.0f - RTB_FILTER_SLACK);
      cnt[r] = 0;
      hit[r].t = __longlong_as_double(0x7ff0000000000000ll);  // t_max = infinity
      hit[r].id = -1;
      tmin[r] = 0.001 * (double)L.dlen[r];  // main.cc:19 t_min, in unit-direction units
      L.nrays[r] += (L.dead[r] || L.need_new[r]) ? 0.f : 1.f;
    }

#pragma unroll 4
    for (uint32_t s = 0; s < n; ++s) {
      const float4 S = fast[s];
      bool h[R], any = false;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        // b = c.d - o.d ; q = 2 c.o + (r^2 - |c|^2 + slack) ; line hits iff b^2 + q >= |o|^2
        float b = fmaf(S.x, L.dx[r], fmaf(S.y, L.dy[r], fmaf(S.z, L.dz[r], nod[r])));
        float q = fmaf(S.x, ox2[r], fmaf(S.y, oy2[r], fmaf(S.z, oz2[r], S.w)));
        float e = fmaf(b, b, q);
        h[r] = e >= oot[r];
        any = any || h[r];
      }
      if (any) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          if (h[r]) {
            my_cand[(r * kCandCap + cnt[r]) * kBlock] = s;
            if (++cnt[r] == kCandCap) {
              hit[r] = resolve_candidates(my_cand + r * kCandCap * kBlock, kBlock, cnt[r], exact,
                                          L.ox[r], L.oy[r], L.oz[r], L.dx[r], L.dy[r], L.dz[r],
                                          tmin[r], hit[r]);
              cnt[r] = 0;
            }
          }
        }
      }
    }

    // ------------------------------------------------------------ (3) resolve + shade
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (L.dead[r] || L.need_new[r]) continue;
      if (cnt[r])
        hit[r] = resolve_candidates(my_cand + r * kCandCap * kBlock, kBlock, cnt[r], exact,
                                    L.ox[r], L.oy[r], L.oz[r], L.dx[r], L.dy[r], L.dz[r], tmin[r],
                                    hit[r]);
      if (hit[r].id < 0) {
        // sky, src/cpu/main.cc:27-29 (direction already unit length)
        float t = 0.5f * (L.dy[r] + 1.0f);
        float w = 1.0f - t;
        L.ar[r] += L.tr[r] * (w + t * 0.5f);
        L.ag[r] += L.tg[r] * (w + t * 0.7f);
        L.ab[r] += L.tb[r] * (w + t * 1.0f);
        L.need_new[r] = true;
        continue;
      }
      // hit record, src/cpu/sphere.h:43-48 + hittable.h:16-19, in f64 then narrowed
      const int id = hit[r].id;
      const double4 S = exact[id];
      const double t = hit[r].t;
      double px = (double)L.ox[r] + t * (double)L.dx[r];
      double py = (double)L.oy[r] + t * (double)L.dy[r];
      double pz = (double)L.oz[r] + t * (double)L.dz[r];
      double ir = 1.0 / S.w;
      double onx = (px - S.x) * ir, ony = (py - S.y) * ir, onz = (pz - S.z) * ir;
      bool front = ((double)L.dx[r] * onx + (double)L.dy[r] * ony + (double)L.dz[r] * onz) < 0.0;
      float nx = (float)(front ? onx : -onx), ny = (float)(front ? ony : -ony),
            nz = (float)(front ? onz : -onz);
      L.ox[r] = (float)px; L.oy[r] = (float)py; L.oz[r] = (float)pz;

      L.slot[r] += 1;
      const float4 xi = philox_uniforms(L.pixel[r], P.sample_offset + L.cur[r], L.slot[r], P.seed);
      const float4 M = __ldg(&P.scene.mat[id]);
      const uint32_t kind = __ldg(&P.scene.kind[id]);
      const float dx = L.dx[r], dy = L.dy[r], dz = L.dz[r];
      float sx, sy, sz;  // scattered direction (un-normalised, as in the reference)
      bool alive = true;
      if (kind == 0u) {
        // lambertian::scatter, src/cpu/material.h:19-32
        float ux, uy, uz;
        unit_vector_from(xi.x, xi.y, ux, uy, uz);
        sx = nx + ux; sy = ny + uy; sz = nz + uz;
        if (fabsf(sx) < 1e-8f && fabsf(sy) < 1e-8f && fabsf(sz) < 1e-8f) { sx = nx; sy = ny; sz = nz; }
        L.tr[r] *= M.x; L.tg[r] *= M.y; L.tb[r] *= M.z;
      } else if (kind == 1u) {
        // metal::scatter, src/cpu/material.h:40-46 (fuzz point drawn INSIDE the unit ball)
        float dn = dx * nx + dy * ny + dz * nz;
        sx = dx - 2.0f * dn * nx; sy = dy - 2.0f * dn * ny; sz = dz - 2.0f * dn * nz;
        if (M.w > 0.0f) {
          float ux, uy, uz;
          unit_vector_from(xi.x, xi.y, ux, uy, uz);
          float k = M.w * cbrtf(xi.z);
          sx += k * ux; sy += k * uy; sz += k * uz;
        }
        alive = (sx * nx + sy * ny + sz * nz) > 0.0f;
        L.tr[r] *= M.x; L.tg[r] *= M.y; L.tb[r] *= M.z;
      } else {
        // dielectric::scatter + reflectance + refract, src/cpu/material.h:57-87, vec3.h:126-131
        float ratio = front ? (1.0f / M.w) : M.w;
        float dn = dx * nx + dy * ny + dz * nz;
        float cos_t = fminf(-dn, 1.0f);
        float sin_t = sqrtf(fmaxf(0.0f, 1.0f - cos_t * cos_t));
        bool reflect_it = ratio * sin_t > 1.0f;
        if (!reflect_it) {
          float r0 = (1.0f - ratio) / (1.0f + ratio);
          r0 *= r0;
          float x = 1.0f - cos_t, x2 = x * x;
          reflect_it = (r0 + (1.0f - r0) * (x2 * x2 * x)) > xi.x;
        }
        if (reflect_it) {
          sx = dx - 2.0f * dn * nx; sy = dy - 2.0f * dn * ny; sz = dz - 2.0f * dn * nz;
        } else {
          float ex = ratio * (dx + cos_t * nx), ey = ratio * (dy + cos_t * ny),
                ez = ratio * (dz + cos_t * nz);
          float k = -sqrtf(fabsf(1.0f - (ex * ex + ey * ey + ez * ez)));
          sx = ex + k * nx; sy = ey + k * ny; sz = ez + k * nz;
        }
      }
      L.depth[r] -= 1;
      if (!alive || L.depth[r] <= 0) {  // absorbed (material.h:45) or bounce limit (main.cc:16)
        L.need_new[r] = true;
        continue;
      }
      float len = sqrtf(sx * sx + sy * sy + sz * sz), inv = 1.0f / len;
      L.dx[r] = sx * inv; L.dy[r] = sy * inv; L.dz[r] = sz * inv; L.dlen[r] = len;
    }
  }
}

// ---------------------------------------------------------------- slice tree + finalise
// Canonical pairwise tree over RT_SLICES = 8 slices: ((s0+s1)+(s2+s3))+((s4+s5)+(s6+s7)).
// A GPU that rendered an aligned power-of-two block of slices sums exactly its sub-tree;
// the finalise kernel adds the remaining upper levels over the per-GPU partials in rank
// order.  Every fp32 addition is therefore the same, in the same order, for 1/2/4/8 GPUs.
__device__ __forceinline__ float4 add4(float4 a, float4 b) {
  return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}

template <int N>
__device__ __forceinline__ float4 tree_sum(const float4 *const *src, size_t i) {
  if constexpr (N == 1) {
    return __ldcs(src[0] + i);
  } else {
    return add4(tree_sum<N / 2>(src, i), tree_sum<N / 2>(src + N / 2, i));
  }
}

struct PtrPack {
  const float4 *p[8];
};

template <int N>
__global__ void __launch_bounds__(256) reduce_local_kernel(const __grid_constant__ PtrPack src,
                                                           float4 *__restrict__ dst,
                                                           uint32_t npix) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < npix;
       i += (size_t)gridDim.x * blockDim.x)
    dst[i] = tree_sum<N>(src.p, i);
}

// write_color, src/cpu/color.h:8-23: x = sqrt(sum/spp); int(256 * clamp(x, 0, 0.999)).
__device__ __forceinline__ uint32_t quantise(float sum, float scale) {
  float x = sqrtf(scale * sum);
  x = fminf(fmaxf(x, 0.0f), 0.999f);
  return (uint32_t)(int)(256.0f * x);
}

// One thread = 4 consecutive pixels = 12 output bytes = three 32-bit stores (coalesced: a
// warp writes 384 contiguous bytes).  Ray counts (w channel) are block-reduced to one atomic.
template <int N>
__global__ void __launch_bounds__(256) finalize_kernel(const __grid_constant__ PtrPack src,
                                                       uint32_t npix, float scale,
                                                       uint32_t *__restrict__ rgb_words,
                                                       float *__restrict__ linear,
                                                       unsigned long long *__restrict__ rays) {
  const uint32_t quads = (npix + 3) / 4;
  double my_rays = 0.0;
  for (uint32_t qd = blockIdx.x * blockDim.x + threadIdx.x; qd < quads;
       qd += gridDim.x * blockDim.x) {
    uint32_t c[12];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      uint32_t px = qd * 4 + k;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (px < npix) {
        v = tree_sum<N>(src.p, px);
        if (linear) {
          linear[(size_t)px * 3 + 0] = v.x; linear[(size_t)px * 3 + 1] = v.y;
          linear[(size_t)px * 3 + 2] = v.z;
        }
      }
      my_rays += (double)v.w;
      c[k * 3 + 0] = quantise(v.x, scale);
      c[k * 3 + 1] = quantise(v.y, scale);
      c[k * 3 + 2] = quantise(v.z, scale);
    }
    // rgb_words is padded to a multiple of 12 bytes by the host
    rgb_words[qd * 3 + 0] = c[0] | (c[1] << 8) | (c[2] << 16) | (c[3] << 24);
    rgb_words[qd * 3 + 1] = c[4] | (c[5] << 8) | (c[6] << 16) | (c[7] << 24);
    rgb_words[qd * 3 + 2] = c[8] | (c[9] << 8) | (c[10] << 16) | (c[11] << 24);
  }
  // block reduce of the ray count
  __shared__ double s_r[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) my_rays += __shfl_xor_sync(0xffffffffu, my_rays, o);
  if ((threadIdx.x & 31) == 0) s_r[threadIdx.x >> 5] = my_rays;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += s_r[k];
    atomicAdd(rays, (unsigned long long)(t + 0.5));
  }
}

// ---------------------------------------------------------------- primary hits (f32 leg)
// Pixel-centre (xi = 0.5), zero-lens rays through the PRODUCTION intersection: f32 ray,
// f32 filter, f64 candidate resolution.  Outputs are double for a uniform host interface.
__global__ void __launch_bounds__(128) primary_f32_kernel(const __grid_constant__ SceneView sc,
                                                          const __grid_constant__ CameraF32 cam,
                                                          uint32_t W, uint32_t H, int32_t *ids,
                                                          double *tt, double *normal,
                                                          uint8_t *ffo) {
  uint32_t px = blockIdx.x * blockDim.x + threadIdx.x;
  if (px >= W * H) return;
  uint32_t row = px / W, i = px - row * W, j = H - 1 - row;
  float u = ((float)i + 0.5f) * cam.inv_wm1, v = ((float)j + 0.5f) * cam.inv_hm1;
  float ox = cam.origin[0], oy = cam.origin[1], oz = cam.origin[2];
  float ddx = cam.llc_minus_origin[0] + u * cam.horizontal[0] + v * cam.vertical[0];
  float ddy = cam.llc_minus_origin[1] + u * cam.horizontal[1] + v * cam.vertical[1];
  float ddz = cam.llc_minus_origin[2] + u * cam.horizontal[2] + v * cam.vertical[2];
  float len = sqrtf(ddx * ddx + ddy * ddy + ddz * ddz), inv = 1.0f / len;
  float dx = ddx * inv, dy = ddy * inv, dz = ddz * inv;
  float nod = -(ox * dx + oy * dy + oz * dz);
  float oot = (ox * ox + oy * oy + oz * oz) * (1.0f - RTB_FILTER_SLACK);
  double tbest = __longlong_as_double(0x7ff0000000000000ll), tmin = 0.001 * (double)len;
  int id = -1;
  for (uint32_t s = 0; s < sc.n; ++s) {
    float4 S = __ldg(&sc.fast[s]);
    float b = fmaf(S.x, dx, fmaf(S.y, dy, fmaf(S.z, dz, nod)));
    float q = fmaf(S.x, 2.0f * ox, fmaf(S.y, 2.0f * oy, fmaf(S.z, 2.0f * oz, S.w)));
    if (fmaf(b, b, q) >= oot) {
      if (exact_hit(sc.exact[s], ox, oy, oz, dx, dy, dz, tmin, tbest)) id = (int)s;
    }
  }
  ids[px] = id;
  if (id < 0) {
    tt[px] = 0; ffo[px] = 0;
    normal[(size_t)px * 3] = normal[(size_t)px * 3 + 1] = normal[(size_t)px * 3 + 2] = 0;
    return;
  }
  double4 S = sc.exact[id];
  double qx = (double)ox + tbest * dx, qy = (double)oy + tbest * dy, qz = (double)oz + tbest * dz;
  double ir = 1.0 / S.w;
  double onx = (qx - S.x) * ir, ony = (qy - S.y) * ir, onz = (qz - S.z) * ir;
  bool front = (dx * onx + dy * ony + dz * onz) < 0.0;
  tt[px] = tbest / (double)len;  // back to the reference's un-normalised parameter
  ffo[px] = front;
  normal[(size_t)px * 3 + 0] = front ? onx : -onx;
  normal[(size_t)px * 3 + 1] = front ? ony : -ony;
  normal[(size_t)px * 3 + 2] = front ? onz : -onz;
}

// ---------------------------------------------------------------- FP32 peak micro-kernel
// 16 independent FFMA chains per thread, 3 register operands each (like the filter loop).
__global__ void __launch_bounds__(256) ffma_peak_kernel(float *out, int iters, float b, float c,
                                                        long long *cycles) {
  float a[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) a[k] = (float)(threadIdx.x + k);
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int k = 0; k < 16; ++k) a[k] = fmaf(a[k], b, c);
  }
  long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 16; ++k) s += a[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0;
}

cudaError_t launch_ffma_peak(float *out, int blocks, int iters, long long *cycles,
                             cudaStream_t st) {
  ffma_peak_kernel<<<blocks, 256, 0, st>>>(out, iters, 0.999f, 0.001f, cycles);
  return cudaGetLastError();
}

// ---------------------------------------------------------------- launchers
template <int R, bool SMEM>
static cudaError_t launch_path_t(const PathParams &p, int sm_count, size_t smem, cudaStream_t st,
                                 PathLaunchInfo *info) {
  auto kern = path_kernel<R, SMEM>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int per_sm = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kBlock, smem);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) return cudaErrorLaunchOutOfResources;
  // persistent grid: one full wave, a multiple of the SM count
  int grid = sm_count * per_sm;
  if (info) {
    info->rays_per_lane = R; info->grid = grid; info->block = kBlock;
    info->blocks_per_sm = per_sm; info->smem_bytes = smem; info->scene_in_smem = SMEM;
  }
  kern<<<grid, kBlock, smem, st>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_path(const PathParams &p, int rays_per_lane, int sm_count,
                        size_t smem_optin_max, cudaStream_t st, PathLaunchInfo *info) {
  int R = rays_per_lane ? rays_per_lane : 2;
  size_t cand = (size_t)kBlock * R * kCandCap * sizeof(uint32_t);
  size_t scene = (size_t)p.scene.n * 48;
  // keep room for >= 2 resident blocks when the scene is small
  bool in_smem = scene + cand + 1024 <= smem_optin_max;
  size_t smem = cand + (in_smem ? scene : 0);
#define RTB_DISPATCH(RR)                                                               \
  return in_smem ? launch_path_t<RR, true>(p, sm_count, smem, st, info)                \
                 : launch_path_t<RR, false>(p, sm_count, smem, st, info)
  switch (R) {
    case 1: RTB_DISPATCH(1);
    case 2: RTB_DISPATCH(2);
    case 4: RTB_DISPATCH(4);
    default: return cudaErrorInvalidValue;
  }
#undef RTB_DISPATCH
}

static int grid_for(uint32_t work_items, int sm_count, int per_sm = 8) {
  int need = (int)((work_items + 255) / 256);
  int cap = sm_count * per_sm;
  return need < cap ? (need > 0 ? need : 1) : cap;
}

cudaError_t launch_reduce_local(const float4 *accum, uint32_t npix, uint32_t slice_begin,
                                uint32_t slice_end, float4 *dst, int sm_count, cudaStream_t st) {
  uint32_t cnt = slice_end - slice_begin;
  PtrPack pk{};
  for (uint32_t k = 0; k < cnt && k < 8; ++k) pk.p[k] = accum + (size_t)(slice_begin + k) * npix;
  int grid = grid_for(npix, sm_count);
  switch (cnt) {
    case 1: reduce_local_kernel<1><<<grid, 256, 0, st>>>(pk, dst, npix); break;
    case 2: reduce_local_kernel<2><<<grid, 256, 0, st>>>(pk, dst, npix); break;
    case 4: reduce_local_kernel<4><<<grid, 256, 0, st>>>(pk, dst, npix); break;
    case 8: reduce_local_kernel<8><<<grid, 256, 0, st>>>(pk, dst, npix); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_finalize(const float4 *const *partials, uint32_t count, uint32_t npix,
                            float scale, uint32_t *rgb_words, float *linear,
                            unsigned long long *rays, int sm_count, cudaStream_t st) {
  PtrPack pk{};
  for (uint32_t k = 0; k < count && k < 8; ++k) pk.p[k] = partials[k];
  int grid = grid_for((npix + 3) / 4, sm_count);
  switch (count) {
    case 1: finalize_kernel<1><<<grid, 256, 0, st>>>(pk, npix, scale, rgb_words, linear, rays); break;
    case 2: finalize_kernel<2><<<grid, 256, 0, st>>>(pk, npix, scale, rgb_words, linear, rays); break;
    case 4: finalize_kernel<4><<<grid, 256, 0, st>>>(pk, npix, scale, rgb_words, linear, rays); break;
    case 8: finalize_kernel<8><<<grid, 256, 0, st>>>(pk, npix, scale, rgb_words, linear, rays); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_primary_f32(const SceneView &sc, const CameraF32 &cam, uint32_t W, uint32_t H,
                               int32_t *ids, double *t, double *normal, uint8_t *ff,
                               cudaStream_t st) {
  uint32_t npix = W * H;
  primary_f32_kernel<<<(npix + 127) / 128, 128, 0, st>>>(sc, cam, W, H, ids, t, normal, ff);
  return cudaGetLastError();
}

}  // namespace rtb
