// csrc/rt_wave.cu — the path-tracing hot path (sm_100a): everything executed per (pixel, sample)
// of src/cpu/main.cc:111-123.
//
// Path state lives in a pool of slots in device memory (L2-resident); two device functions
// alternate over it:
//
//   shade_slot     one slot.  Finishes the ray segment traced by the previous pass: f64 resolution
//                  of the filter's candidates (sphere::hit, src/cpu/sphere.h:24-51), sky or material
//                  scatter (main.cc:12-30, material.h), then, if the path ended, path regeneration:
//                  next sample of the slot's (pixel, slice) item or the next item of the frame's
//                  work queue, camera::get_ray (camera.h:28-35).
//   filter_rays    R slots of one lane.  Conservative f32 line-sphere test of R rays against ALL
//                  spheres in reference order (hittable_list::hit visits every object,
//                  hittable_list.h:28-43): packed FFMA2, sphere rows through the constant bank
//                  into uniform registers.  Survivors are recorded as {group, row mask} entries in
//                  the slot's candidate planes.  The lane carries no path state, only rays.
//   bvh_trace      optional replacement of filter_rays: closest hit through the host-built BVH.
//
// fused_kernel is the product: a persistent kernel in which every lane owns R slots for the whole
// frame and alternates the two phases.  shade_kernel / filter_kernel run the same functions as a
// two-kernel wavefront (A/B variant, RT_FLAG_WAVEFRONT).
#include "rt_device.cuh"
#include "rt_kernels.h"

#include <cstring>

namespace rtb {

#ifndef RTW_BLOCK
#define RTW_BLOCK 256
#endif
#ifndef RTW_MINB
#define RTW_MINB 3
#endif
constexpr int kWBlock = RTW_BLOCK;

__device__ __forceinline__ float2 wffma2(float2 a, float2 b, float2 c) {
  unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a),
                     rb = *reinterpret_cast<unsigned long long *>(&b),
                     rc = *reinterpret_cast<unsigned long long *>(&c), rd;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  return *reinterpret_cast<float2 *>(&rd);
}
__device__ __forceinline__ float2 wbc(float v) { return make_float2(v, v); }
#define WPK(arr, r) (((r)&1) ? (arr)[(r) >> 1].y : (arr)[(r) >> 1].x)

// ================================================================ filter
// R rays per lane (even), G sphere rows per decision group.
template <int R, int G, bool CROWS>
__device__ __forceinline__ void filter_rays(const WaveParams &P, const ConstRows &CR, const uint32_t base,
                                            const uint32_t stride) {
  static_assert(R % 2 == 0, "rays are packed in pairs");
  constexpr int NP = R / 2;
  const uint32_t npad = P.scene.n_pad;
  const float INF = __int_as_float(0x7f800000);
  {
    float2 DX[NP], DY[NP], DZ[NP], OX2[NP], OY2[NP], OZ2[NP], NOD[NP], OOT[NP];
    uint32_t cnt[R], off[R];  // survivors so far / element offset of the next candidate plane
#pragma unroll
    for (int r = 0; r < R; ++r) {
      off[r] = base + r * stride;
      const float4 o = P.ray_o[base + r * stride];  // {ox, oy, oz, |d| or < 0 = no ray}
      const float4 d = P.ray_d[base + r * stride];  // unit direction
      WPK(DX, r) = d.x; WPK(DY, r) = d.y; WPK(DZ, r) = d.z;
      WPK(OX2, r) = 2.0f * o.x; WPK(OY2, r) = 2.0f * o.y; WPK(OZ2, r) = 2.0f * o.z;
      WPK(NOD, r) = -(o.x * d.x + o.y * d.y + o.z * d.z);
      // idle slots never pass the filter
      WPK(OOT, r) = o.w > 0.0f ? (o.x * o.x + o.y * o.y + o.z * o.z) * (1.0f - RTB_FILTER_SLACK) : INF;
      cnt[r] = 0;
    }

    // e = (c.d - o.d)^2 + (2 c.o + w) for one group of G rows; every FFMA2 multiplies by exactly
    // one sphere scalar (a uniform register when CROWS)
    auto group = [&](const float4 (&S)[G], float2 (&e)[G][NP]) {
#pragma unroll
      for (int k = 0; k < G; ++k)
#pragma unroll
        for (int p = 0; p < NP; ++p) {
          float2 b = wffma2(wbc(S[k].x), DX[p], wffma2(wbc(S[k].y), DY[p], wffma2(wbc(S[k].z), DZ[p], NOD[p])));
          float2 e1 = wffma2(b, b, wbc(S[k].w));
          e[k][p] = wffma2(wbc(S[k].x), OX2[p], wffma2(wbc(S[k].y), OY2[p], wffma2(wbc(S[k].z), OZ2[p], e1)));
        }
    };
    // decision for one group: max over rows, one compare per ray, one vote, one branch
    uint32_t *const cand = P.cand_idx;
    const uint32_t nslots = P.n_slots;
    auto decide = [&](const float2 (&e)[G][NP], uint32_t s0) {
      float mx[NP], my[NP];
      bool any = false;
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        mx[p] = e[0][p].x; my[p] = e[0][p].y;
#pragma unroll
        for (int k = 1; k < G; ++k) { mx[p] = fmaxf(mx[p], e[k][p].x); my[p] = fmaxf(my[p], e[k][p].y); }
        any = any | (mx[p] >= OOT[p].x) | (my[p] >= OOT[p].y);  // no short-circuit: independent compares
      }
      if (__any_sync(0xffffffffu, any)) {
        // rare path, narrowed ray by ray: a ray whose group maximum passed records the group and
        // the mask of its rows that passed as ONE candidate entry {s0 << kCandShift | mask}
#pragma unroll
        for (int r = 0; r < R; ++r)
          if (((r & 1) ? my[r >> 1] : mx[r >> 1]) >= WPK(OOT, r)) {
            uint32_t m = 0;
#pragma unroll
            for (int k = 0; k < G; ++k) m |= WPK(e[k], r) >= WPK(OOT, r) ? (1u << k) : 0u;
            if (cnt[r] < kCandPlanes) cand[off[r]] = s0 << kCandShift | m;
            off[r] += nslots;
            cnt[r] += 1;
          }
      }
    };

    const float4 *frow = P.scene.fast;
    float4 S[G];
    float2 ea[G][NP], eb[G][NP];
#pragma unroll
    for (int k = 0; k < G; ++k) S[k] = CROWS ? CR.row[k] : __ldg(frow + k);
    group(S, ea);
#pragma unroll
    for (int k = 0; k < G; ++k) S[k] = CROWS ? CR.row[G + k] : __ldg(frow + G + k);
    uint32_t npad_r = npad;
    asm volatile("" : "+r"(npad_r));
    // software pipeline: the FFMA2s of group g+1 are issued before the decision on group g, so
    // the max/compare/vote/branch chain never sits at the end of a dependency tail.  Two groups
    // per trip (ea/eb ping-pong).  The row array carries 2G padding rows (w = -inf).
    for (uint32_t s0 = 0; s0 < npad_r; s0 += 2 * G) {
      group(S, eb);
#pragma unroll
      for (int k = 0; k < G; ++k) S[k] = CROWS ? CR.row[s0 + 2 * G + k] : __ldg(frow + s0 + 2 * G + k);
      decide(ea, s0);
      group(S, ea);
#pragma unroll
      for (int k = 0; k < G; ++k) S[k] = CROWS ? CR.row[s0 + 3 * G + k] : __ldg(frow + s0 + 3 * G + k);
      decide(eb, s0 + G);
    }
#pragma unroll
    for (int r = 0; r < R; ++r) P.cand_cnt[base + r * stride] = cnt[r];
  }
}

template <int R, int G, bool CROWS>
__global__ void __launch_bounds__(kWBlock, RTW_MINB)
filter_kernel(const __grid_constant__ WaveParams P, const __grid_constant__ ConstRows CR) {
  const uint32_t stride = P.n_slots / R;  // slot of ray r = base + r * stride
  // warp-uniform trip count (stride is a multiple of the block size): the sphere loop must stay
  // convergent so that its rows can live in uniform registers
  for (uint32_t blk = blockIdx.x; blk < stride / kWBlock; blk += gridDim.x)
    filter_rays<R, G, CROWS>(P, CR, blk * kWBlock + threadIdx.x, stride);
}

// ================================================================ shade
// sqrt of a non-negative double: rsqrt seed in f32, two Newton steps in f64 (no f64 division)
__device__ __forceinline__ double wave_dsqrt(double x) {
  float xf = fmaxf((float)x, 1e-30f);
  double y = (double)rsqrtf(xf);
  double h = 0.5 * y;
  double s = x * y;
  s = fma(fma(-s, s, x), h, s);
  s = fma(fma(-s, s, x), h, s);
  return s;
}

struct WRay {
  double ox, oy, oz, dx, dy, dz, a, tminA;
  double lo;  // smallest acceptable root numerator: tminA, or the next double above it when strict
};
__device__ __forceinline__ double next_up_pos(double x) {  // x > 0 finite
  return __longlong_as_double(__double_as_longlong(x) + 1ll);
}

// sphere::hit on numerators (root = N / a, a common to the ray): src/cpu/sphere.h:24-51 with the
// inclusive upper bound of hittable_list.h:34-39 (a later sphere with an equal root wins), or —
// strict — the v4 interval of src/gpu/sphere.h:31-36 (t_min < root < t_max: the earlier sphere
// wins).  Both are ONE pair of compares, N < lo || !(N < hi): the bounds carry the rule
// (lo = t_min or the next double above it; hi = the best root, or the next double above it when
// an equal root may replace it).
__device__ __forceinline__ bool wave_exact_hit(const double4 S, const WRay &r, double &bestN,
                                               double &hi, const bool strict) {
  double ocx = r.ox - S.x, ocy = r.oy - S.y, ocz = r.oz - S.z;
  double half_b = ocx * r.dx + ocy * r.dy + ocz * r.dz;
  double c = ocx * ocx + ocy * ocy + ocz * ocz - S.w * S.w;
  double disc = half_b * half_b - r.a * c;
  if (disc < 0.0) return false;
  double sq = wave_dsqrt(disc);
  double N = -half_b - sq;
  if (N < r.lo || !(N < hi)) {
    N = -half_b + sq;
    if (N < r.lo || !(N < hi)) return false;
  }
  bestN = N;
  hi = strict ? N : next_up_pos(N);
  return true;
}

// One slot, one wave.  (1) Finish the ray segment traced by the last filter pass: f64 resolution
// of its candidates, sky on a miss.  (2) ONE Philox block per slot and wave: the camera sample of
// a new path (miss / empty slot) or the scatter sample of the hit.  (3) camera::get_ray or
// material::scatter.  A path that ends at a hit (absorbed, bounce limit) leaves the slot without a
// ray for one wave and regenerates in the next one.  Returns 0 once the slot is retired.
template <bool BVH, bool F32HIT = false>
__device__ __forceinline__ uint32_t shade_slot(const WaveParams &P, const uint32_t i,
                                               const double4 *exact, const float4 *mat,
                                               const float2 *aux, const uint32_t npix,
                                               const uint64_t total_items) {
  float4 T = P.thr[i];  // {tr, tg, tb, depth}
  int depth = __float_as_int(T.w);
  if (depth < 0) return 0u;  // slot retired: the frame's work queue is empty
  uint4 st = P.it[i];   // {item = slice << kItemShift | pixel (0xffffffff = none), samp, samp_end, -}
  float4 A = P.acc[i];  // {sum r, sum g, sum b, ray segments} of the slot's item
  float4 RO = make_float4(0.f, 0.f, 0.f, -1.f), RD = make_float4(0.f, 1.f, 0.f, 0.f);
  int id = -1;
  double hpx = 0, hpy = 0, hpz = 0;  // hit point (f64)

  if (depth > 0) {
    // ---------------------------------------------------------- resolve
    RO = P.ray_o[i];
    RD = P.ray_d[i];
    A.w += 1.0f;  // one ray_color invocation (src/cpu/main.cc:12)
    WRay ray;
    ray.ox = RO.x; ray.oy = RO.y; ray.oz = RO.z; ray.dx = RD.x; ray.dy = RD.y; ray.dz = RD.z;
    ray.a = ray.dx * ray.dx + ray.dy * ray.dy + ray.dz * ray.dz;
    ray.tminA = 0.001 * (double)RO.w * ray.a;  // main.cc:19 t_min in unit-direction units, times a
    double bestN = __longlong_as_double(0x7ff0000000000000ll);
    double hi = bestN;
    if (BVH) {
      // the traversal already resolved the closest hit: {id, numerator of the root}
      id = (int)P.cand_idx[i];
      bestN = __hiloint2double((int)P.cand_idx[(size_t)2 * P.n_slots + i],
                               (int)P.cand_idx[(size_t)1 * P.n_slots + i]);
    } else if (F32HIT) {
      // RT_FLAG_F32_HIT: sphere::hit in f32 on the reference's un-normalised direction, op for op
      // src/gpu/sphere.h:15-47 (no self-skip, f32 hit point) -- reproduces the cancellation of
      // |oc|^2 - r^2 on the r = 1000 ground sphere that the reference's CUDA program shows
      const uint32_t cnt = P.cand_cnt[i];
      const bool strict = P.model_v4 != 0u;
      const float ddx = RD.x * RO.w, ddy = RD.y * RO.w, ddz = RD.z * RO.w;
      const float a = ddx * ddx + ddy * ddy + ddz * ddz;
      float bestT = __int_as_float(0x7f800000);
      auto hit32 = [&](const uint32_t s) {
        const double4 S = exact[s];
        const float r = (float)S.w;
        const float ocx = RO.x - (float)S.x, ocy = RO.y - (float)S.y, ocz = RO.z - (float)S.z;
        const float hb = ocx * ddx + ocy * ddy + ocz * ddz;
        const float c = ocx * ocx + ocy * ocy + ocz * ocz - r * r;
        const float disc = hb * hb - a * c;
        if (disc < 0.0f) return;
        const float sq = __fsqrt_rn(disc);  // IEEE like the reference's build (this TU uses approximate / and sqrt)
        float root = __fdiv_rn(-hb - sq, a);
        if (strict ? !(0.001f < root && root < bestT) : (root < 0.001f || bestT < root)) {
          root = __fdiv_rn(-hb + sq, a);
          if (strict ? !(0.001f < root && root < bestT) : (root < 0.001f || bestT < root)) return;
        }
        bestT = root;
        id = (int)s;
      };
      if (cnt <= kCandPlanes) {
        uint32_t k = 0, m = 0, s0 = 0;
        for (;;) {
          if (m == 0u) {
            if (k == cnt) break;
            const uint32_t e = P.cand_idx[(size_t)k * P.n_slots + i];
            ++k;
            s0 = e >> kCandShift;
            m = e & ((1u << kCandShift) - 1u);
          }
          hit32(s0 + (uint32_t)__ffs((int)m) - 1u);
          m &= m - 1u;
        }
      } else {
        for (uint32_t s = 0; s < P.scene.n; ++s) hit32(s);
      }
      if (id >= 0) {  // ray::at in f32 (src/gpu/ray.h), kept as the slot's f64 hit point
        hpx = (double)fmaf(bestT, ddx, RO.x); hpy = (double)fmaf(bestT, ddy, RO.y); hpz = (double)fmaf(bestT, ddz, RO.z);
      }
    } else {
      const uint32_t cnt = P.cand_cnt[i];
      const bool strict = P.model_v4 != 0u;
      ray.lo = strict ? next_up_pos(ray.tminA) : ray.tminA;
      const uint32_t skip = (uint32_t)__float_as_int(RD.w);  // the sphere the ray just left, or -1
      if (cnt <= kCandPlanes) {
        // entries {first row of a group << kCandShift | mask of its rows that passed}, ascending;
        // one exact test per trip so that lanes with different entry shapes stay together
        uint32_t k = 0, m = 0, s0 = 0;
        for (;;) {
          if (m == 0u) {
            if (k == cnt) break;
            const uint32_t e = P.cand_idx[(size_t)k * P.n_slots + i];
            ++k;
            s0 = e >> kCandShift;
            m = e & ((1u << kCandShift) - 1u);
          }
          const uint32_t s = s0 + (uint32_t)__ffs((int)m) - 1u;
          m &= m - 1u;
          if (s != skip && wave_exact_hit(exact[s], ray, bestN, hi, strict)) id = (int)s;
        }
      } else {
        // candidate planes overflowed (a ray along a long row of spheres): scan every sphere, the
        // same conservative f32 test first so that f64 work is spent on its survivors only
        const float nod = -(RO.x * RD.x + RO.y * RD.y + RO.z * RD.z);
        const float ox2 = 2.0f * RO.x, oy2 = 2.0f * RO.y, oz2 = 2.0f * RO.z;
        const float oot = (RO.x * RO.x + RO.y * RO.y + RO.z * RO.z) * (1.0f - RTB_FILTER_SLACK);
        const float4 *fast = P.scene.fast;
        for (uint32_t s = 0; s < P.scene.n; ++s) {
          const float4 S = __ldg(fast + s);
          const float b = fmaf(S.x, RD.x, fmaf(S.y, RD.y, fmaf(S.z, RD.z, nod)));
          const float e = fmaf(S.x, ox2, fmaf(S.y, oy2, fmaf(S.z, oz2, fmaf(b, b, S.w))));
          // RT_FLAG_NO_FILTER: no pre-test at all -- hittable_list::hit as written (hittable_list.h:28-43)
          if ((e >= oot || P.no_filter) && s != skip && wave_exact_hit(exact[s], ray, bestN, hi, strict)) id = (int)s;
        }
      }
    }
    if (id < 0) {
      // sky, src/cpu/main.cc:27-29 (direction already unit length)
      float t = 0.5f * (RD.y + 1.0f);
      float w = 1.0f - t;
      A.x += T.x * (w + t * 0.5f);
      A.y += T.y * (w + t * 0.7f);
      A.z += T.z * (w + t * 1.0f);
      depth = 0;
    } else if (!F32HIT) {
      const double t = bestN * (2.0 - ray.a);  // 1/a = 2 - a to within eps^2 (a = 1 + eps)
      hpx = ray.ox + t * ray.dx; hpy = ray.oy + t * ray.dy; hpz = ray.oz + t * ray.dz;
    }
  }

  const bool regen = depth == 0;
  if (regen) {
    // ------------------------------------------------------------ next sample / next item
    while (st.y == st.z) {  // item exhausted: flush it and fetch the next
      if (st.x != 0xffffffffu) {
        P.accum[(size_t)(st.x >> kItemShift) * npix + (st.x & kItemPixelMask)] = A;
        st.x = 0xffffffffu;
      }
      // warp-aggregated pop: the lanes that need an item right now share ONE atomic (single-address
      // atomics top out near 1 G/s, and a frame ends on one-sample items)
      const unsigned am = __activemask();
      const unsigned lane = threadIdx.x & 31u;
      const int leader = __ffs((int)am) - 1;
      unsigned long long itq = 0;
      if ((int)lane == leader) itq = atomicAdd(P.counter, (unsigned long long)__popc(am));
      itq = __shfl_sync(am, itq, leader) + (unsigned long long)__popc(am & ((1u << lane) - 1u));
      if (itq >= total_items) { depth = -1; break; }
      // hand-out order: first every slice of the heavy pixels, then the other pixels slice by
      // slice (largest slices first, heaviest pixels first within a slice)
      uint32_t ord, px;
      if (total_items <= 0xffffffffull) {
        // 32-bit arithmetic (a 64-bit division is ~100 instructions at a few active lanes)
        const uint32_t q32 = (uint32_t)itq;
        if (P.order) {
          const uint32_t nh = __ldg(P.heavy);
          const uint32_t heavy_items = nh * P.n_orders;
          if (q32 < heavy_items) {
            ord = q32 / nh;
            px = __ldg(P.order + (q32 - ord * nh));
          } else {
            const uint32_t q = q32 - heavy_items, nl = npix - nh;
            ord = q / nl;
            px = __ldg(P.order + nh + (q - ord * nl));
          }
        } else {
          ord = q32 / npix;
          px = q32 - ord * npix;
        }
      } else if (P.order) {
        const uint32_t nh = __ldg(P.heavy);
        const uint64_t heavy_items = (uint64_t)nh * P.n_orders;
        if (itq < heavy_items) {
          ord = (uint32_t)(itq / nh);
          px = __ldg(P.order + (uint32_t)(itq - (uint64_t)ord * nh));
        } else {
          const uint64_t q = itq - heavy_items;
          const uint32_t nl = npix - nh;
          ord = (uint32_t)(q / nl);
          px = __ldg(P.order + nh + (uint32_t)(q - (uint64_t)ord * nl));
        }
      } else {
        ord = (uint32_t)(itq / npix);
        px = (uint32_t)(itq - (uint64_t)ord * npix);
      }
      const uint32_t sl = slice_in_order(ord, P.slice_begin, P.slice_end);
      st.x = sl << kItemShift | px;
      slice_range(sl, P.spp, st.y, st.z);
      A = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }

  if (depth >= 0) {
    // ------------------------------------------------------------ the wave's random numbers
    // stream: key = (pixel, sample), counter slot 0 = camera, k >= 1 = scatter at the k-th hit
    const uint32_t px = st.x & kItemPixelMask;
    const uint32_t hitno = P.max_depth - (uint32_t)depth + 1u;
    const float4 xi = philox_uniforms(px, P.sample_offset + (regen ? st.y : st.y - 1u),
                                      regen ? 0u : hitno, P.seed);
    if (regen) {
      // camera::get_ray (src/cpu/camera.h:28-35) with pixel jitter (main.cc:116-117)
      const uint32_t row = px / P.width, ii = px - row * P.width;
      const uint32_t j = P.height - 1 - row;
      st.y += 1;
      float u = ((float)ii + xi.x) * P.cam.inv_wm1;
      float v = ((float)j + xi.y) * P.cam.inv_hm1;
      float rad = P.cam.lens_radius * sqrtf(xi.z), sn, cs;
      sincospif(2.0f * xi.w, &sn, &cs);
      float rdx = rad * cs, rdy = rad * sn;
      float offx = P.cam.u[0] * rdx + P.cam.v[0] * rdy;
      float offy = P.cam.u[1] * rdx + P.cam.v[1] * rdy;
      float offz = P.cam.u[2] * rdx + P.cam.v[2] * rdy;
      RO.x = P.cam.origin[0] + offx;
      RO.y = P.cam.origin[1] + offy;
      RO.z = P.cam.origin[2] + offz;
      float ddx = P.cam.llc_minus_origin[0] + u * P.cam.horizontal[0] + v * P.cam.vertical[0] - offx;
      float ddy = P.cam.llc_minus_origin[1] + u * P.cam.horizontal[1] + v * P.cam.vertical[1] - offy;
      float ddz = P.cam.llc_minus_origin[2] + u * P.cam.horizontal[2] + v * P.cam.vertical[2] - offz;
      float len = sqrtf(ddx * ddx + ddy * ddy + ddz * ddz), inv = 1.0f / len;
      RD.x = ddx * inv; RD.y = ddy * inv; RD.z = ddz * inv;
      RD.w = __int_as_float(-1);  // no sphere to skip
      RO.w = len;
      T.x = T.y = T.z = 1.0f;
      depth = (int)P.max_depth;
    } else {
      // hit record, src/cpu/sphere.h:43-48 + hittable.h:16-19: point in f64, then narrowed
      const float dx = RD.x, dy = RD.y, dz = RD.z;
      const double4 S = exact[id];
      const float2 ax = aux[id];  // {1/r (signed), kind}
      float onx = (float)(hpx - S.x) * ax.x, ony = (float)(hpy - S.y) * ax.x,
            onz = (float)(hpz - S.z) * ax.x;
      const bool front = (dx * onx + dy * ony + dz * onz) < 0.0f;
      const float nx = front ? onx : -onx, ny = front ? ony : -ony, nz = front ? onz : -onz;
      RO.x = (float)hpx; RO.y = (float)hpy; RO.z = (float)hpz;
      const float4 M = mat[id];
      const uint32_t kind = (uint32_t)__float_as_int(ax.y);
      const float dn = dx * nx + dy * ny + dz * nz;
      float sx, sy, sz;  // scattered direction (un-normalised, as in the reference)
      bool alive = true;
      if (kind != 2u) {
        // lambertian::scatter (material.h:19-32) and metal::scatter (material.h:40-46) share the
        // unit vector; the metal fuzz point is drawn INSIDE the unit ball
        float ux, uy, uz;
        unit_vector_from(xi.x, xi.y, ux, uy, uz);
        if (kind == 0u) {
          sx = nx + ux; sy = ny + uy; sz = nz + uz;
          if (fabsf(sx) < 1e-8f && fabsf(sy) < 1e-8f && fabsf(sz) < 1e-8f) { sx = nx; sy = ny; sz = nz; }
        } else {
          sx = dx - 2.0f * dn * nx; sy = dy - 2.0f * dn * ny; sz = dz - 2.0f * dn * nz;
          if (M.w > 0.0f) {
            // src/cpu/material.h:43 draws the fuzz point inside the unit ball, the v4 model
            // (src/gpu/material.h:52) on the unit sphere
            float k = P.model_v4 ? M.w : M.w * cbrtf(xi.z);
            sx += k * ux; sy += k * uy; sz += k * uz;
          }
          alive = (sx * nx + sy * ny + sz * nz) > 0.0f;
        }
        T.x *= M.x; T.y *= M.y; T.z *= M.z;
      } else {
        // dielectric::scatter + reflectance + refract, src/cpu/material.h:57-87, vec3.h:126-131
        float ratio = front ? (1.0f / M.w) : M.w;
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
      depth -= 1;
      if (!alive) depth = 0;  // absorbed (material.h:45); depth 0 = bounce limit (main.cc:16)
      if (depth > 0) {
        float len = sqrtf(sx * sx + sy * sy + sz * sz), inv = 1.0f / len;
        RD.x = sx * inv; RD.y = sy * inv; RD.z = sz * inv;
        RO.w = len;
        // A ray that starts on a sphere and points away from its centre side never meets that
        // sphere again: (p - c).d > 0 (clearly, not within rounding of the tangent plane).  Its
        // line still passes the filter, so the resolve pass is told to skip that candidate; this
        // also removes the f32-origin self-hits the t_min = 0.001 rule (main.cc:19) cannot catch
        // at grazing angles on the r = 1000 ground sphere.
        const float away = (RD.x * onx + RD.y * ony + RD.z * onz) * ((float)S.w > 0.f ? 1.f : -1.f);
        RD.w = __int_as_float(away > 1e-3f && !F32HIT ? id : -1);
      } else {
        RO.w = -1.0f;  // no ray this wave: the path regenerates in the next one
      }
    }
  } else {
    RO.w = -1.0f;
  }

  P.ray_o[i] = RO;
  if (depth > 0) P.ray_d[i] = RD;
  T.w = __int_as_float(depth);
  P.thr[i] = T;
  P.acc[i] = A;
  P.it[i] = st;
  return depth >= 0 ? 1u : 0u;  // still part of the frame (with or without a ray this wave)
}

// per-hit scene rows -> shared memory: 1-D bulk (TMA) copies completing on an mbarrier
__device__ __forceinline__ void stage_hit_rows(const WaveParams &P, unsigned char *smem_raw,
                                               uint64_t *bar, const double4 *&exact,
                                               const float4 *&mat, const float2 *&aux) {
  const uint32_t npad = P.scene.n_pad;
  double4 *s_exact = reinterpret_cast<double4 *>(smem_raw);
  float4 *s_mat = reinterpret_cast<float4 *>(smem_raw + (size_t)npad * 32);
  float2 *s_aux = reinterpret_cast<float2 *>(smem_raw + (size_t)npad * 48);
  if (threadIdx.x == 0) mbar_init(bar, 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar, npad * 56u);
    const uint32_t kChunk = 32768u;
    const unsigned char *src[3] = {reinterpret_cast<const unsigned char *>(P.scene.exact),
                                   reinterpret_cast<const unsigned char *>(P.scene.mat),
                                   reinterpret_cast<const unsigned char *>(P.scene.aux)};
    unsigned char *dst[3] = {reinterpret_cast<unsigned char *>(s_exact),
                             reinterpret_cast<unsigned char *>(s_mat),
                             reinterpret_cast<unsigned char *>(s_aux)};
    const uint32_t len[3] = {npad * 32u, npad * 16u, npad * 8u};
    for (int a = 0; a < 3; ++a)
      for (uint32_t off = 0; off < len[a]; off += kChunk)
        bulk_g2s(dst[a] + off, src[a] + off, min(kChunk, len[a] - off), bar);
  }
  mbar_wait(bar, 0);
  exact = s_exact; mat = s_mat; aux = s_aux;
}

template <bool SMEM>
__global__ void __launch_bounds__(256, 4)
shade_kernel(const __grid_constant__ WaveParams P) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t s_bar;
  const double4 *exact = P.scene.exact;
  const float4 *mat = P.scene.mat;
  const float2 *aux = P.scene.aux;
  if (SMEM) stage_hit_rows(P, smem_raw, &s_bar, exact, mat, aux);
  const uint32_t npix = P.width * P.height;
  const uint64_t total_items = (uint64_t)npix * P.n_orders;
  uint32_t live_lanes = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < P.n_slots; i += gridDim.x * blockDim.x)
    live_lanes += shade_slot<false>(P, i, exact, mat, aux, npix, total_items);
  // live rays of the next wave (the host stops the frame when this stays 0)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) live_lanes += __shfl_xor_sync(0xffffffffu, live_lanes, o);
  if ((threadIdx.x & 31) == 0 && live_lanes) atomicAdd(P.live, live_lanes);
}

// ================================================================ BVH traversal (optional mode)
// Closest hit of R rays through the host-built hierarchy (rt_bvh.cu) with the REFERENCE's result:
// the linear scan of hittable_list::hit accepts sphere s when its root N_s (near root if >= t_min,
// else far root) is <= the best so far, so the winner is the smallest root and, among equal roots,
// the LARGEST index (sphere.h:38-40 + hittable_list.h:34-39).  That rule does not depend on the
// visiting order, so leaves may be visited in any order.  Boxes are tested in f32 with outward
// slack; a node is skipped only if it lies entirely beyond the best root or before t_min.
// Writes {id, root numerator} to candidate planes 0..2 of the slot.
template <int R>
__device__ __forceinline__ void bvh_trace(const WaveParams &P, const double4 *exact, const uint32_t base,
                                          const uint32_t stride) {
  const BvhNode *nodes = P.scene.bvh_nodes;
  const uint32_t *prim = P.scene.bvh_prim;
  const float4 *fast = P.scene.fast;
  const bool strict = P.model_v4 != 0u;
#pragma unroll 1
  for (int r = 0; r < R; ++r) {
    const uint32_t slot = base + r * stride;
    const float4 o = P.ray_o[slot];
    if (!(o.w > 0.0f)) continue;  // no ray this wave
    const float4 d = P.ray_d[slot];
    WRay ray;
    ray.ox = o.x; ray.oy = o.y; ray.oz = o.z; ray.dx = d.x; ray.dy = d.y; ray.dz = d.z;
    ray.a = ray.dx * ray.dx + ray.dy * ray.dy + ray.dz * ray.dz;
    ray.tminA = 0.001 * (double)o.w * ray.a;
    double bestN = __longlong_as_double(0x7ff0000000000000ll);
    int id = -1;
    const uint32_t skip = (uint32_t)__float_as_int(d.w);
    const float tmin_f = 0.001f * o.w * 0.9999f;
    float tbest = __int_as_float(0x7f800000);  // upper bound of the best root (unit-direction t)
    const float ix = 1.0f / (fabsf(d.x) > 1e-30f ? d.x : copysignf(1e-30f, d.x));
    const float iy = 1.0f / (fabsf(d.y) > 1e-30f ? d.y : copysignf(1e-30f, d.y));
    const float iz = 1.0f / (fabsf(d.z) > 1e-30f ? d.z : copysignf(1e-30f, d.z));
    const float nod = -(o.x * d.x + o.y * d.y + o.z * d.z);
    const float ox2 = 2.0f * o.x, oy2 = 2.0f * o.y, oz2 = 2.0f * o.z;
    const float oot = (o.x * o.x + o.y * o.y + o.z * o.z) * (1.0f - RTB_FILTER_SLACK);
    // entry distance of the ray into a node's box, or +inf when the box can be skipped
    auto entry = [&](const float4 a, const float4 b) {  // a = {lo.xyz, hi.x}, b = {hi.yz, first, count}
      float t0 = (a.x - o.x) * ix, t1 = (a.w - o.x) * ix;
      float tn = fminf(t0, t1), tf = fmaxf(t0, t1);
      t0 = (a.y - o.y) * iy; t1 = (b.x - o.y) * iy;
      tn = fmaxf(tn, fminf(t0, t1)); tf = fminf(tf, fmaxf(t0, t1));
      t0 = (a.z - o.z) * iz; t1 = (b.y - o.z) * iz;
      tn = fmaxf(tn, fminf(t0, t1)); tf = fminf(tf, fmaxf(t0, t1));
      tf = tf * 1.00001f + 1e-6f;  // robust: never lose a box to rounding
      const bool hit = tn <= tf && tf >= tmin_f && tn <= tbest;
      return hit ? tn : __int_as_float(0x7f800000);
    };
    uint32_t stack[kBvhMaxDepth];
    int sp = 0;
    uint32_t node = 0;
    {
      const float4 *q = reinterpret_cast<const float4 *>(nodes);
      if (entry(__ldg(q), __ldg(q + 1)) == __int_as_float(0x7f800000)) node = 0xffffffffu;
    }
    while (node != 0xffffffffu) {
      const float4 *q = reinterpret_cast<const float4 *>(nodes + node);
      const float4 nb = __ldg(q + 1);
      const uint32_t first = (uint32_t)__float_as_int(nb.z), count = (uint32_t)__float_as_int(nb.w);
      if (count) {
        for (uint32_t k = 0; k < count; ++k) {
          const uint32_t s = __ldg(prim + first + k);
          if (s == skip) continue;  // the sphere the ray just left (see shade_slot)
          const float4 S = __ldg(fast + s);
          const float b = fmaf(S.x, d.x, fmaf(S.y, d.y, fmaf(S.z, d.z, nod)));
          const float e = fmaf(S.x, ox2, fmaf(S.y, oy2, fmaf(S.z, oz2, fmaf(b, b, S.w))));
          if (e >= oot) {
            const double4 E = exact[s];
            double ocx = ray.ox - E.x, ocy = ray.oy - E.y, ocz = ray.oz - E.z;
            double half_b = ocx * ray.dx + ocy * ray.dy + ocz * ray.dz;
            double c = ocx * ocx + ocy * ocy + ocz * ocz - E.w * E.w;
            double disc = half_b * half_b - ray.a * c;
            if (disc >= 0.0) {
              double sq = wave_dsqrt(disc);
              double N = -half_b - sq;
              bool take;
              if (strict) {  // v4: t_min < root, equal roots keep the SMALLER index
                if (N <= ray.tminA) N = -half_b + sq;
                take = N > ray.tminA && (N < bestN || (N == bestN && (int)s < id));
              } else {
                if (N < ray.tminA) N = -half_b + sq;
                take = N >= ray.tminA && (N < bestN || (N == bestN && (int)s > id));
              }
              if (take) {
                bestN = N; id = (int)s;
                tbest = (float)(N * 1.00001) + 1e-6f;
              }
            }
          }
        }
        node = sp ? stack[--sp] : 0xffffffffu;
      } else {
        // inner node: test both children, descend into the nearer one first
        const float4 *c0 = reinterpret_cast<const float4 *>(nodes + first);
        const float t0 = entry(__ldg(c0), __ldg(c0 + 1));
        const float t1 = entry(__ldg(c0 + 2), __ldg(c0 + 3));
        const float INF = __int_as_float(0x7f800000);
        if (t0 == INF && t1 == INF) {
          node = sp ? stack[--sp] : 0xffffffffu;
        } else if (t1 == INF) {
          node = first;
        } else if (t0 == INF) {
          node = first + 1;
        } else {
          const bool near0 = t0 <= t1;
          stack[sp++] = near0 ? first + 1 : first;
          node = near0 ? first : first + 1;
        }
      }
    }
    P.cand_idx[slot] = (uint32_t)id;
    P.cand_idx[(size_t)1 * P.n_slots + slot] = (uint32_t)__double2loint(bestN);
    P.cand_idx[(size_t)2 * P.n_slots + slot] = (uint32_t)__double2hiint(bestN);
  }
}

// ================================================================ fused persistent form
// Every lane owns R slots for the whole frame and alternates shade (one slot at a time, rolled
// loop, state in the pool) and filter (R rays at once, registers hold nothing but rays).  No
// kernel boundary, no inter-lane dependency; warps of an SM drift apart, so the latency-bound
// shade phases of some warps overlap the FMA-pipe-bound filter phases of the others.
// CTAs per SM: the brute-force form needs its 78 registers (3 x 256 threads); the BVH form fits 64
// and is bound by load latency, so it takes a fourth CTA.
constexpr int kBvhMinB = RTW_MINB + 1;
template <int R, int G, bool CROWS, bool SMEM, bool BVH, bool F32HIT = false>
__global__ void __launch_bounds__(kWBlock, BVH ? kBvhMinB : RTW_MINB)
fused_kernel(const __grid_constant__ WaveParams P, const __grid_constant__ ConstRows CR) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t s_bar;
  const double4 *exact = P.scene.exact;
  const float4 *mat = P.scene.mat;
  const float2 *aux = P.scene.aux;
  if (SMEM) stage_hit_rows(P, smem_raw, &s_bar, exact, mat, aux);
  const uint32_t npix = P.width * P.height;
  const uint64_t total_items = (uint64_t)npix * P.n_orders;
  const uint32_t stride = P.n_slots / R;
  const uint32_t base = blockIdx.x * kWBlock + threadIdx.x;  // grid * block == stride
#pragma unroll 1
  for (int r = 0; r < R; ++r) {
    const uint32_t i = base + r * stride;
    P.thr[i] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));  // depth 0: needs a path
    P.acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    P.it[i] = make_uint4(0xffffffffu, 0u, 0u, 0u);
  }
  for (;;) {
    uint32_t live = 0;
#pragma unroll 1
    for (int r = 0; r < R; ++r) live += shade_slot<BVH, F32HIT>(P, base + r * stride, exact, mat, aux, npix, total_items);
#ifdef RTW_CTA_SYNC
    // CTA-wide barrier + vote: the warps of a CTA walk the sphere rows in step (one stream of
    // constant-cache lines per CTA instead of one per warp)
    if (!__syncthreads_or(live != 0u)) break;
#else
    if (__all_sync(0xffffffffu, live == 0u)) break;  // warp-uniform exit keeps the loop convergent
#endif
    if (BVH) bvh_trace<R>(P, exact, base, stride);
    else if (P.no_filter) {
      // soundness switch: no candidates are recorded, every slot takes the full-scan branch
#pragma unroll
      for (int r = 0; r < R; ++r) P.cand_cnt[base + r * stride] = kCandPlanes + 1u;
    } else filter_rays<R, G, CROWS>(P, CR, base, stride);
  }
}

__global__ void __launch_bounds__(256) wave_init_kernel(const __grid_constant__ WaveParams P) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < P.n_slots; i += gridDim.x * blockDim.x) {
    P.thr[i] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));  // depth 0: needs a path
    P.acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    P.it[i] = make_uint4(0xffffffffu, 0u, 0u, 0u);
    P.ray_o[i] = make_float4(0.f, 0.f, 0.f, -1.0f);
    P.ray_d[i] = make_float4(0.f, 1.f, 0.f, 0.f);
    P.cand_cnt[i] = 0;
  }
}

// ================================================================ host side
size_t wave_pool_bytes(uint32_t n_slots) {
  // ray_o, ray_d, thr, acc (float4) + it (uint4) + cand_cnt + cand_idx planes
  return (size_t)n_slots * (5 * 16 + 4 + 4 * (size_t)kCandPlanes);
}

uint32_t wave_default_slots(int sm_count) {
  // a whole number of filter-kernel rounds: SMs x resident lanes x rays per lane x rounds
  // (also the capacity the fused kernel needs at R = 4)
  return (uint32_t)sm_count * RTW_MINB * kWBlock * 2u * 2u;
}

void wave_carve(WaveParams &p, void *pool, uint32_t n_slots) {
  unsigned char *b = static_cast<unsigned char *>(pool);
  size_t n = n_slots;
  p.n_slots = n_slots;
  p.ray_o = reinterpret_cast<float4 *>(b); b += n * 16;
  p.ray_d = reinterpret_cast<float4 *>(b); b += n * 16;
  p.thr = reinterpret_cast<float4 *>(b); b += n * 16;
  p.acc = reinterpret_cast<float4 *>(b); b += n * 16;
  p.it = reinterpret_cast<uint4 *>(b); b += n * 16;
  p.cand_cnt = reinterpret_cast<uint32_t *>(b); b += n * 4;
  p.cand_idx = reinterpret_cast<uint32_t *>(b);
}

template <int R, int G>
static cudaError_t launch_filter_t(const WaveParams &p, const ConstRows &cr, bool crows, int grid,
                                   cudaStream_t st) {
  if (crows)
    filter_kernel<R, G, true><<<grid, kWBlock, 0, st>>>(p, cr);
  else
    filter_kernel<R, G, false><<<grid, kWBlock, 0, st>>>(p, cr);
  return cudaGetLastError();
}

cudaError_t run_wavefront(const WaveParams &p, const float4 *host_rows, int sm_count,
                          size_t smem_optin_max, cudaStream_t st, unsigned int *host_live,
                          WaveRunInfo *info) {
  static thread_local ConstRows cr;  // 16 KB kernel parameter; frames may render from several host threads
  const bool crows = host_rows && p.scene.n_pad + 2 * kRowTail <= (uint32_t)kConstRows &&
                     !env_cfg().no_const_rows;
  if (crows) memcpy(cr.row, host_rows, (size_t)(p.scene.n_pad + kRowTail) * sizeof(float4));
  const int rpl = env_cfg().rays_per_lane;
  if (rpl != 2 && rpl != 4) return cudaErrorInvalidValue;
  if (p.n_slots % (kWBlock * 4u)) return cudaErrorInvalidValue;

  // shade kernel: per-hit rows in shared memory when they fit
  const size_t scene_smem = (size_t)p.scene.n_pad * 56;
  const bool in_smem = scene_smem + 1024 <= smem_optin_max / 4;
  cudaError_t e;
  if (in_smem) {
    e = cudaFuncSetAttribute(shade_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)scene_smem);
    if (e != cudaSuccess) return e;
  }
  int shade_per_sm = 0;
  e = in_smem ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&shade_per_sm, shade_kernel<true>, 256, scene_smem)
              : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&shade_per_sm, shade_kernel<false>, 256, 0);
  if (e != cudaSuccess) return e;
  if (shade_per_sm < 1) return cudaErrorLaunchOutOfResources;
  const int shade_grid = sm_count * shade_per_sm;
  const int filter_grid = sm_count * RTW_MINB;

  auto shade = [&]() -> cudaError_t {
    if (in_smem) shade_kernel<true><<<shade_grid, 256, scene_smem, st>>>(p);
    else shade_kernel<false><<<shade_grid, 256, 0, st>>>(p);
    return cudaGetLastError();
  };
  auto filter = [&]() -> cudaError_t {
    return rpl == 4 ? launch_filter_t<4, RTW_G>(p, cr, crows, filter_grid, st)
                    : launch_filter_t<2, RTW_G>(p, cr, crows, filter_grid, st);
  };

  uint32_t launches = 0, waves = 0;
  wave_init_kernel<<<sm_count * 4, 256, 0, st>>>(p);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  ++launches;

  // every sample needs at least one wave of its slot: no need to poll before that many waves
  uint64_t total_samples = 0;
  for (uint32_t k = p.slice_begin; k < p.slice_end; ++k) {
    uint32_t a, b;
    slice_range(k, p.spp, a, b);
    total_samples += b - a;
  }
  total_samples *= (uint64_t)p.width * p.height;
  const uint64_t min_waves = total_samples / p.n_slots;
  const uint32_t poll_every = 4;

  for (;;) {
    const bool poll = waves >= min_waves && (waves - min_waves) % poll_every == 0;
    if (poll && (e = cudaMemsetAsync(p.live, 0, sizeof(unsigned int), st)) != cudaSuccess) return e;
    if ((e = shade()) != cudaSuccess) return e;
    ++launches;
    if (poll) {
      if ((e = cudaMemcpyAsync(host_live, p.live, sizeof(unsigned int), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
      if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
      if (*host_live == 0) break;
    }
    if ((e = filter()) != cudaSuccess) return e;
    ++launches;
    ++waves;
  }
  if (info) {
    info->launches = launches; info->waves = waves; info->n_slots = p.n_slots;
    info->rays_per_lane = rpl; info->filter_grid = filter_grid; info->shade_grid = shade_grid;
    info->rows_in_constant_bank = crows; info->scene_in_smem = in_smem;
  }
  return cudaSuccess;
}

template <int R, int G>
static cudaError_t launch_fused_t(WaveParams &p, void *pool, uint32_t pool_slots, const ConstRows &cr,
                                  bool crows, bool in_smem, bool bvh, bool f32hit, size_t smem,
                                  int sm_count, int *grid_out, cudaStream_t st) {
#define RTW_FUSED(CR_, SM_, BV_, ...)                                                              \
  do {                                                                                             \
    auto k = fused_kernel<R, G, CR_, SM_, BV_, ##__VA_ARGS__>;                                     \
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e != cudaSuccess) return e;                                                                \
    int per_sm = 0;                                                                                \
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, kWBlock, smem);                  \
    if (e != cudaSuccess) return e;                                                                \
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;                                          \
    /* persistent grid: every CTA must be resident (one full wave, a multiple of the SM count) */  \
    int want = BV_ ? kBvhMinB : RTW_MINB;                                                          \
    while (want > 1 && (size_t)sm_count * want * kWBlock * R > pool_slots) --want;                 \
    const int grid = sm_count * (per_sm < want ? per_sm : want);                                   \
    const uint32_t n_slots = (uint32_t)grid * kWBlock * (uint32_t)R;                               \
    if (n_slots > pool_slots) return cudaErrorInvalidValue;                                        \
    wave_carve(p, pool, n_slots);                                                                  \
    *grid_out = grid;                                                                              \
    k<<<grid, kWBlock, smem, st>>>(p, cr);                                                         \
  } while (0)
  if (f32hit && crows && in_smem) RTW_FUSED(true, true, false, true);
  else if (f32hit) RTW_FUSED(false, false, false, true);
  else if (bvh && in_smem) RTW_FUSED(false, true, true);
  else if (bvh) RTW_FUSED(false, false, true);
  else if (crows && in_smem) RTW_FUSED(true, true, false);
  else if (crows) RTW_FUSED(true, false, false);
  else if (in_smem) RTW_FUSED(false, true, false);
  else RTW_FUSED(false, false, false);
#undef RTW_FUSED
  return cudaGetLastError();
}

cudaError_t run_fused(WaveParams p, void *pool, uint32_t pool_slots, const float4 *host_rows,
                      bool use_bvh, bool f32_hit, int sm_count, size_t smem_optin_max, cudaStream_t st,
                      WaveRunInfo *info) {
  if (use_bvh && !p.scene.bvh_nodes) return cudaErrorInvalidValue;
  if (use_bvh && f32_hit) return cudaErrorInvalidValue;
  static thread_local ConstRows cr;  // 16 KB kernel parameter; frames may render from several host threads
  const bool crows = host_rows && p.scene.n_pad + 2 * kRowTail <= (uint32_t)kConstRows &&
                     !env_cfg().no_const_rows;
  if (crows) memcpy(cr.row, host_rows, (size_t)(p.scene.n_pad + kRowTail) * sizeof(float4));
  const int rpl = env_cfg().rays_per_lane;
  if (rpl != 2 && rpl != 4) return cudaErrorInvalidValue;
  const size_t scene_smem = (size_t)p.scene.n_pad * 56;
  const bool in_smem = (scene_smem + 1024) * RTW_MINB <= smem_optin_max;
  const size_t smem = in_smem ? scene_smem : 0;
  int grid = 0;
  cudaError_t e = rpl == 4 ? launch_fused_t<4, RTW_G>(p, pool, pool_slots, cr, crows, in_smem, use_bvh, f32_hit, smem, sm_count, &grid, st)
                           : launch_fused_t<2, RTW_G>(p, pool, pool_slots, cr, crows, in_smem, use_bvh, f32_hit, smem, sm_count, &grid, st);
  const uint32_t n_slots = p.n_slots;
  if (info) {
    info->launches = 1; info->waves = 0; info->n_slots = n_slots; info->rays_per_lane = rpl;
    info->filter_grid = grid; info->shade_grid = 0; info->rows_in_constant_bank = crows;
    info->scene_in_smem = in_smem;
  }
  return e;
}

}  // namespace rtb
