// csrc/rt_path.cu — the persistent path-tracing megakernel (sm_100a).
//
// Replaces the pixel/sample loops of src/cpu/main.cc:111-123 and everything they call
// (camera::get_ray, ray_color, hittable_list::hit, sphere::hit, material::scatter).
//
// Shape of the kernel (full text in DESIGN.md):
//  * Work item = (pixel, canonical sample slice).  A lane owns R items at a time ("slots"),
//    sums each item's samples sequentially in f32 and writes accum[slice][pixel] once, so the
//    image does not depend on scheduling or on the GPU count.  A slot whose path ends starts
//    its next sample at once (path regeneration), so lanes never idle inside a frame.
//  * Closest hit = brute force over all spheres in reference order, in two stages:
//      filter   f32, conservative LINE-sphere test, 7 FMA per (ray, sphere).  Slots are packed
//               in pairs and every FMA is a Blackwell packed FFMA2 (fma.rn.f32x2): one issue
//               slot does the work for two rays, the sphere row {cx,cy,cz,w} is one broadcast
//               LDS.128 shared by all R slots of the lane.
//      resolve  the handful of survivors per ray (kept in a per-lane shared-memory list, in
//               ascending sphere order) go through the reference's sphere::hit arithmetic in
//               f64, so closest-hit decisions do not suffer f32 cancellation (r = 1000 ground).
//  * Scene rows are staged into shared memory once per CTA with 1-D bulk (TMA) copies that
//    complete on an mbarrier.
#include "rt_device.cuh"
#include "rt_kernels.h"

#include <cstdlib>
#include <cstring>

namespace rtb {

#ifndef RTB_BRANCH_VOTE
#define RTB_BRANCH_VOTE 0  // 1: warp-uniform vote + branch; 0: plain divergent branch
#endif

#ifndef RTB_BLOCK
#define RTB_BLOCK 256
#endif
#ifndef RTB_MINB2
#define RTB_MINB2 2  // resident CTAs per SM the R = 2 kernel is compiled for
#endif
constexpr int kBlock = RTB_BLOCK;
constexpr int kCandCap = 8;  // candidate list capacity per slot before an in-loop flush
#ifndef RTB_GROUP
#define RTB_GROUP 4
#endif
constexpr int kGroup = RTB_GROUP;  // sphere rows per filter group; kRowPad is a multiple of 2*kGroup
static_assert(kRowPad % (2 * kGroup) == 0 && kRowTail >= kGroup, "row padding");

// ---------------------------------------------------------------- packed f32x2 helpers
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a),
                     rb = *reinterpret_cast<unsigned long long *>(&b),
                     rc = *reinterpret_cast<unsigned long long *>(&c), rd;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  return *reinterpret_cast<float2 *>(&rd);
}
__device__ __forceinline__ float2 bc(float v) { return make_float2(v, v); }  // -> .F32 operand

// element r of a pair-packed array (r is a compile-time constant after unrolling)
#define PK(arr, r) (((r)&1) ? (arr)[(r) >> 1].y : (arr)[(r) >> 1].x)

// ---------------------------------------------------------------- f64 candidate resolution
// sphere::hit (src/cpu/sphere.h:24-51) on numerators: root = N / a with a = |d|^2 common to
// all spheres of a ray, so roots are compared as N against t_min*a and bestN (no division).
// Inclusive upper bound: a later sphere with an equal root replaces the earlier one
// (hittable_list.h:34-39 + sphere.h:38-40).
// sqrt of a non-negative double: f32 reciprocal-square-root seed + two Heron steps in f64
// (relative error ~2^-45 after the first, below 1 ulp after the second); no IEEE fix-up code.
__device__ __forceinline__ double fast_dsqrt(double x) {
  float xf = fmaxf((float)x, 1e-30f);
  double y = (double)rsqrtf(xf);   // ~ 1/sqrt(x), rel. error 2^-22
  double h = 0.5 * y;
  double s = x * y;
  s = fma(fma(-s, s, x), h, s);
  s = fma(fma(-s, s, x), h, s);
  return s;
}

struct RayD {
  double ox, oy, oz, dx, dy, dz, a, tminA;
};

__device__ __forceinline__ bool exact_hit_num(const double4 S, const RayD &r, double &bestN) {
  double ocx = r.ox - S.x, ocy = r.oy - S.y, ocz = r.oz - S.z;
  double half_b = ocx * r.dx + ocy * r.dy + ocz * r.dz;
  double c = ocx * ocx + ocy * ocy + ocz * ocz - S.w * S.w;
  double disc = half_b * half_b - r.a * c;
  if (disc < 0.0) return false;
  double sq = fast_dsqrt(disc);
  double N = -half_b - sq;
  if (N < r.tminA || bestN < N) {
    N = -half_b + sq;
    if (N < r.tminA || bestN < N) return false;
  }
  bestN = N;
  return true;
}

__device__ __forceinline__ RayD make_rayd(float ox, float oy, float oz, float dx, float dy,
                                          float dz, float dlen) {
  RayD r;
  r.ox = ox; r.oy = oy; r.oz = oz; r.dx = dx; r.dy = dy; r.dz = dz;
  r.a = r.dx * r.dx + r.dy * r.dy + r.dz * r.dz;
  r.tminA = 0.001 * (double)dlen * r.a;  // main.cc:19 t_min, in unit-direction units, times a
  return r;
}

__device__ __forceinline__ void resolve_list(const uint32_t *list, uint32_t cnt,
                                             const double4 *exact, const RayD &ray,
                                             double &bestN, int &id) {
  for (uint32_t k = 0; k < cnt; ++k) {
    uint32_t s = list[k * kBlock];
    if (exact_hit_num(exact[s], ray, bestN)) id = (int)s;
  }
}

// Rare path: a slot's candidate list filled up inside the sphere loop.  The running hit of
// the slot lives in shared memory (hit_n / hit_id) so the hot loop carries no f64 state.
__device__ __noinline__ void flush_overflow(const uint32_t *list, const double4 *exact, float ox,
                                            float oy, float oz, float dx, float dy, float dz,
                                            float dlen, double *hit_n, int *hit_id, bool first) {
  RayD ray = make_rayd(ox, oy, oz, dx, dy, dz, dlen);
  double bestN = first ? __longlong_as_double(0x7ff0000000000000ll) : *hit_n;
  int id = first ? -1 : *hit_id;
  resolve_list(list, kCandCap, exact, ray, bestN, id);
  *hit_n = bestN;
  *hit_id = id;
}

// ---------------------------------------------------------------- the kernel
// R = slots (rays in flight) per lane, even; SMEM = scene rows live in shared memory.
// CROWS = filter rows are read from the kernel-parameter (constant) bank through uniform
// registers (LDCU): every FFMA2 of the filter then has exactly one sphere operand, a uniform
// register, and reads at most two register pairs -> the register file is no longer the limit
// (tools/microbench.cu, "FFMA2 uniform-register multiplier").
template <int R, bool SMEM, bool CROWS>
__global__ void __launch_bounds__(kBlock, (R <= 2 ? RTB_MINB2 : 1))
path_kernel(const __grid_constant__ PathParams P, const __grid_constant__ ConstRows CR) {
  static_assert(R % 2 == 0, "slots are packed in pairs");
  constexpr int NP = R / 2;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const uint32_t n = P.scene.n, npad = P.scene.n_pad;
  const uint32_t tid = threadIdx.x;

  // ---- shared memory carve-up (scene part only when SMEM)
  const size_t scene_bytes = SMEM ? (size_t)npad * 72 + kRowTail * 16 : 0;
  // [exact npad*32][mat npad*16][aux npad*8][fast (npad + kRowTail)*16]
  double4 *s_exact = reinterpret_cast<double4 *>(smem_raw);
  float4 *s_mat = reinterpret_cast<float4 *>(smem_raw + (size_t)npad * 32);
  float2 *s_aux = reinterpret_cast<float2 *>(smem_raw + (size_t)npad * 48);
  float4 *s_fast = reinterpret_cast<float4 *>(smem_raw + (size_t)npad * 56);
  uint32_t *s_cand = reinterpret_cast<uint32_t *>(smem_raw + scene_bytes);
  double *s_hit_n = reinterpret_cast<double *>(smem_raw + scene_bytes +
                                               (size_t)R * kCandCap * kBlock * 4);
  int *s_hit_id = reinterpret_cast<int *>(smem_raw + scene_bytes +
                                          (size_t)R * kCandCap * kBlock * 4 + (size_t)R * kBlock * 8);
  __shared__ __align__(8) uint64_t s_bar;

  const float4 *fast = SMEM ? s_fast : P.scene.fast;
  const double4 *exact = SMEM ? s_exact : P.scene.exact;
  const float4 *mat = SMEM ? s_mat : P.scene.mat;
  const float2 *aux = SMEM ? s_aux : P.scene.aux;

  if (SMEM) {
    // Scene rows -> shared memory: one elected thread issues 1-D bulk (TMA) copies that
    // complete on an mbarrier; everyone waits on the barrier phase.
    if (tid == 0) mbar_init(&s_bar, 1);
    __syncthreads();
    if (tid == 0) {
      mbar_expect_tx(&s_bar, npad * 72u + kRowTail * 16u);
      const uint32_t kChunk = 32768u;
      const unsigned char *src[4] = {reinterpret_cast<const unsigned char *>(P.scene.fast),
                                     reinterpret_cast<const unsigned char *>(P.scene.exact),
                                     reinterpret_cast<const unsigned char *>(P.scene.mat),
                                     reinterpret_cast<const unsigned char *>(P.scene.aux)};
      unsigned char *dst[4] = {reinterpret_cast<unsigned char *>(s_fast),
                               reinterpret_cast<unsigned char *>(s_exact),
                               reinterpret_cast<unsigned char *>(s_mat),
                               reinterpret_cast<unsigned char *>(s_aux)};
      const uint32_t len[4] = {(npad + kRowTail) * 16u, npad * 32u, npad * 16u, npad * 8u};
      for (int a = 0; a < 4; ++a)
        for (uint32_t off = 0; off < len[a]; off += kChunk)
          bulk_g2s(dst[a] + off, src[a] + off, min(kChunk, len[a] - off), &s_bar);
    }
    mbar_wait(&s_bar, 0);
  }

  const uint32_t npix = P.width * P.height;
  const uint64_t total_items = (uint64_t)npix * (P.slice_end - P.slice_begin);
  uint32_t *my_cand = s_cand + tid;  // element (r, k) at my_cand[(r*kCandCap + k) * kBlock]

  // ---- lane state
  float2 OX[NP], OY[NP], OZ[NP], DX[NP], DY[NP], DZ[NP];  // ray origin / unit direction
  float dlen[R];                  // |direction| of the reference's un-normalised ray
  float tr[R], tg[R], tb[R];      // path throughput (product of attenuations)
  float ar[R], ag[R], ab[R];      // item accumulator (sum over the slice's samples)
  float nrays[R];                 // ray segments traced for this item
  uint32_t item[R];               // slice << kItemShift | pixel; 0xffffffff = no item yet
  uint32_t samp[R], samp_end[R];  // next sample / end of the slice (frame-relative)
  int depth[R];                   // > 0 bounces left (main.cc:16); 0 = needs a path; -1 = dead
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    OX[p] = OY[p] = OZ[p] = DX[p] = DZ[p] = make_float2(0.f, 0.f);
    DY[p] = make_float2(1.f, 1.f);
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    dlen[r] = 1.f; tr[r] = tg[r] = tb[r] = 0.f; ar[r] = ag[r] = ab[r] = 0.f; nrays[r] = 0.f;
    item[r] = 0xffffffffu; samp[r] = samp_end[r] = 0; depth[r] = 0;
  }

  for (;;) {
    // ------------------------------------------------------------ (1) regenerate paths
    bool all_dead = true;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (depth[r] == 0) {
        while (samp[r] == samp_end[r]) {  // item exhausted: flush it and fetch the next
          if (item[r] != 0xffffffffu) {
            P.accum[(size_t)(item[r] >> kItemShift) * npix + (item[r] & kItemPixelMask)] =
                make_float4(ar[r], ag[r], ab[r], nrays[r]);
            item[r] = 0xffffffffu;
          }
          unsigned long long it = atomicAdd(P.counter, 1ull);
          if (it >= total_items) { depth[r] = -1; break; }
          uint32_t sl = (uint32_t)(it / npix);
          uint32_t px = (uint32_t)(it - (uint64_t)sl * npix);
          sl += P.slice_begin;
          item[r] = sl << kItemShift | px;
          // canonical slice k covers samples [floor(k*spp/S), floor((k+1)*spp/S)), S = RT_SLICES
          samp[r] = (uint32_t)(((uint64_t)sl * P.spp) / RT_SLICES_DEV);
          samp_end[r] = (uint32_t)(((uint64_t)(sl + 1) * P.spp) / RT_SLICES_DEV);
          ar[r] = ag[r] = ab[r] = 0.f; nrays[r] = 0.f;
        }
        if (depth[r] == 0) {
          // camera::get_ray (src/cpu/camera.h:28-35) with pixel jitter (main.cc:116-117)
          const uint32_t px = item[r] & kItemPixelMask;
          const uint32_t row = px / P.width, i = px - row * P.width;
          const uint32_t j = P.height - 1 - row;
          const float4 xi = philox_uniforms(px, P.sample_offset + samp[r], 0u, P.seed);
          samp[r] += 1;
          float u = ((float)i + xi.x) * P.cam.inv_wm1;
          float v = ((float)j + xi.y) * P.cam.inv_hm1;
          float rad = P.cam.lens_radius * sqrtf(xi.z), sn, cs;
          sincospif(2.0f * xi.w, &sn, &cs);
          float rdx = rad * cs, rdy = rad * sn;
          float offx = P.cam.u[0] * rdx + P.cam.v[0] * rdy;
          float offy = P.cam.u[1] * rdx + P.cam.v[1] * rdy;
          float offz = P.cam.u[2] * rdx + P.cam.v[2] * rdy;
          PK(OX, r) = P.cam.origin[0] + offx;
          PK(OY, r) = P.cam.origin[1] + offy;
          PK(OZ, r) = P.cam.origin[2] + offz;
          float ddx = P.cam.llc_minus_origin[0] + u * P.cam.horizontal[0] +
                      v * P.cam.vertical[0] - offx;
          float ddy = P.cam.llc_minus_origin[1] + u * P.cam.horizontal[1] +
                      v * P.cam.vertical[1] - offy;
          float ddz = P.cam.llc_minus_origin[2] + u * P.cam.horizontal[2] +
                      v * P.cam.vertical[2] - offz;
          float len = sqrtf(ddx * ddx + ddy * ddy + ddz * ddz), inv = 1.0f / len;
          PK(DX, r) = ddx * inv; PK(DY, r) = ddy * inv; PK(DZ, r) = ddz * inv;
          dlen[r] = len;
          tr[r] = tg[r] = tb[r] = 1.0f;
          depth[r] = (int)P.max_depth;
        }
      }
      all_dead = all_dead && depth[r] < 0;
    }
    if (__all_sync(0xffffffffu, all_dead)) break;  // warp-uniform exit keeps the loop convergent

    // ------------------------------------------------------------ (2) filter: all spheres
    // hittable_list::hit (src/cpu/hittable_list.h:28-43) visits every sphere in order.  The
    // line o + t d (|d| = 1) meets sphere (c, r) iff  (c.d - o.d)^2 + 2 c.o + (r^2 - |c|^2)
    // >= |o|^2 ; the row holds {c, r^2 - |c|^2 + slack} and the test is evaluated with slack on
    // both sides so that f32 rounding can only ADD candidates (filter error bound).
    float2 NOD[NP], OX2[NP], OY2[NP], OZ2[NP], OOT[NP];
    uint32_t cnt = 0;  // 4 bits per slot: candidates in the slot's list
    uint32_t flushed = 0;  // bit r: slot r already has a running hit in s_hit_*
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      OX2[p] = make_float2(2.0f * OX[p].x, 2.0f * OX[p].y);
      OY2[p] = make_float2(2.0f * OY[p].x, 2.0f * OY[p].y);
      OZ2[p] = make_float2(2.0f * OZ[p].x, 2.0f * OZ[p].y);
      NOD[p] = make_float2(-(OX[p].x * DX[p].x + OY[p].x * DY[p].x + OZ[p].x * DZ[p].x),
                           -(OX[p].y * DX[p].y + OY[p].y * DY[p].y + OZ[p].y * DZ[p].y));
      OOT[p] = make_float2(
          (OX[p].x * OX[p].x + OY[p].x * OY[p].x + OZ[p].x * OZ[p].x) * (1.0f - RTB_FILTER_SLACK),
          (OX[p].y * OX[p].y + OY[p].y * OY[p].y + OZ[p].y * OZ[p].y) * (1.0f - RTB_FILTER_SLACK));
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const bool live = depth[r] > 0;
      if (!live) PK(OOT, r) = __int_as_float(0x7f800000);  // idle slots never pass the filter
      nrays[r] += live ? 1.f : 0.f;
    }

    // Groups of kGroup rows: all FFMA2 chains of a group are independent (ILP hides the FMA
    // latency), the per-slot maximum over the group is compared once, and ONE branch per
    // group leads to the rare path.  The next group's rows are fetched before the branch
    // (software prefetch; the arrays carry one extra padding group), so the LDS latency is
    // off the loop-carried path.
#ifdef RTB_LOOP_REPEAT  // timing experiment only: run the filter pass several times
    for (int rep = 0; rep < RTB_LOOP_REPEAT; ++rep)
#endif
    {
#ifdef RTB_LOOP_REPEAT
      cnt = 0; flushed = 0;
#endif
      const float4 *frow = fast;
      float4 S[kGroup];
#pragma unroll
      for (int k = 0; k < kGroup; ++k) S[k] = CROWS ? CR.row[k] : frow[k];
      uint32_t npad_r = npad;
      asm volatile("" : "+r"(npad_r));  // keep the bound in a register (no per-iteration LDCU)
#pragma unroll 2
      for (uint32_t s0 = 0; s0 < npad_r; s0 += kGroup) {
        float2 e[kGroup][NP];
#pragma unroll
        for (int k = 0; k < kGroup; ++k) {
#pragma unroll
          for (int p = 0; p < NP; ++p) {
            // e = (c.d - o.d)^2 + (2 c.o + w), as one chain in which every FFMA2 multiplies
            // by exactly one sphere scalar
            float2 b = ffma2(bc(S[k].x), DX[p],
                             ffma2(bc(S[k].y), DY[p], ffma2(bc(S[k].z), DZ[p], NOD[p])));
            float2 e1 = ffma2(b, b, bc(S[k].w));
            e[k][p] = ffma2(bc(S[k].x), OX2[p],
                            ffma2(bc(S[k].y), OY2[p], ffma2(bc(S[k].z), OZ2[p], e1)));
          }
        }
        frow += kGroup;
#pragma unroll
        for (int k = 0; k < kGroup; ++k) S[k] = CROWS ? CR.row[s0 + kGroup + k] : frow[k];
        bool any = false;
#pragma unroll
        for (int p = 0; p < NP; ++p) {
          float mx = e[0][p].x, my = e[0][p].y;
#pragma unroll
          for (int k = 1; k < kGroup; ++k) { mx = fmaxf(mx, e[k][p].x); my = fmaxf(my, e[k][p].y); }
          any = any || mx >= OOT[p].x || my >= OOT[p].y;
        }
        if (CROWS ? __any_sync(0xffffffffu, any) : __builtin_expect(any, 0)) {
          // bit (k*R + r): sphere s0+k survives for slot r; ascending bits = ascending sphere
          // order within every slot
          uint32_t m = 0;
#pragma unroll
          for (int k = 0; k < kGroup; ++k)
#pragma unroll
            for (int r = 0; r < R; ++r) m |= (PK(e[k], r) >= PK(OOT, r) ? 1u : 0u) << (k * R + r);
          while (m) {
            const uint32_t bit = (uint32_t)__ffs((int)m) - 1u;
            m &= m - 1u;
            const uint32_t r = bit % R, c = (cnt >> (4u * r)) & 15u;
            my_cand[(r * kCandCap + c) * kBlock] = s0 + bit / R;
            cnt += 1u << (4u * r);
            if (c + 1u == kCandCap) {  // list full: resolve it now (rare)
#pragma unroll
              for (int rr = 0; rr < R; ++rr)
                if (rr == (int)r)
                  flush_overflow(my_cand + rr * kCandCap * kBlock, exact, PK(OX, rr), PK(OY, rr),
                                 PK(OZ, rr), PK(DX, rr), PK(DY, rr), PK(DZ, rr), dlen[rr],
                                 s_hit_n + rr * kBlock + tid, s_hit_id + rr * kBlock + tid,
                                 !(flushed >> rr & 1u));
              flushed |= 1u << r;
              cnt &= ~(15u << (4u * r));
            }
          }
        }
      }
    }

    // ------------------------------------------------------------ (3) resolve + shade
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (depth[r] <= 0) continue;
      const float ox = PK(OX, r), oy = PK(OY, r), oz = PK(OZ, r);
      const float dx = PK(DX, r), dy = PK(DY, r), dz = PK(DZ, r);
      const RayD ray = make_rayd(ox, oy, oz, dx, dy, dz, dlen[r]);
      double bestN = __longlong_as_double(0x7ff0000000000000ll);  // t_max = infinity
      int id = -1;
      if (flushed >> r & 1u) { bestN = s_hit_n[r * kBlock + tid]; id = s_hit_id[r * kBlock + tid]; }
      resolve_list(my_cand + r * kCandCap * kBlock, (cnt >> (4 * r)) & 15u, exact, ray, bestN, id);
      if (id < 0) {
        // sky, src/cpu/main.cc:27-29 (direction already unit length)
        float t = 0.5f * (dy + 1.0f);
        float w = 1.0f - t;
        ar[r] += tr[r] * (w + t * 0.5f);
        ag[r] += tg[r] * (w + t * 0.7f);
        ab[r] += tb[r] * (w + t * 1.0f);
        depth[r] = 0;
        continue;
      }
      // hit record, src/cpu/sphere.h:43-48 + hittable.h:16-19: point in f64, then narrowed
      const double4 S = exact[id];
      // t = bestN / a with a = 1 + eps (|eps| ~ 1e-7): 1/a = 2 - a to within eps^2
      const double t = bestN * (2.0 - ray.a);
      const double px = ray.ox + t * ray.dx, py = ray.oy + t * ray.dy, pz = ray.oz + t * ray.dz;
      const float2 ax = aux[id];  // {1/r (signed), kind}
      float onx = (float)(px - S.x) * ax.x, ony = (float)(py - S.y) * ax.x,
            onz = (float)(pz - S.z) * ax.x;
      const bool front = (dx * onx + dy * ony + dz * onz) < 0.0f;
      const float nx = front ? onx : -onx, ny = front ? ony : -ony, nz = front ? onz : -onz;
      PK(OX, r) = (float)px; PK(OY, r) = (float)py; PK(OZ, r) = (float)pz;

      const uint32_t slot = P.max_depth - (uint32_t)depth[r] + 1u;  // k-th hit of this path
      const float4 xi =
          philox_uniforms(item[r] & kItemPixelMask, P.sample_offset + samp[r] - 1u, slot, P.seed);
      const float4 M = mat[id];
      const uint32_t kind = (uint32_t)__float_as_int(ax.y);
      float sx, sy, sz;  // scattered direction (un-normalised, as in the reference)
      bool alive = true;
      if (kind == 0u) {
        // lambertian::scatter, src/cpu/material.h:19-32
        float ux, uy, uz;
        unit_vector_from(xi.x, xi.y, ux, uy, uz);
        sx = nx + ux; sy = ny + uy; sz = nz + uz;
        if (fabsf(sx) < 1e-8f && fabsf(sy) < 1e-8f && fabsf(sz) < 1e-8f) { sx = nx; sy = ny; sz = nz; }
        tr[r] *= M.x; tg[r] *= M.y; tb[r] *= M.z;
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
        tr[r] *= M.x; tg[r] *= M.y; tb[r] *= M.z;
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
      depth[r] -= 1;
      if (!alive) depth[r] = 0;  // absorbed (material.h:45); depth 0 = bounce limit (main.cc:16)
      if (depth[r] == 0) continue;
      float len = sqrtf(sx * sx + sy * sy + sz * sz), inv = 1.0f / len;
      PK(DX, r) = sx * inv; PK(DY, r) = sy * inv; PK(DZ, r) = sz * inv;
      dlen[r] = len;
    }
  }
}

// ---------------------------------------------------------------- launcher
template <int R, bool SMEM, bool CROWS>
static cudaError_t launch_path_t(const PathParams &p, const ConstRows &cr, int sm_count,
                                 size_t smem, cudaStream_t st, PathLaunchInfo *info) {
  auto kern = path_kernel<R, SMEM, CROWS>;
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
    info->rows_in_constant_bank = CROWS;
  }
  kern<<<grid, kBlock, smem, st>>>(p, cr);
  return cudaGetLastError();
}

cudaError_t launch_path(const PathParams &p, const float4 *host_rows, int rays_per_lane,
                        int sm_count, size_t smem_optin_max, cudaStream_t st,
                        PathLaunchInfo *info) {
  int R = rays_per_lane ? rays_per_lane : 2;
  size_t lists = (size_t)kBlock * R * (kCandCap * 4 + 8 + 4);
  size_t scene = (size_t)p.scene.n_pad * 72 + kRowTail * 16;
  bool in_smem = scene + lists + 1024 <= smem_optin_max;
  size_t smem = lists + (in_smem ? scene : 0);
  static ConstRows cr;  // rows beyond the scene are never read
  bool crows = in_smem && host_rows && p.scene.n_pad + kRowTail <= (uint32_t)kConstRows &&
               !getenv("RT_B200_NO_CONST_ROWS");
  if (crows) memcpy(cr.row, host_rows, (size_t)(p.scene.n_pad + kRowTail) * sizeof(float4));
#define RTB_DISPATCH(RR)                                                                  \
  return crows     ? launch_path_t<RR, true, true>(p, cr, sm_count, smem, st, info)       \
         : in_smem ? launch_path_t<RR, true, false>(p, cr, sm_count, smem, st, info)      \
                   : launch_path_t<RR, false, false>(p, cr, sm_count, smem, st, info)
  switch (R) {
    case 2: RTB_DISPATCH(2);
    case 4: RTB_DISPATCH(4);
    default: return cudaErrorInvalidValue;
  }
#undef RTB_DISPATCH
}

}  // namespace rtb
