// csrc/rt_wave.cu — the path-tracing hot path (sm_100a): everything executed per (pixel, sample)
// of src/cpu/main.cc:111-123.
//
// fused_kernel is a persistent kernel.  A WARP owns 64 path slots for the whole frame (two per
// lane).  The rays of a wave and their closest-hit records live in the warp's shared memory; the
// rest of a slot's state (throughput, item sums, item) in a pool in device memory (L2-resident).
// Every wave runs three phases, the divergent ones compacted across the warp's 64 slots:
//
//   shade     one slot per lane and step (two steps).  Finishes the ray segment traced by the
//             previous wave from its closest-hit record: sky or material scatter (main.cc:12-30,
//             material.h), then, if the path ended, path regeneration: next sample of the slot's
//             (pixel, slice) item or the next item of the frame's work queue (table-driven
//             decode), camera::get_ray (camera.h:28-35).
//   filter    two rays per lane.  Conservative f32 line-sphere test against ALL spheres in
//             reference order (hittable_list::hit visits every object, hittable_list.h:28-43):
//             packed FFMA2, sphere rows through the constant bank into uniform registers.  A
//             decision group of 4 rows that some ray passes is recorded as ONE entry
//             {slot, group} in the warp's candidate list, appended with ballots (no per-lane
//             counters, no divergent code).
//   resolve   the warp walks its candidate list 32 entries at a time, whatever slot they belong
//             to: f32 re-test of the group's 4 rows, then sphere::hit in f64 (sphere.h:24-51) on
//             the survivors, and the slot's closest-hit record is updated under the reference's
//             tie rule.  Lanes are balanced by ENTRY, not by slot.
//   bvh_trace optional replacement of filter + resolve: closest hit through the host-built BVH.
//
// Three kernels share these phases and differ in the filter:
//   fused_tc_kernel   (default, 32 .. 960 spheres) the filter as a 3xTF32 contraction on the tensor cores:
//                     filter_rays_tc, one CTA of five warpgroups per SM, sphere operand resident in shared
//                     memory, entries name single spheres;
//   fused_tcs_kernel  (more than 960 spheres) the same with the operand streamed through two chunk buffers;
//   fused_kernel      the FFMA2 loop above (RT_FLAG_SIMT_FILTER, fewer than 32 spheres, RT_FLAG_NO_FILTER) and
//                     the BVH form.
#include "rt_device.cuh"
#include "rt_kernels.h"
#include "rt_tc.cuh"

#include <cstring>

namespace rtb {

#ifndef RTW_BLOCK
#define RTW_BLOCK 256
#endif
#ifndef RTW_MINB
#define RTW_MINB 3
#endif
constexpr int kWBlock = RTW_BLOCK;
constexpr int kWarps = kWBlock / 32;
constexpr unsigned FULL = 0xffffffffu;
constexpr unsigned long long kInfKey = 0x7ff0000000000000ull;  // +inf as a double: no hit yet
constexpr uint32_t kNoItem = 0xffffffffu;

// Shared memory of one warp: the 64 rays of the wave, their closest-hit records, the candidate list.
struct __align__(16) WarpState {
  float4 ro[64];               // {ox, oy, oz, |d| of the reference's un-normalised ray; < 0 = no ray}
  float4 rd[64];               // {unit direction, index of the sphere the ray just left (int bits) or -1}
  unsigned long long key[64];  // root numerator of the closest hit so far (bits of a positive double)
  int id[64];                  // its sphere, -1 = none
  uint32_t list[kListSmem];    // {slot << kGroupBits | decision group}; the resolve does not depend on the order
  uint32_t cnt;                // entries appended by the tensor-core filter (lanes claim ranges with one atomic)
  uint32_t pad_[3];
};

// @phase helpers
__device__ __forceinline__ float2 wffma2(float2 a, float2 b, float2 c) {
  unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a),
                     rb = *reinterpret_cast<unsigned long long *>(&b),
                     rc = *reinterpret_cast<unsigned long long *>(&c), rd;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  return *reinterpret_cast<float2 *>(&rd);
}
__device__ __forceinline__ float2 wbc(float v) { return make_float2(v, v); }

// ================================================================ filter
// @phase filter-prologue
// Two rays per lane (slots lane and lane + 32), G sphere rows per decision group.  Returns the
// number of candidate entries the warp recorded (entries beyond the capacity are counted, not stored).
template <int G, bool CROWS>
__device__ __forceinline__ uint32_t filter_rays(const WaveParams &P, const ConstRows &CR, WarpState &W,
                                                uint32_t *glist, const uint32_t lane) {
  const uint32_t npad = P.scene.n_pad;
  const float INF = __int_as_float(0x7f800000);
  float2 DX, DY, DZ, OX2, OY2, OZ2, NOD, OOT;
  {
    const float4 o0 = W.ro[lane], d0 = W.rd[lane], o1 = W.ro[lane + 32], d1 = W.rd[lane + 32];
    DX = make_float2(d0.x, d1.x); DY = make_float2(d0.y, d1.y); DZ = make_float2(d0.z, d1.z);
    OX2 = make_float2(2.0f * o0.x, 2.0f * o1.x); OY2 = make_float2(2.0f * o0.y, 2.0f * o1.y);
    OZ2 = make_float2(2.0f * o0.z, 2.0f * o1.z);
    NOD = make_float2(-(o0.x * d0.x + o0.y * d0.y + o0.z * d0.z), -(o1.x * d1.x + o1.y * d1.y + o1.z * d1.z));
    // idle slots never pass the filter
    OOT.x = o0.w > 0.0f ? (o0.x * o0.x + o0.y * o0.y + o0.z * o0.z) * (1.0f - RTB_FILTER_SLACK) : INF;
    OOT.y = o1.w > 0.0f ? (o1.x * o1.x + o1.y * o1.y + o1.z * o1.z) * (1.0f - RTB_FILTER_SLACK) : INF;
  }
  uint32_t wcnt = 0;  // entries recorded so far (warp-uniform)
  const uint32_t lt = (1u << lane) - 1u;
  const uint32_t e0 = lane << kGroupBits, e1 = (lane + 32u) << kGroupBits;
  constexpr uint32_t gcap = 64u * kListGlobalPerSlot;

  // @phase main-loop
  // e = (c.d - o.d)^2 + (2 c.o + w) for one group of G rows; every FFMA2 multiplies by exactly
  // one sphere scalar (a uniform register when CROWS)
  auto group = [&](const float4 (&S)[G], float2 (&e)[G]) {
#pragma unroll
    for (int k = 0; k < G; ++k) {
      float2 b = wffma2(wbc(S[k].x), DX, wffma2(wbc(S[k].y), DY, wffma2(wbc(S[k].z), DZ, NOD)));
      float2 e1_ = wffma2(b, b, wbc(S[k].w));
      e[k] = wffma2(wbc(S[k].x), OX2, wffma2(wbc(S[k].y), OY2, wffma2(wbc(S[k].z), OZ2, e1_)));
    }
  };
  // decision for one group: max over rows, one compare per ray, one vote, one branch
  auto decide = [&](const float2 (&e)[G], uint32_t s0) {
    float mx = e[0].x, my = e[0].y;
#pragma unroll
    for (int k = 1; k < G; ++k) { mx = fmaxf(mx, e[k].x); my = fmaxf(my, e[k].y); }
    const bool p0 = mx >= OOT.x, p1 = my >= OOT.y;  // no short-circuit: independent compares
    if (__any_sync(FULL, p0 | p1)) {
      // @phase rare
      // rare path, warp-convergent: the rays whose group maximum passed append ONE entry
      // {slot, group} each to the warp's list; positions come from two ballots
      const unsigned b0 = __ballot_sync(FULL, p0), b1 = __ballot_sync(FULL, p1);
      const uint32_t g = s0 / (uint32_t)G;
      const uint32_t n0 = __popc(b0);
      const uint32_t pos0 = wcnt + __popc(b0 & lt), pos1 = wcnt + n0 + __popc(b1 & lt);
      if (wcnt <= kListSmem - 64u) {
        // the 64 entries a group can add fit the shared-memory list: no capacity checks
        if (p0) W.list[pos0] = e0 | g;
        if (p1) W.list[pos1] = e1 | g;
      } else {
        if (p0) {
          if (pos0 < kListSmem) W.list[pos0] = e0 | g;
          else if (pos0 - kListSmem < gcap) glist[pos0 - kListSmem] = e0 | g;
        }
        if (p1) {
          if (pos1 < kListSmem) W.list[pos1] = e1 | g;
          else if (pos1 - kListSmem < gcap) glist[pos1 - kListSmem] = e1 | g;
        }
      }
      wcnt += n0 + __popc(b1);
      // @phase main-loop
    }
  };

  const float4 *frow = P.scene.fast;
  float4 S[G];
  float2 ea[G], eb[G];
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
  return wcnt;
}

// ================================================================ resolve
// @phase resolve
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

__device__ __forceinline__ double next_up_pos(double x) {  // x > 0 finite
  return __longlong_as_double(__double_as_longlong(x) + 1ll);
}

// A ray as the exact test sees it: f32 origin and unit direction widened to double.
struct WRay {
  double ox, oy, oz, dx, dy, dz, a;
  double lo;  // smallest acceptable root numerator: t_min * a, or the next double above it when strict
};
__device__ __forceinline__ WRay widen_ray(const float4 o, const float4 d, const bool strict) {
  WRay r;
  r.ox = o.x; r.oy = o.y; r.oz = o.z; r.dx = d.x; r.dy = d.y; r.dz = d.z;
  r.a = r.dx * r.dx + r.dy * r.dy + r.dz * r.dz;
  const double tminA = 0.001 * (double)o.w * r.a;  // main.cc:19 t_min in unit-direction units, times a
  r.lo = strict ? next_up_pos(tminA) : tminA;
  return r;
}

// sphere::hit on numerators (root = N / a, a common to the ray), src/cpu/sphere.h:24-51, in the
// order-independent form: the root a sphere offers is its near root if that lies in the interval,
// else its far root (a far root can only be accepted when the near one lies before t_min, since
// near <= far); the closest-hit record then keeps the smallest root and, among equal roots, the
// LATER sphere (inclusive upper bound of hittable_list.h:34-39 + sphere.h:38-40) -- or, strict, the
// EARLIER one (the v4 interval of src/gpu/sphere.h:31-36: t_min < root < t_max).
__device__ __forceinline__ bool exact_root(const double4 S, const WRay &r, unsigned long long &key) {
  double ocx = r.ox - S.x, ocy = r.oy - S.y, ocz = r.oz - S.z;
  double half_b = ocx * r.dx + ocy * r.dy + ocz * r.dz;
  double c = ocx * ocx + ocy * ocy + ocz * ocz - S.w * S.w;
  double disc = half_b * half_b - r.a * c;
  if (disc < 0.0) return false;
  double sq = wave_dsqrt(disc);
  double N = -half_b - sq;
  if (N < r.lo) {
    N = -half_b + sq;
    if (N < r.lo) return false;
  }
  key = (unsigned long long)__double_as_longlong(N);
  return true;
}

// RT_FLAG_F32_HIT: sphere::hit in f32 on the reference's un-normalised direction, op for op
// src/gpu/sphere.h:15-47 (no self-skip, f32 hit point) -- reproduces the cancellation of
// |oc|^2 - r^2 on the r = 1000 ground sphere that the reference's CUDA program shows.  The key is
// the bit pattern of the (positive) f32 root.
__device__ __forceinline__ bool exact_root_f32(const double4 S, const float4 o, const float4 d,
                                               const bool strict, unsigned long long &key) {
  const float ddx = d.x * o.w, ddy = d.y * o.w, ddz = d.z * o.w;
  const float a = ddx * ddx + ddy * ddy + ddz * ddz;
  const float r = (float)S.w;
  const float ocx = o.x - (float)S.x, ocy = o.y - (float)S.y, ocz = o.z - (float)S.z;
  const float hb = ocx * ddx + ocy * ddy + ocz * ddz;
  const float c = ocx * ocx + ocy * ocy + ocz * ocz - r * r;
  const float disc = hb * hb - a * c;
  if (disc < 0.0f) return false;
  const float sq = __fsqrt_rn(disc);  // IEEE like the reference's build (this TU uses approximate / and sqrt)
  float root = __fdiv_rn(-hb - sq, a);
  if (strict ? !(0.001f < root) : root < 0.001f) {
    root = __fdiv_rn(-hb + sq, a);
    if (strict ? !(0.001f < root) : root < 0.001f) return false;
  }
  key = (unsigned long long)__float_as_uint(root);
  return true;
}

// closest-hit record of a slot <- one accepted root (called by one lane per slot at a time)
__device__ __forceinline__ void record_root(WarpState &W, const uint32_t slot, const unsigned long long key,
                                            const int s, const bool strict) {
  const unsigned long long cur = W.key[slot];
  const int cid = W.id[slot];
  if (key < cur || (key == cur && (strict ? s < cid : s > cid))) {
    W.key[slot] = key;
    W.id[slot] = s;
  }
}

// The warp walks its candidate list 32 entries at a time.  An entry names a slot and a decision
// group; the lane re-runs the f32 test on the group's G rows (same arithmetic as the filter), then
// one exact test per round on the rows that pass.  Lanes that hold accepted roots of the SAME
// slot in a round update its record one after the other (a single turn almost always).
template <int G, bool F32HIT, bool EXACT = false>
__device__ __forceinline__ void resolve_entries(const WaveParams &P, WarpState &W, const uint32_t *glist,
                                                const uint32_t n_entries, const double4 *exact,
                                                const uint32_t lane) {
  const bool strict = P.model_v4 != 0u;
  const float4 *fast = P.scene.fast;
  const uint32_t lt = (1u << lane) - 1u;
  for (uint32_t first = 0; first < n_entries; first += 32) {  // warp-uniform trip count
    const uint32_t idx = first + lane;
    uint32_t m = 0, slot = 0, s0 = 0;
    float4 o = make_float4(0.f, 0.f, 0.f, -1.f), d = make_float4(0.f, 1.f, 0.f, 0.f);
    if (idx < n_entries) {
      const uint32_t ent = idx < kListSmem ? W.list[idx] : glist[idx - kListSmem];
      slot = ent >> kGroupBits;
      o = W.ro[slot];
      d = W.rd[slot];
      if (EXACT) {
        // entries of the tensor-core filter name ONE sphere (the skip rule is applied there)
        s0 = ent & ((1u << kGroupBits) - 1u);
        m = 1u;
      } else {
        s0 = (ent & ((1u << kGroupBits) - 1u)) * (uint32_t)G;
        const uint32_t skip = (uint32_t)__float_as_int(d.w);  // the sphere the ray just left, or -1
        const float nod = -(o.x * d.x + o.y * d.y + o.z * d.z);
        const float ox2 = 2.0f * o.x, oy2 = 2.0f * o.y, oz2 = 2.0f * o.z;
        const float oot = (o.x * o.x + o.y * o.y + o.z * o.z) * (1.0f - RTB_FILTER_SLACK);
#pragma unroll
        for (int k = 0; k < G; ++k) {
          const float4 S = __ldg(fast + s0 + k);
          const float b = fmaf(S.x, d.x, fmaf(S.y, d.y, fmaf(S.z, d.z, nod)));
          const float e = fmaf(S.x, ox2, fmaf(S.y, oy2, fmaf(S.z, oz2, fmaf(b, b, S.w))));
          if (e >= oot && s0 + (uint32_t)k != skip) m |= 1u << k;
        }
      }
    }
    const WRay ray = widen_ray(o, d, strict);
    while (__any_sync(FULL, m != 0u)) {
      bool hit = false;
      unsigned long long key = 0;
      int s = -1;
      if (m) {
        s = (int)(s0 + (uint32_t)__ffs((int)m) - 1u);
        m &= m - 1u;
        hit = F32HIT ? exact_root_f32(exact[s], o, d, strict, key) : exact_root(exact[s], ray, key);
      }
      // lanes with accepted roots of one slot take turns (lanes without one match nobody)
      const unsigned peers = __match_any_sync(FULL, hit ? slot : 64u + lane);
      const uint32_t turn = __popc(peers & lt);
      const uint32_t turns = __reduce_max_sync(FULL, turn);
      for (uint32_t t = 0; t <= turns; ++t) {
        if (hit && turn == t) record_root(W, slot, key, s, strict);
        __syncwarp();
      }
    }
  }
}

// Fallback: every lane scans ALL spheres for its own two slots.  Taken when the warp's candidate
// list overflowed (rays along long rows of spheres) -- with the f32 test first -- and under
// RT_FLAG_NO_FILTER, without it: hittable_list::hit as written (hittable_list.h:28-43).
template <bool F32HIT>
__device__ __forceinline__ void resolve_scan(const WaveParams &P, WarpState &W, const double4 *exact,
                                             const uint32_t lane) {
  const bool strict = P.model_v4 != 0u;
  const float4 *fast = P.scene.fast;
#pragma unroll 1
  for (int r = 0; r < 2; ++r) {
    const uint32_t slot = lane + 32u * r;
    const float4 o = W.ro[slot], d = W.rd[slot];
    if (!(o.w > 0.0f)) continue;
    const uint32_t skip = (uint32_t)__float_as_int(d.w);
    const float nod = -(o.x * d.x + o.y * d.y + o.z * d.z);
    const float ox2 = 2.0f * o.x, oy2 = 2.0f * o.y, oz2 = 2.0f * o.z;
    const float oot = (o.x * o.x + o.y * o.y + o.z * o.z) * (1.0f - RTB_FILTER_SLACK);
    const WRay ray = widen_ray(o, d, strict);
    for (uint32_t s = 0; s < P.scene.n; ++s) {
      if (!P.no_filter) {
        const float4 S = __ldg(fast + s);
        const float b = fmaf(S.x, d.x, fmaf(S.y, d.y, fmaf(S.z, d.z, nod)));
        const float e = fmaf(S.x, ox2, fmaf(S.y, oy2, fmaf(S.z, oz2, fmaf(b, b, S.w))));
        if (!(e >= oot)) continue;
      }
      if (s == skip) continue;
      unsigned long long key;
      if (F32HIT ? exact_root_f32(exact[s], o, d, strict, key) : exact_root(exact[s], ray, key))
        record_root(W, slot, key, (int)s, strict);
    }
  }
}

// ================================================================ shade
// @phase fetch
// q / d and q % d for quotients below 2^16: float estimate (within 1), fixed up exactly
__device__ __forceinline__ void divmod_small(const uint32_t q, const uint32_t d, uint32_t &quo, uint32_t &rem) {
  uint32_t e = (uint32_t)(__uint2float_rz(q) * __frcp_rn(__uint2float_rn(d)));
  int r = (int)(q - e * d);
  if (r < 0) { e -= 1u; r += (int)d; }
  else if ((uint32_t)r >= d) { e += 1u; r -= (int)d; }
  quo = e; rem = (uint32_t)r;
}

// One slot, one wave.  (1) Finish the ray segment traced by the last wave from the slot's
// closest-hit record: sky on a miss.  (2) ONE Philox block per slot and wave: the camera sample of
// a new path (miss / empty slot) or the scatter sample of the hit.  (3) camera::get_ray or
// material::scatter.  A path that ends at a hit (absorbed, bounce limit) leaves the slot without a
// ray for one wave and regenerates in the next one.  Returns 0 once the slot is retired.
template <bool F32HIT>
__device__ __forceinline__ uint32_t shade_slot(const WaveParams &P, WarpState &W, const uint32_t slot,
                                               const uint32_t i, const double4 *exact, const float4 *mat,
                                               const float2 *aux, const uint32_t npix,
                                               const uint32_t total_items) {
  // @phase slot-state
  float4 T = P.thr[i];  // {tr, tg, tb, depth}
  int depth = __float_as_int(T.w);
  if (depth < 0) return 0u;  // slot retired: the frame's work queue is empty
  uint4 st = P.it[i];   // {item = slice << kItemShift | pixel (kNoItem = none), samp, samp_end, -}
  float4 A = P.acc[i];  // {sum r, sum g, sum b, ray segments} of the slot's item
  float4 RO = make_float4(0.f, 0.f, 0.f, -1.f), RD = make_float4(0.f, 1.f, 0.f, 0.f);
  int id = -1;
  double hpx = 0, hpy = 0, hpz = 0;  // hit point (f64)

  if (depth > 0) {
    // @phase sky-hitpoint
    RO = W.ro[slot];
    RD = W.rd[slot];
    A.w += 1.0f;  // one ray_color invocation (src/cpu/main.cc:12)
    id = W.id[slot];
    if (id < 0) {
      // sky, src/cpu/main.cc:27-29 (direction already unit length)
      float t = 0.5f * (RD.y + 1.0f);
      float w = 1.0f - t;
      A.x += T.x * (w + t * 0.5f);
      A.y += T.y * (w + t * 0.7f);
      A.z += T.z * (w + t * 1.0f);
      depth = 0;
    } else if (F32HIT) {
      // ray::at in f32 (src/gpu/ray.h), kept as the slot's f64 hit point
      const float bestT = __uint_as_float((uint32_t)W.key[slot]);
      const float ddx = RD.x * RO.w, ddy = RD.y * RO.w, ddz = RD.z * RO.w;
      hpx = (double)fmaf(bestT, ddx, RO.x); hpy = (double)fmaf(bestT, ddy, RO.y); hpz = (double)fmaf(bestT, ddz, RO.z);
    } else {
      const double bestN = __longlong_as_double((long long)W.key[slot]);
      const double ox = RO.x, oy = RO.y, oz = RO.z, dx = RD.x, dy = RD.y, dz = RD.z;
      const double a = dx * dx + dy * dy + dz * dz;
      const double t = bestN * (2.0 - a);  // 1/a = 2 - a to within eps^2 (a = 1 + eps)
      hpx = ox + t * dx; hpy = oy + t * dy; hpz = oz + t * dz;
    }
  }

  const bool regen = depth == 0;
  if (regen) {
    // @phase fetch
    // ------------------------------------------------------------ next sample / next item
    while (st.y == st.z) {  // item exhausted: flush it and fetch the next
      if (st.x != kNoItem) {
        P.accum[(size_t)(st.x >> kItemShift) * npix + (st.x & kItemPixelMask)] = A;
        st.x = kNoItem;
      }
      // warp-aggregated pop: the lanes that need an item right now share ONE atomic (single-address
      // atomics top out near 1 G/s, and a frame ends on one-sample items)
      const unsigned am = __activemask();
      const unsigned lane = threadIdx.x & 31u;
      const int leader = __ffs((int)am) - 1;
      unsigned long long itq = 0;
      if ((int)lane == leader) itq = atomicAdd(P.counter, (unsigned long long)__popc(am));
      itq = __shfl_sync(am, itq, leader) + (unsigned long long)__popc(am & ((1u << lane) - 1u));
      if (itq >= (unsigned long long)total_items) { depth = -1; break; }
      // hand-out order: first every slice of the heavy pixels, then the other pixels slice by
      // slice (largest slices first, heaviest pixels first within a slice).  The queue holds fewer
      // than 2^32 items (2^26 pixels x 64 slices), quotients are slice positions (< 64).
      uint32_t q = (uint32_t)itq, ord, px;
      if (P.order) {
        const uint32_t nh = __ldg(P.heavy);
        const uint32_t heavy_items = nh * P.n_orders;
        uint32_t dvs = nh, off = 0;
        if (q >= heavy_items) { q -= heavy_items; dvs = npix - nh; off = nh; }
        divmod_small(q, dvs, ord, px);
        px = __ldg(P.order + off + px);
      } else {
        divmod_small(q, npix, ord, px);
      }
      const uint32_t sl = P.order_slice[ord];
      st.x = sl << kItemShift | px;
      st.y = P.slice_bounds[sl];
      st.z = P.slice_bounds[sl + 1u];
      A = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }

  if (depth >= 0) {
    // @phase philox
    // ------------------------------------------------------------ the wave's random numbers
    // stream: key = (pixel, sample), counter slot 0 = camera, k >= 1 = scatter at the k-th hit
    const uint32_t px = st.x & kItemPixelMask;
    const uint32_t hitno = P.max_depth - (uint32_t)depth + 1u;
    const float4 xi = philox_uniforms(px, P.sample_offset + (regen ? st.y : st.y - 1u),
                                      regen ? 0u : hitno, P.seed);
    if (regen) {
      // @phase camera
      // camera::get_ray (src/cpu/camera.h:28-35) with pixel jitter (main.cc:116-117)
      uint32_t row, ii;
      divmod_small(px, P.width, row, ii);  // rows < 2^16 (checked by rt_frame_create)
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
      // @phase scatter
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
        } else if (kind >= 3u) {
          // the book's earlier diffuse rules: a point INSIDE the unit ball, added to the normal
          // (archive/listing33/main.cc:12-13) or flipped into the normal's hemisphere
          // (archive/listing40/main.cc:26, random_in_hemisphere of listing39/vec3.h:110-116)
          const float k = cbrtf(xi.z);
          ux *= k; uy *= k; uz *= k;
          if (kind == 3u) {
            sx = nx + ux; sy = ny + uy; sz = nz + uz;
          } else {
            const bool same = (ux * nx + uy * ny + uz * nz) > 0.0f;
            sx = same ? ux : -ux; sy = same ? uy : -uy; sz = same ? uz : -uz;
          }
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

  // @phase slot-state
  W.ro[slot] = RO;
  if (depth > 0) W.rd[slot] = RD;
  W.key[slot] = kInfKey;
  W.id[slot] = -1;
  T.w = __int_as_float(depth);
  P.thr[i] = T;
  P.acc[i] = A;
  P.it[i] = st;
  return depth >= 0 ? 1u : 0u;  // still part of the frame (with or without a ray this wave)
}

// @phase stage
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

// ================================================================ BVH traversal (optional mode)
// @phase bvh
// Closest hit of the lane's two rays through the host-built hierarchy (rt_bvh.cu) with the
// REFERENCE's result: the linear scan of hittable_list::hit keeps the smallest root and, among
// equal roots, the LARGEST index (sphere.h:38-40 + hittable_list.h:34-39) -- a rule that does not
// depend on the visiting order, so leaves may be visited in any order.  Boxes are tested in f32
// with outward slack; a node is skipped only if it lies entirely beyond the best root or before
// t_min.  Writes the slot's closest-hit record.
__device__ __forceinline__ void bvh_trace(const WaveParams &P, WarpState &W, const double4 *exact,
                                          const uint32_t lane) {
  const BvhNode *nodes = P.scene.bvh_nodes;
  const uint32_t *prim = P.scene.bvh_prim;
  const float4 *fast = P.scene.fast;
  const bool strict = P.model_v4 != 0u;
#pragma unroll 1
  for (int r = 0; r < 2; ++r) {
    const uint32_t slot = lane + 32u * r;
    const float4 o = W.ro[slot];
    if (!(o.w > 0.0f)) continue;  // no ray this wave
    const float4 d = W.rd[slot];
    const WRay ray = widen_ray(o, d, strict);
    unsigned long long best = kInfKey;
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
            unsigned long long key;
            if (exact_root(exact[s], ray, key) &&
                (key < best || (key == best && (strict ? (int)s < id : (int)s > id)))) {
              best = key; id = (int)s;
              tbest = (float)(__longlong_as_double((long long)key) * 1.00001) + 1e-6f;
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
    W.key[slot] = best;
    W.id[slot] = id;
  }
}

// ================================================================ the persistent kernel
// @phase kernel-top
// CTAs per SM: the brute-force form needs its registers for the filter loop (3 x 256 threads); the
// BVH form is bound by load latency and takes a fourth CTA.
constexpr int kBvhMinB = RTW_MINB + 1;
template <int G, bool CROWS, bool SMEM, bool BVH, bool F32HIT = false>
__global__ void __launch_bounds__(kWBlock, BVH ? kBvhMinB : RTW_MINB)
fused_kernel(const __grid_constant__ WaveParams P, const __grid_constant__ ConstRows CR) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t s_bar;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  WarpState &W = reinterpret_cast<WarpState *>(smem_raw)[warp];
  const double4 *exact = P.scene.exact;
  const float4 *mat = P.scene.mat;
  const float2 *aux = P.scene.aux;
  if (SMEM) stage_hit_rows(P, smem_raw + kWarps * sizeof(WarpState), &s_bar, exact, mat, aux);
  const uint32_t npix = P.width * P.height;
  const uint32_t total_items = npix * P.n_orders;  // < 2^32: 2^26 pixels x 64 slices at most
  const uint32_t stride = P.n_slots / 2u;
  const uint32_t base = blockIdx.x * kWBlock + threadIdx.x;  // grid * block == stride
  uint32_t *glist = P.glist + (size_t)(blockIdx.x * kWarps + warp) * (64u * kListGlobalPerSlot);
#pragma unroll 1
  for (int r = 0; r < 2; ++r) {
    const uint32_t i = base + r * stride, slot = lane + 32u * r;
    P.thr[i] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));  // depth 0: needs a path
    P.acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    P.it[i] = make_uint4(kNoItem, 0u, 0u, 0u);
    W.ro[slot] = make_float4(0.f, 0.f, 0.f, -1.f);
    W.rd[slot] = make_float4(0.f, 1.f, 0.f, __int_as_float(-1));
    W.key[slot] = kInfKey;
    W.id[slot] = -1;
  }
  for (;;) {
    uint32_t live = 0;
#pragma unroll 1
    for (int r = 0; r < 2; ++r)
      live += shade_slot<F32HIT>(P, W, lane + 32u * r, base + r * stride, exact, mat, aux, npix, total_items);
    if (__all_sync(FULL, live == 0u)) break;  // warp-uniform exit keeps the loop convergent
    __syncwarp();  // the wave's rays and the cleared records are visible to the whole warp
    if (BVH) {
      bvh_trace(P, W, exact, lane);
    } else {
      uint32_t n_entries = 0xffffffffu;  // RT_FLAG_NO_FILTER: no list, every lane scans every sphere
      if (!P.no_filter) n_entries = filter_rays<G, CROWS>(P, CR, W, glist, lane);
      __syncwarp();
      if (n_entries <= kListSmem + 64u * kListGlobalPerSlot) resolve_entries<G, F32HIT>(P, W, glist, n_entries, exact, lane);
      else resolve_scan<F32HIT>(P, W, exact, lane);
    }
    __syncwarp();
  }
}

// ================================================================ tensor-core filter (tcgen05)
// @phase tc-filter
// The brute-force filter as one tf32 contraction (formulation and operand layout: rt_tc.cu).  A
// WARPGROUP works together: its 128 threads hold the 128 rays of an MMA's M dimension, thread <-> TMEM
// lane.  Per block of rays (each lane's slot 0, then slot 1): every thread writes its ray's 32 feature
// slots to its lane of the A columns; per tile of 64 spheres one elected thread issues the four K
// steps and commits them to the group's mbarrier; every thread then reads ITS ray's 64 values back,
// the group barrier frees D for the next tile, and the sign bits become candidate entries
// {slot, sphere} appended to the warp's list with ballots -- while the next tile's MMAs run.
#ifndef RTW_TC_BLOCK
#define RTW_TC_BLOCK 640
#endif

constexpr int kTcBlock = RTW_TC_BLOCK, kTcWarps = kTcBlock / 32, kTcGroups = kTcBlock / 128;
constexpr uint32_t kTcGroupCols = 32u + (uint32_t)kTcTile;  // TMEM columns of a warpgroup: ray features + one accumulator tile
static_assert(kTcGroups * kTcGroupCols <= 512u, "the warpgroups' tensor-memory shares exceed the SM's 512 columns");

__device__ __forceinline__ void group_sync(const uint32_t gid) {
  asm volatile("bar.sync %0, 128;" ::"r"(gid + 1u) : "memory");
}
__device__ __forceinline__ bool group_or(const uint32_t gid, const bool v) {
  uint32_t r;
  asm volatile(
      "{\n\t.reg .pred p, q;\n\tsetp.ne.u32 q, %2, 0;\n\tbar.red.or.pred p, %1, 128, q;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(r)
      : "r"(gid + 1u), "r"((uint32_t)v)
      : "memory");
  return r != 0u;
}

__device__ __forceinline__ uint32_t filter_rays_tc(const WaveParams &P, WarpState &W, uint32_t *glist,
                                                   const uint32_t lane, const uint32_t gid, const bool issuer,
                                                   uint64_t *bar, uint32_t &parity, const uint32_t tmem_g,
                                                   const uint64_t desc0) {
  const uint32_t tiles = P.scene.tc_tiles;
  const uint32_t lane_base = (((threadIdx.x >> 5) & 3u) * 32u) << 16;
  const uint32_t a_tmem = tmem_g, d_tmem = tmem_g + 32u;
  constexpr uint32_t gcap = 64u * kListGlobalPerSlot;
  // the elected thread starts the four K steps of tile `t` (descriptors of consecutive K slabs are
  // kTcSlabUnits apart in the address field) and commits them to the group's mbarrier
  auto issue = [&](const uint32_t t) {
    if (issuer) {
      tc_fence_after();
      const uint64_t dt = desc0 + (uint64_t)(t * 4u) * (uint64_t)kTcSlabUnits;
#pragma unroll
      for (uint32_t j = 0; j < 4u; ++j) tc_mma_ts(d_tmem, a_tmem + j * 8u, dt + (uint64_t)j * (uint64_t)kTcSlabUnits, kTcIdesc, j);
      tc_commit(bar);
    }
  };
  if (lane == 0) W.cnt = 0u;  // ordered before the first append by the group barriers below
#pragma unroll 1
  for (uint32_t blk = 0; blk < 2u; ++blk) {
    const uint32_t slot = lane + 32u * blk;
    const float4 o = W.ro[slot], d = W.rd[slot];
    // the sphere the ray just left (or -1) never passes: its tile and the two words that clear its bit
    const uint32_t skip = (uint32_t)__float_as_int(d.w);
    const uint32_t skip_t = skip / (uint32_t)kTcTile, skip_b = skip - skip_t * (uint32_t)kTcTile;
    const uint32_t keep0 = skip_b < 32u ? ~(1u << skip_b) : ~0u;
    const uint32_t keep1 = skip_b < 32u ? ~0u : ~(1u << (skip_b - 32u));
    {
      uint32_t sv[kTcK];
      tc_ray_slots(o, d, sv);
      TC_ST32(a_tmem + lane_base, sv);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    tc_fence_before();
    group_sync(gid);  // the 128 rays' features are in tensor memory
    issue(0u);
#pragma unroll 1
    for (uint32_t t = 0; t < tiles; ++t) {
      // bounded wait: a tensor-core fault must not hang the GPU; after one expiry (bit 31) the thread
      // stops waiting, the frame finishes with garbage and the host reports the status word
      if (!(parity & 0x80000000u) && !mbar_wait_bounded(bar, parity & 1u)) {
        atomicExch(P.status, 1u);
        parity |= 0x80000000u;
      }
      parity ^= 1u;
      tc_fence_after();
      uint32_t v[kTcTile];
      uint32_t p0, p1;
      // the row comes back in two halves: the sign bits of the first are folded while the second is in flight
      TC_LD32(d_tmem + lane_base, v, 0);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      TC_LD_TILE_HI(d_tmem + lane_base, v);
      p0 = tc_pass_bits_lo(v);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      tc_fence_before();
      group_sync(gid);  // D is free: the next tile's MMAs run under the mask and append work below
      if (t + 1u < tiles) issue(t + 1u);
      p1 = tc_pass_bits_hi(v);
      if (t == skip_t) { p0 &= keep0; p1 &= keep1; }
      const uint32_t ebase = (slot << kGroupBits) + t * (uint32_t)kTcTile;
      // candidates -> the warp's list: the lane claims a range with one shared-memory atomic
      if ((p0 | p1) != 0u) {
        uint32_t pos = atomicAdd(&W.cnt, (uint32_t)(__popc(p0) + __popc(p1)));
        auto put = [&](uint32_t p, const uint32_t eb) {
          while (p != 0u) {
            const uint32_t e = eb + (uint32_t)__ffs((int)p) - 1u;
            p &= p - 1u;
            if (pos < kListSmem) W.list[pos] = e;
            else if (pos - kListSmem < gcap) glist[pos - kListSmem] = e;
            ++pos;
          }
        };
        put(p0, ebase);
        put(p1, ebase + 32u);
      }
    }
  }
  __syncwarp();
  return W.cnt;
}

// The persistent kernel with the tensor-core filter: ONE CTA of kTcBlock threads (five warpgroups) per
// SM -- the sphere operand (8 KB per 64 spheres) lives once per SM in shared memory and the SM's
// 512 TMEM columns are split between the warpgroups (32 columns of ray features + one accumulator tile each).
// Shade and resolve are the warp-level phases of fused_kernel, unchanged.
template <bool SMEM, bool F32HIT>
__global__ void __launch_bounds__(kTcBlock, 1)
fused_tc_kernel(const __grid_constant__ WaveParams P) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ __align__(8) uint64_t s_mma[kTcGroups];
  __shared__ uint32_t s_tmem;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, gid = warp >> 2;
  const uint32_t op_floats = P.scene.tc_tiles * 4u * (uint32_t)kTcSlabFloats;
  float *s_op = reinterpret_cast<float *>(smem_raw);
  unsigned char *after_op = smem_raw + (size_t)op_floats * 4u;
  WarpState &W = reinterpret_cast<WarpState *>(after_op)[warp];
  {
    const float4 *src = reinterpret_cast<const float4 *>(P.scene.tc_op);
    float4 *dst = reinterpret_cast<float4 *>(s_op);
    for (uint32_t i = threadIdx.x; i < op_floats / 4u; i += kTcBlock) dst[i] = __ldg(src + i);
  }
  if (threadIdx.x < kTcGroups) mbar_init(&s_mma[threadIdx.x], 1);
  if (warp == 0) tmem_alloc(&s_tmem, 512);
  // the operand was written through the generic proxy; the tensor core reads it through the async one
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  const double4 *exact = P.scene.exact;
  const float4 *mat = P.scene.mat;
  const float2 *aux = P.scene.aux;
  if (SMEM) stage_hit_rows(P, after_op + kTcWarps * sizeof(WarpState), &s_bar, exact, mat, aux);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_g = s_tmem + gid * kTcGroupCols;
  const bool issuer = (threadIdx.x & 127u) == 0u;
  const uint32_t npix = P.width * P.height;
  const uint32_t total_items = npix * P.n_orders;  // < 2^32: 2^26 pixels x 64 slices at most
  const uint32_t stride = P.n_slots / 2u;
  const uint32_t base = blockIdx.x * kTcBlock + threadIdx.x;  // grid * block == stride
  uint32_t *glist = P.glist + (size_t)(blockIdx.x * kTcWarps + warp) * (64u * kListGlobalPerSlot);
#pragma unroll 1
  for (int r = 0; r < 2; ++r) {
    const uint32_t i = base + r * stride, slot = lane + 32u * r;
    P.thr[i] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));  // depth 0: needs a path
    P.acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    P.it[i] = make_uint4(kNoItem, 0u, 0u, 0u);
    W.ro[slot] = make_float4(0.f, 0.f, 0.f, -1.f);
    W.rd[slot] = make_float4(0.f, 1.f, 0.f, __int_as_float(-1));
    W.key[slot] = kInfKey;
    W.id[slot] = -1;
  }
  const uint64_t desc0 = tc_smem_desc(s_op);
  uint32_t parity = 0;
  for (;;) {
    uint32_t live = 0;
#pragma unroll 1
    for (int r = 0; r < 2; ++r)
      live += shade_slot<F32HIT>(P, W, lane + 32u * r, base + r * stride, exact, mat, aux, npix, total_items);
    if (!group_or(gid, live != 0u)) break;  // the four warps of a group leave together
    __syncwarp();
    const uint32_t n_entries = filter_rays_tc(P, W, glist, lane, gid, issuer, &s_mma[gid], parity, tmem_g, desc0);
    __syncwarp();
    if (n_entries <= kListSmem + 64u * kListGlobalPerSlot) resolve_entries<RTW_G, F32HIT, true>(P, W, glist, n_entries, exact, lane);
    else resolve_scan<F32HIT>(P, W, exact, lane);
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(s_tmem, 512);
}

// ================================================================ tensor-core filter, streamed operand
// Scenes beyond kTcMaxSpheres: the sphere operand (128 B per sphere) no longer fits a CTA's shared memory, so
// every warpgroup streams it through kTcsBufs tile buffers of its own with bulk (TMA) copies: while the MMAs of
// tile t run, the next tiles are already in the other buffers and the copy of tile t + kTcsBufs starts as soon
// as the group has read the accumulator of tile t (whose MMAs -- the readers of that buffer -- are then complete).
// The group's issuer thread is the producer; the groups stay independent of each other, as in the resident form.
constexpr uint32_t kTcsTileBytes = 4u * (uint32_t)kTcSlabFloats * 4u;  // 8 KB per 64 spheres
#ifndef RTW_TCS_BUFS
#define RTW_TCS_BUFS 2
#endif
constexpr uint32_t kTcsBufs = RTW_TCS_BUFS;  // tile buffers per warpgroup (copies run kTcsBufs - 1 rounds ahead)

__device__ __forceinline__ uint32_t filter_rays_tc_stream(const WaveParams &P, WarpState &W, uint32_t *glist,
                                                          const uint32_t lane, const uint32_t gid, const bool issuer,
                                                          uint64_t *bar, uint32_t &parity, const uint32_t tmem_g,
                                                          unsigned char *bufs, uint64_t *full, uint32_t &fpar) {
  const uint32_t tiles = P.scene.tc_tiles;
  const uint32_t lane_base = (((threadIdx.x >> 5) & 3u) * 32u) << 16;
  const uint32_t a_tmem = tmem_g, d_tmem = tmem_g + 32u;
  constexpr uint32_t gcap = 64u * kListGlobalPerSlot;
  const uint64_t desc_b0 = tc_smem_desc(bufs);
  // producer side (issuer only): tile t of the scene -> buffer t % kTcsBufs
  auto load = [&](const uint32_t t) {
    const uint32_t b = t % kTcsBufs;
    mbar_expect_tx(full + b, kTcsTileBytes);
    bulk_g2s(bufs + (size_t)b * kTcsTileBytes,
             reinterpret_cast<const unsigned char *>(P.scene.tc_op) + (size_t)t * kTcsTileBytes, kTcsTileBytes, full + b);
  };
  // the tile's bytes have landed, then its four K steps; bit b of fpar is the phase of buffer b
  auto issue = [&](const uint32_t t) {
    if (issuer) {
      const uint32_t b = t % kTcsBufs;
      mbar_wait(full + b, (fpar >> b) & 1u);
      fpar ^= 1u << b;
      tc_fence_after();
      const uint64_t dt = desc_b0 + (uint64_t)(b * 4u) * (uint64_t)kTcSlabUnits;
#pragma unroll
      for (uint32_t j = 0; j < 4u; ++j) tc_mma_ts(d_tmem, a_tmem + j * 8u, dt + (uint64_t)j * (uint64_t)kTcSlabUnits, kTcIdesc, j);
      tc_commit(bar);
    }
  };
  if (lane == 0) W.cnt = 0u;  // ordered before the first append by the group barriers below
#pragma unroll 1
  for (uint32_t blk = 0; blk < 2u; ++blk) {
    // the MMAs of the block before are complete: both buffers are free
    if (issuer)
      for (uint32_t t = 0; t < kTcsBufs && t < tiles; ++t) load(t);
    const uint32_t slot = lane + 32u * blk;
    const float4 o = W.ro[slot], d = W.rd[slot];
    const uint32_t skip = (uint32_t)__float_as_int(d.w);
    const uint32_t skip_t = skip / (uint32_t)kTcTile, skip_b = skip - skip_t * (uint32_t)kTcTile;
    const uint32_t keep0 = skip_b < 32u ? ~(1u << skip_b) : ~0u;
    const uint32_t keep1 = skip_b < 32u ? ~0u : ~(1u << (skip_b - 32u));
    {
      uint32_t sv[kTcK];
      tc_ray_slots(o, d, sv);
      TC_ST32(a_tmem + lane_base, sv);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    tc_fence_before();
    group_sync(gid);
    issue(0u);
#pragma unroll 1
    for (uint32_t t = 0; t < tiles; ++t) {
      if (!(parity & 0x80000000u) && !mbar_wait_bounded(bar, parity & 1u)) {
        atomicExch(P.status, 1u);
        parity |= 0x80000000u;
      }
      parity ^= 1u;
      tc_fence_after();
      uint32_t vv[kTcTile];
      uint32_t p0, p1;
      TC_LD32(d_tmem + lane_base, vv, 0);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      TC_LD_TILE_HI(d_tmem + lane_base, vv);
      p0 = tc_pass_bits_lo(vv);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      tc_fence_before();
      group_sync(gid);
      // D is free and the MMAs of tile t are complete: their buffer takes tile t + 2, tile t + 1 starts
      if (issuer && t + kTcsBufs < tiles) load(t + kTcsBufs);
      if (t + 1u < tiles) issue(t + 1u);
      p1 = tc_pass_bits_hi(vv);
      if (t == skip_t) { p0 &= keep0; p1 &= keep1; }
      const uint32_t ebase = (slot << kGroupBits) + t * (uint32_t)kTcTile;
      if ((p0 | p1) != 0u) {
        uint32_t pos = atomicAdd(&W.cnt, (uint32_t)(__popc(p0) + __popc(p1)));
        auto put = [&](uint32_t p, const uint32_t eb) {
          while (p != 0u) {
            const uint32_t e = eb + (uint32_t)__ffs((int)p) - 1u;
            p &= p - 1u;
            if (pos < kListSmem) W.list[pos] = e;
            else if (pos - kListSmem < gcap) glist[pos - kListSmem] = e;
            ++pos;
          }
        };
        put(p0, ebase);
        put(p1, ebase + 32u);
      }
    }
  }
  __syncwarp();
  return W.cnt;
}

template <bool F32HIT>
__global__ void __launch_bounds__(kTcBlock, 1)
fused_tcs_kernel(const __grid_constant__ WaveParams P) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t s_mma[kTcGroups];
  __shared__ __align__(8) uint64_t s_full[kTcGroups][kTcsBufs];
  __shared__ uint32_t s_tmem;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, gid = warp >> 2;
  WarpState &W = reinterpret_cast<WarpState *>(smem_raw + (size_t)kTcGroups * kTcsBufs * kTcsTileBytes)[warp];
  if (threadIdx.x < kTcGroups) {
    mbar_init(&s_mma[threadIdx.x], 1);
    for (uint32_t b = 0; b < kTcsBufs; ++b) mbar_init(&s_full[threadIdx.x][b], 1);
  }
  if (warp == 0) tmem_alloc(&s_tmem, 512);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the barriers are seen by the bulk copies
  const double4 *exact = P.scene.exact;
  const float4 *mat = P.scene.mat;
  const float2 *aux = P.scene.aux;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_g = s_tmem + gid * kTcGroupCols;
  const bool issuer = (threadIdx.x & 127u) == 0u;
  const uint32_t npix = P.width * P.height;
  const uint32_t total_items = npix * P.n_orders;
  const uint32_t stride = P.n_slots / 2u;
  const uint32_t base = blockIdx.x * kTcBlock + threadIdx.x;
  uint32_t *glist = P.glist + (size_t)(blockIdx.x * kTcWarps + warp) * (64u * kListGlobalPerSlot);
#pragma unroll 1
  for (int r = 0; r < 2; ++r) {
    const uint32_t i = base + r * stride, slot = lane + 32u * r;
    P.thr[i] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));
    P.acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    P.it[i] = make_uint4(kNoItem, 0u, 0u, 0u);
    W.ro[slot] = make_float4(0.f, 0.f, 0.f, -1.f);
    W.rd[slot] = make_float4(0.f, 1.f, 0.f, __int_as_float(-1));
    W.key[slot] = kInfKey;
    W.id[slot] = -1;
  }
  unsigned char *bufs = smem_raw + (size_t)gid * kTcsBufs * kTcsTileBytes;
  uint32_t parity = 0, fpar = 0;
  for (;;) {
    uint32_t live = 0;
#pragma unroll 1
    for (int r = 0; r < 2; ++r)
      live += shade_slot<F32HIT>(P, W, lane + 32u * r, base + r * stride, exact, mat, aux, npix, total_items);
    if (!group_or(gid, live != 0u)) break;  // the four warps of a group leave together
    __syncwarp();
    const uint32_t n_entries = filter_rays_tc_stream(P, W, glist, lane, gid, issuer, &s_mma[gid], parity, tmem_g,
                                                     bufs, s_full[gid], fpar);
    __syncwarp();
    if (n_entries <= kListSmem + 64u * kListGlobalPerSlot) resolve_entries<RTW_G, F32HIT, true>(P, W, glist, n_entries, exact, lane);
    else resolve_scan<F32HIT>(P, W, exact, lane);
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(s_tmem, 512);
}

// ================================================================ host side
size_t wave_pool_bytes(uint32_t n_slots) {
  // thr, acc (float4) + it (uint4) + the candidate entries beyond the shared-memory lists
  return (size_t)n_slots * (3 * 16 + 4 * (size_t)kListGlobalPerSlot);
}

uint32_t wave_default_slots(int sm_count) {
  // SMs x resident CTAs (the BVH form takes one more) x lanes x 2 slots per lane
  return (uint32_t)sm_count * (RTW_MINB + 1) * kWBlock * 2u;
}

static void wave_carve(WaveParams &p, void *pool, uint32_t n_slots) {
  unsigned char *b = static_cast<unsigned char *>(pool);
  size_t n = n_slots;
  p.n_slots = n_slots;
  p.thr = reinterpret_cast<float4 *>(b); b += n * 16;
  p.acc = reinterpret_cast<float4 *>(b); b += n * 16;
  p.it = reinterpret_cast<uint4 *>(b); b += n * 16;
  p.glist = reinterpret_cast<uint32_t *>(b);
}

template <int G>
static cudaError_t launch_fused_t(WaveParams &p, void *pool, uint32_t pool_slots, const ConstRows &cr,
                                  bool crows, bool in_smem, bool bvh, bool f32hit, size_t smem,
                                  int sm_count, int *grid_out, cudaStream_t st) {
#define RTW_FUSED(CR_, SM_, BV_, ...)                                                              \
  do {                                                                                             \
    auto k = fused_kernel<G, CR_, SM_, BV_, ##__VA_ARGS__>;                                        \
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e != cudaSuccess) return e;                                                                \
    int per_sm = 0;                                                                                \
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, kWBlock, smem);                  \
    if (e != cudaSuccess) return e;                                                                \
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;                                          \
    /* persistent grid: every CTA must be resident (one full wave, a multiple of the SM count) */  \
    int want = BV_ ? kBvhMinB : RTW_MINB;                                                          \
    while (want > 1 && (size_t)sm_count * want * kWBlock * 2 > pool_slots) --want;                 \
    const int grid = sm_count * (per_sm < want ? per_sm : want);                                   \
    const uint32_t n_slots = (uint32_t)grid * kWBlock * 2u;                                        \
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

// one CTA of kTcBlock threads per SM; shared memory = sphere operand | warp states | per-hit rows
static cudaError_t launch_fused_tc(WaveParams &p, void *pool, uint32_t pool_slots, bool f32hit, int sm_count,
                                   size_t smem_optin_max, bool *in_smem_out, int *grid_out, cudaStream_t st) {
  const size_t op_bytes = (size_t)p.scene.tc_tiles * 4u * (size_t)kTcSlabFloats * 4u;
  const size_t warp_smem = (size_t)kTcWarps * sizeof(WarpState);
  const size_t scene_smem = (size_t)p.scene.n_pad * 56;
  const bool in_smem = op_bytes + warp_smem + scene_smem <= smem_optin_max;
  const size_t smem = op_bytes + warp_smem + (in_smem ? scene_smem : 0);
  if (smem > smem_optin_max) return cudaErrorInvalidValue;
  const int grid = sm_count;
  const uint32_t n_slots = (uint32_t)grid * kTcBlock * 2u;
  if (n_slots > pool_slots) return cudaErrorInvalidValue;
  wave_carve(p, pool, n_slots);
  *grid_out = grid; *in_smem_out = in_smem;
#define RTW_TC(SM_, F32_)                                                                          \
  do {                                                                                             \
    auto k = fused_tc_kernel<SM_, F32_>;                                                           \
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e != cudaSuccess) return e;                                                                \
    k<<<grid, kTcBlock, smem, st>>>(p);                                                            \
  } while (0)
  if (in_smem && f32hit) RTW_TC(true, true);
  else if (in_smem) RTW_TC(true, false);
  else if (f32hit) RTW_TC(false, true);
  else RTW_TC(false, false);
#undef RTW_TC
  return cudaGetLastError();
}

// streamed form: two tile buffers per warpgroup | warp states (the per-hit rows stay in L1/L2)
static cudaError_t launch_fused_tcs(WaveParams &p, void *pool, uint32_t pool_slots, bool f32hit, int sm_count,
                                    size_t smem_optin_max, int *grid_out, cudaStream_t st) {
  const size_t smem = (size_t)kTcGroups * kTcsBufs * kTcsTileBytes + (size_t)kTcWarps * sizeof(WarpState);
  if (smem > smem_optin_max) return cudaErrorInvalidValue;
  const int grid = sm_count;
  const uint32_t n_slots = (uint32_t)grid * kTcBlock * 2u;
  if (n_slots > pool_slots) return cudaErrorInvalidValue;
  wave_carve(p, pool, n_slots);
  *grid_out = grid;
  cudaError_t e;
  if (f32hit) {
    auto k = fused_tcs_kernel<true>;
    if ((e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
    k<<<grid, kTcBlock, smem, st>>>(p);
  } else {
    auto k = fused_tcs_kernel<false>;
    if ((e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
    k<<<grid, kTcBlock, smem, st>>>(p);
  }
  return cudaGetLastError();
}

cudaError_t run_fused(WaveParams p, void *pool, uint32_t pool_slots, const float4 *host_rows,
                      bool use_bvh, bool f32_hit, bool use_tc, int sm_count, size_t smem_optin_max,
                      cudaStream_t st, WaveRunInfo *info) {
  if (use_bvh && !p.scene.bvh_nodes) return cudaErrorInvalidValue;
  if (use_bvh && f32_hit) return cudaErrorInvalidValue;
  if (use_tc) {
    if (use_bvh || !p.scene.tc_op || p.scene.tc_tiles == 0) return cudaErrorInvalidValue;
    bool in_smem = false;
    int grid = 0;
    cudaError_t e = p.scene.n > kTcMaxSpheres
                        ? launch_fused_tcs(p, pool, pool_slots, f32_hit, sm_count, smem_optin_max, &grid, st)
                        : launch_fused_tc(p, pool, pool_slots, f32_hit, sm_count, smem_optin_max, &in_smem, &grid, st);
    if (info) {
      info->launches = 1; info->n_slots = p.n_slots; info->rays_per_lane = 2; info->grid = grid;
      info->rows_in_constant_bank = false; info->scene_in_smem = in_smem; info->tensor_core_filter = true;
    }
    return e;
  }
  static thread_local ConstRows cr;  // 16 KB kernel parameter; frames may render from several host threads
  const bool crows = !use_bvh && host_rows && p.scene.n_pad + 2 * kRowTail <= (uint32_t)kConstRows &&
                     !env_cfg().no_const_rows;
  if (crows) memcpy(cr.row, host_rows, (size_t)(p.scene.n_pad + kRowTail) * sizeof(float4));
  // shared memory of a CTA: the warps' states, then (when every resident CTA can afford it) the
  // per-hit scene rows; an SM offers its opt-in maximum plus the 1 KB each CTA reserves
  const size_t warp_smem = (size_t)kWarps * sizeof(WarpState);
  const size_t scene_smem = (size_t)p.scene.n_pad * 56;
  const size_t ctas = use_bvh ? kBvhMinB : RTW_MINB;
  const bool in_smem = (warp_smem + scene_smem + 1024) * ctas <= smem_optin_max + 1024;
  const size_t smem = warp_smem + (in_smem ? scene_smem : 0);
  int grid = 0;
  cudaError_t e = launch_fused_t<RTW_G>(p, pool, pool_slots, cr, crows, in_smem, use_bvh, f32_hit, smem,
                                        sm_count, &grid, st);
  if (info) {
    info->launches = 1; info->n_slots = p.n_slots; info->rays_per_lane = 2; info->grid = grid;
    info->rows_in_constant_bank = crows; info->scene_in_smem = in_smem; info->tensor_core_filter = false;
  }
  return e;
}

}  // namespace rtb
