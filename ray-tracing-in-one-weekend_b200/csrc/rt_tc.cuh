// csrc/rt_tc.cuh — tcgen05 building blocks of the tensor-core form of the brute-force filter
// (see rt_tc.cu for the formulation): TMEM allocation, the tf32 MMA with its A operand in tensor
// memory, descriptors, the ray-side feature split.  Shared by the probe kernel (rt_tc.cu) and the
// path kernel (rt_wave.cu).
#pragma once

#include "rt_device.cuh"

namespace rtb {

constexpr int kTcK = 32;          // K slots (4 MMA steps of 8)
#ifndef RTW_TC_TILE
#define RTW_TC_TILE 64
#endif
constexpr int kTcTile = RTW_TC_TILE;  // spheres per MMA (N): 48 or 64 (a warpgroup's TMEM share holds 32 + N columns)
static_assert(kTcTile == 48 || kTcTile == 64, "accumulator tile of 48 or 64 spheres");
constexpr int kTcSlabFloats = kTcTile * 8;   // one [tile x 8 K] slab of the sphere operand
constexpr int kTcSlabUnits = kTcTile * 2;    // ... in 16-byte units of the descriptor's address field
constexpr float kTcKappa = 1.52587890625e-05f;  // 2^-16: bound of the contraction's error in units of A_s + B_r

// ---------------------------------------------------------------- device helpers (tcgen05)
__device__ __forceinline__ float d_tf32_trunc(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

__device__ __forceinline__ void tmem_alloc(uint32_t *slot_in_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(slot_in_smem)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[tmem] x B[smem descriptor], kind::tf32, one CTA
__device__ __forceinline__ void tc_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(bar)) : "memory");
}
// shared-memory matrix descriptor: K-major, no swizzle, LBO = 128 B, SBO = 256 B, version 1 (Blackwell)
__device__ __forceinline__ uint64_t tc_smem_desc(const void *p) {
  const uint64_t a = (uint64_t)((smem_addr(p) >> 4) & 0x3fffu);
  return a | (8ull << 16) | (16ull << 32) | (1ull << 46);
}
// instruction descriptor: D = f32, A = B = tf32, both K-major, M = 128, N = 64
constexpr uint32_t kTcIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kTcTile >> 3) << 17) | ((128u >> 4) << 24);

// bounded wait: a wrong descriptor must not hang the GPU box.  try_wait suspends the thread in
// hardware for a system-dependent time, so the loop spins a handful of times; 2^22 expiries
// (seconds) mean the MMA will never complete.
__device__ __forceinline__ bool mbar_wait_bounded(uint64_t *bar, uint32_t parity) {
  for (uint32_t spins = 0; spins < (1u << 22); ++spins) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(ok)
        : "r"(smem_addr(bar)), "r"(parity)
        : "memory");
    if (ok) return true;
  }
  return false;
}

// sign bits of 32 accumulator values -> one word (bit b = sign of v[o + b]).  Eight independent
// chains of funnel shifts (one per byte) instead of one chain of 32: with 16 warps per SM the
// latency of a dependent chain is not hidden.
template <int BYTES = 4>
__device__ __forceinline__ uint32_t tc_sign_mask(const uint32_t (&v)[kTcTile], const int o) {
  uint32_t m[BYTES];
#pragma unroll
  for (int k = 0; k < BYTES; ++k) m[k] = 0u;
#pragma unroll
  for (int b = 7; b >= 0; --b)
#pragma unroll
    for (int k = 0; k < BYTES; ++k) m[k] = __funnelshift_l(v[o + 8 * k + b], m[k], 1);
  uint32_t r = m[0];
#pragma unroll
  for (int k = 1; k < BYTES; ++k) r |= m[k] << (8 * k);
  return r;
}
// pass bits of a tile: low word = spheres 0..31, high word = spheres 32..kTcTile-1 (sign bit 1 = miss)
__device__ __forceinline__ void tc_pass_bits(const uint32_t (&v)[kTcTile], uint32_t &p0, uint32_t &p1) {
  p0 = ~tc_sign_mask<4>(v, 0);
  p1 = kTcTile == 64 ? ~tc_sign_mask<4>(v, 32) : (~tc_sign_mask<2>(v, 32) & 0xffffu);
}

__device__ __forceinline__ uint32_t tc_pass_bits_lo(const uint32_t (&v)[kTcTile]) { return ~tc_sign_mask<4>(v, 0); }
__device__ __forceinline__ uint32_t tc_pass_bits_hi(const uint32_t (&v)[kTcTile]) {
  return kTcTile == 64 ? ~tc_sign_mask<4>(v, 32) : (~tc_sign_mask<2>(v, 32) & 0xffffu);
}

// psi(o, d) -> the ray's 32 K slots {hi, lo, hi} x 9, k (hi, lo), 1, 1, 0
__device__ __forceinline__ void tc_ray_slots(const float4 o, const float4 d, uint32_t (&slot)[kTcK]) {
  const float od = o.x * d.x + o.y * d.y + o.z * d.z;
  const float o2 = o.x * o.x + o.y * o.y + o.z * o.z;
  float f[9] = {d.x * d.x, d.y * d.y, d.z * d.z, 2.0f * d.x * d.y, 2.0f * d.x * d.z, 2.0f * d.y * d.z,
                2.0f * (o.x - od * d.x), 2.0f * (o.y - od * d.y), 2.0f * (o.z - od * d.z)};
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    const float hi = d_tf32_trunc(f[i]), lo = d_tf32_trunc(f[i] - hi);
    slot[3 * i] = __float_as_uint(hi); slot[3 * i + 1] = __float_as_uint(lo); slot[3 * i + 2] = __float_as_uint(hi);
  }
  // k = (o.d)^2 - |o|^2 + kappa * 2 |o|^2, plus a few ulps of |o|^2 for the f32 arithmetic above and
  // 2^-19 |k| for the truncating split below: the slack can only ADD candidates.  A slot without a
  // ray (o.w <= 0) gets k = -huge: it passes nothing.
  float k = fmaf(od, od, -o2) + (2.0f * kTcKappa + 4.0e-7f) * o2;
  k += fabsf(k) * 0x1p-19f;
  if (!(o.w > 0.0f)) k = -1.0e30f;
  const float kh = d_tf32_trunc(k), kl = d_tf32_trunc(k - kh);
  slot[27] = __float_as_uint(kh); slot[28] = __float_as_uint(kl);
  slot[29] = __float_as_uint(1.0f); slot[30] = __float_as_uint(1.0f);
  slot[31] = 0u;
}

#define TC_ST32(addr, s)                                                                                   \
  asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, "   \
               "%11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, " \
               "%29, %30, %31, %32};" ::"r"(addr),                                                         \
               "r"(s[0]), "r"(s[1]), "r"(s[2]), "r"(s[3]), "r"(s[4]), "r"(s[5]), "r"(s[6]), "r"(s[7]),      \
               "r"(s[8]), "r"(s[9]), "r"(s[10]), "r"(s[11]), "r"(s[12]), "r"(s[13]), "r"(s[14]),            \
               "r"(s[15]), "r"(s[16]), "r"(s[17]), "r"(s[18]), "r"(s[19]), "r"(s[20]), "r"(s[21]),          \
               "r"(s[22]), "r"(s[23]), "r"(s[24]), "r"(s[25]), "r"(s[26]), "r"(s[27]), "r"(s[28]),          \
               "r"(s[29]), "r"(s[30]), "r"(s[31])                                                          \
               : "memory")

#define TC_LD32(addr, v, o)                                                                                 \
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, " \
               "%12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, "  \
               "%30, %31}, [%32];"                                                                          \
               : "=r"(v[o + 0]), "=r"(v[o + 1]), "=r"(v[o + 2]), "=r"(v[o + 3]), "=r"(v[o + 4]),            \
                 "=r"(v[o + 5]), "=r"(v[o + 6]), "=r"(v[o + 7]), "=r"(v[o + 8]), "=r"(v[o + 9]),            \
                 "=r"(v[o + 10]), "=r"(v[o + 11]), "=r"(v[o + 12]), "=r"(v[o + 13]), "=r"(v[o + 14]),       \
                 "=r"(v[o + 15]), "=r"(v[o + 16]), "=r"(v[o + 17]), "=r"(v[o + 18]), "=r"(v[o + 19]),       \
                 "=r"(v[o + 20]), "=r"(v[o + 21]), "=r"(v[o + 22]), "=r"(v[o + 23]), "=r"(v[o + 24]),       \
                 "=r"(v[o + 25]), "=r"(v[o + 26]), "=r"(v[o + 27]), "=r"(v[o + 28]), "=r"(v[o + 29]),       \
                 "=r"(v[o + 30]), "=r"(v[o + 31])                                                           \
               : "r"(addr)                                                                                  \
               : "memory")


#define TC_LD16(addr, v, o)                                                                                 \
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, " \
               "%12, %13, %14, %15}, [%16];"                                                                \
               : "=r"(v[o + 0]), "=r"(v[o + 1]), "=r"(v[o + 2]), "=r"(v[o + 3]), "=r"(v[o + 4]),            \
                 "=r"(v[o + 5]), "=r"(v[o + 6]), "=r"(v[o + 7]), "=r"(v[o + 8]), "=r"(v[o + 9]),            \
                 "=r"(v[o + 10]), "=r"(v[o + 11]), "=r"(v[o + 12]), "=r"(v[o + 13]), "=r"(v[o + 14]),       \
                 "=r"(v[o + 15])                                                                            \
               : "r"(addr)                                                                                  \
               : "memory")

// columns 32 .. kTcTile-1 of this thread's row
#define TC_LD_TILE_HI(addr, v)                                   \
  do {                                                           \
    if (kTcTile == 64) { TC_LD32((addr) + 32u, v, 32); }         \
    else { TC_LD16((addr) + 32u, v, 32); }                       \
  } while (0)

// this thread's row of an accumulator tile -> registers
#define TC_LD_TILE(addr, v)                                      \
  do {                                                           \
    TC_LD32((addr), v, 0);                                       \
    if (kTcTile == 64) { TC_LD32((addr) + 32u, v, 32); }         \
    else { TC_LD16((addr) + 32u, v, 32); }                       \
  } while (0)

}  // namespace rtb
