// csrc/rt_device.cuh — device-side building blocks shared by the sm_100a kernels.
// Reference citations are relative to the reference root (src/cpu/ is the semantic spec).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace rtb {

// ---------------------------------------------------------------- flattened scene (HBM)
// hittable_list (src/cpu/hittable_list.h:12-26) flattened in insertion order.
struct SceneView {
  const float4 *fast;    // [n_pad + kRowTail] {cx, cy, cz, r^2-|c|^2 + slack}  conservative f32 filter row
  const double4 *exact;  // [n_pad] {cx, cy, cz, signed radius}       f64 candidate resolution
  const float4 *mat;     // [n] {albedo.r, albedo.g, albedo.b, fuzz | ir}
  const float2 *aux;     // [n] {1 / signed radius, material kind (integer bits)}
  const struct BvhNode *bvh_nodes;  // optional hierarchy (rt_bvh.cu); root = node 0
  const uint32_t *bvh_prim;         // sphere indices referenced by the leaves
  const float *tc_op;               // sphere operand of the tensor-core filter (rt_tc.cu): [tc_tiles][4][2048 B]
  uint32_t tc_tiles;                // 64-sphere tiles of it (0 = not built)
  uint32_t n;            // spheres in the hittable_list
  uint32_t n_pad;        // rows allocated: n rounded up to a multiple of kRowPad; padding rows can
                         // never pass the filter (w = -inf)
};

// One node of the optional BVH, 32 bytes.  count > 0: leaf over bvh_prim[first, first+count);
// count == 0: inner node with children first and first + 1.  Bounds are f32, rounded outwards.
struct BvhNode {
  float lo[3], hi[3];
  uint32_t first, count;
};
constexpr unsigned kBvhLeafMax = 4, kBvhMaxDepth = 64;

// camera (src/cpu/camera.h:8-26) derived on the host in f64, narrowed once.
struct CameraF32 {
  float origin[3];
  float llc_minus_origin[3];  // lower_left_corner - origin
  float horizontal[3], vertical[3];
  float u[3], v[3];
  float lens_radius;
  float inv_wm1, inv_hm1;  // 1/(W-1), 1/(H-1)  (src/cpu/main.cc:116-117)
};

struct CameraF64 {
  double origin[3], lower_left[3], horizontal[3], vertical[3];
};

// Relative slack of the f32 line-sphere filter (see DESIGN.md "filter error bound").
#define RTB_FILTER_SLACK 1.9073486328125e-06f /* 32 * 2^-24 */

// @phase fetch
// ---------------------------------------------------------------- canonical sample slices
// The spp samples of a pixel are cut into 64 canonical slices in 8 blocks of 8.  Block b covers
// [floor(b*spp/8), floor((b+1)*spp/8)); inside a block the slices take 13, 12, 10, 9, 8, 6, 3, 1
// sixty-secondths of it (exactly those sample counts at 500 spp, where a block has 62 samples).
// Work items are handed out largest slices first, so a frame ends on its smallest items; and the
// largest item stays below a fifth of a block — with one block per GPU (8 GPUs) an item of a
// pixel that needs ~25 ray segments per sample must still be much shorter than the whole kernel
// (a bounce costs one pass over the spheres, ~36 us: halving sizes 31, 16, 8... put single items
// of the glass-ball pixels on the critical path: 31 x 23 x 36 us = 26 ms in a 21 ms kernel).
__host__ __device__ __forceinline__ void slice_range(uint32_t k, uint32_t spp, uint32_t &begin,
                                                     uint32_t &end) {
  const uint32_t blk = k >> 3, j = k & 7u;
  const uint32_t lo = (uint32_t)(((uint64_t)blk * spp) >> 3);
  const uint32_t hi = (uint32_t)(((uint64_t)(blk + 1) * spp) >> 3);
  // cumulative weights 0, 13, 25, 35, 44, 52, 58, 61, 62 packed one byte each
  const uint64_t cum = 0x3e3d3a342c23190dull;  // bytes: 13, 25, 35, 44, 52, 58, 61, 62
  const uint32_t c0 = j ? (uint32_t)(cum >> (8u * (j - 1u))) & 0xffu : 0u;
  const uint32_t c1 = (uint32_t)(cum >> (8u * j)) & 0xffu;
  const uint32_t n = hi - lo;
  begin = lo + (uint32_t)(((uint64_t)n * c0) / 62u);
  end = lo + (uint32_t)(((uint64_t)n * c1) / 62u);
}
// order_idx-th slice of the aligned block [sb, se) in hand-out order (big slices first)
__host__ __device__ __forceinline__ uint32_t slice_in_order(uint32_t order_idx, uint32_t sb, uint32_t se) {
  const uint32_t cnt = se - sb;
  if ((sb & 7u) || (cnt & 7u)) return sb + order_idx;
  const uint32_t nb = cnt >> 3;
  return sb + (order_idx % nb) * 8u + order_idx / nb;
}

// @phase philox
// ---------------------------------------------------------------- Philox4x32-10
// Stream contract (identical in oracle/rt_oracle.c):
//   key = (pixel index in output order, absolute sample index)
//   ctr = (slot, 0, seed_lo, seed_hi); slot 0 = camera, slot k >= 1 = scatter at k-th hit
//   uniform = (word >> 8) * 2^-24
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2,
                                               uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    uint32_t h0 = __umulhi(M0, c0), l0 = M0 * c0;
    uint32_t h1 = __umulhi(M1, c2), l1 = M1 * c2;
    c0 = h1 ^ c1 ^ k0; c1 = l1;
    c2 = h0 ^ c3 ^ k1; c3 = l0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ float4 philox_uniforms(uint32_t pixel, uint32_t sample,
                                                  uint32_t slot, uint64_t seed) {
  uint4 o = philox4x32_10(slot, 0u, (uint32_t)seed, (uint32_t)(seed >> 32), pixel, sample);
  const float s = 5.9604644775390625e-08f;  // 2^-24
  return make_float4((float)(o.x >> 8) * s, (float)(o.y >> 8) * s, (float)(o.z >> 8) * s,
                     (float)(o.w >> 8) * s);
}

// @phase scatter
// ---------------------------------------------------------------- samplers
// Direct, distribution-equivalent forms of the reference's rejection loops.
__device__ __forceinline__ void unit_vector_from(float x0, float x1, float &ux, float &uy,
                                                 float &uz) {  // ~ src/cpu/vec3.h:114
  float z = 1.0f - 2.0f * x0;
  float rxy = sqrtf(fmaxf(0.0f, 1.0f - z * z));
  float s, c;
  sincospif(2.0f * x1, &s, &c);
  ux = rxy * c; uy = rxy * s; uz = z;
}

// @phase gate
// ---------------------------------------------------------------- exact candidate test
// sphere::hit (src/cpu/sphere.h:24-51) in double on the f32 ray; returns true and updates
// t_best when the root lies in [t_min, t_best] (inclusive upper bound: a later sphere with
// an equal root replaces the earlier one, hittable_list.h:34-39 + sphere.h:38-40).
__device__ __forceinline__ bool exact_hit(const double4 S, double ox, double oy, double oz,
                                          double dx, double dy, double dz, double t_min,
                                          double &t_best) {
  double ocx = ox - S.x, ocy = oy - S.y, ocz = oz - S.z;
  double a = dx * dx + dy * dy + dz * dz;
  double half_b = ocx * dx + ocy * dy + ocz * dz;
  double c = ocx * ocx + ocy * ocy + ocz * ocz - S.w * S.w;
  double disc = half_b * half_b - a * c;
  if (disc < 0.0) return false;
  double sq = sqrt(disc);
  double root = (-half_b - sq) / a;
  if (root < t_min || t_best < root) {
    root = (-half_b + sq) / a;
    if (root < t_min || t_best < root) return false;
  }
  t_best = root;
  return true;
}

// @phase stage
// ---------------------------------------------------------------- bulk copy (TMA 1-D)
__device__ __forceinline__ uint32_t smem_addr(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes,
                                         uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_addr(dst)), "l"(src), "r"(bytes), "r"(smem_addr(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_addr(bar)), "r"(parity)
      : "memory");
}

}  // namespace rtb
