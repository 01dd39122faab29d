// csrc/rt_tc.cu — the line-sphere filter as ONE contraction on the 5th-generation tensor cores
// (tcgen05, sm_100a): probe kernel + host-side operand builder.
//
// The brute-force filter of rt_wave.cu decides, per (ray, sphere), the sign of
//     e = (c.d - o.d)^2 + 2 c.o + (r^2 - |c|^2) - |o|^2            (>= 0: the line meets the sphere)
// Expanded, e is a dot product of 11 SPHERE features with 11 RAY features,
//     phi(c, r) = [cx^2, cy^2, cz^2, cx cy, cx cz, cy cz, cx, cy, cz, w, 1]
//     psi(o, d) = [dx^2, dy^2, dz^2, 2 dx dy, 2 dx dz, 2 dy dz, q_x, q_y, q_z, 1, k]
//     q = 2 (o - (o.d) d),   k = (o.d)^2 - |o|^2,   w = r^2 - |c|^2,
// i.e. a [rays x K] x [K x spheres] matrix product: hittable_list::hit's "every ray against every
// sphere" (src/cpu/hittable_list.h:28-43) is a dense contraction once it is written this way.
// TF32 keeps 11 significant bits, so every feature is split x = hi + lo and the three products
// hi*hi + lo*hi + hi*lo take three K slots (the 3xTF32 scheme): 9 x 3 + 2 + 2 = 31 -> K = 32, four
// tcgen05.mma.kind::tf32 steps of K = 8 with fp32 accumulation in tensor memory.
//
//   rays    = the M operand (128 rows = the 128 TMEM lanes = 128 threads: thread <-> ray), written
//             by the threads themselves with tcgen05.st (A operand read from TMEM);
//   spheres = the N operand, K-major, no swizzle, resident in shared memory for the whole kernel;
//   D       = [128 rays x 64 spheres] fp32 in TMEM; each thread reads ITS ray's row with
//             tcgen05.ld and folds the 64 sign bits into a mask: no FMA left in the epilogue.
//
// Conservativeness: the split, the fp32 feature arithmetic and the accumulation perturb e by at
// most kappa * (A_s + B_r), A_s = 2|c|^2 + |w|, B_r = 2|o|^2 (bound of sum |phi_i psi_i|); kappa A_s
// is folded into w on the host, kappa B_r into k on the device, like the SIMT filter's slack.
#include <cmath>
#include <cstring>
#include <vector>

#include "rt_capi_internal.h"
#include "rt_tc.cuh"

namespace rtb {

// ---------------------------------------------------------------- host: sphere operand
static float tf32_trunc(float x) {
  uint32_t u;
  memcpy(&u, &x, 4);
  u &= 0xffffe000u;  // 1 + 8 + 10 bits
  memcpy(&x, &u, 4);
  return x;
}
static void split_tf32(double x, float &hi, float &lo) {
  hi = tf32_trunc((float)x);
  lo = tf32_trunc((float)(x - (double)hi));
}

// Canonical K-major, no-swizzle layout of one [kTcTile spheres x 8 K] slab (32 B per sphere): core matrices of
// 8 rows x 16 B; the two K halves of a row group 128 B apart (LBO), row groups 256 B apart (SBO).
static size_t slab_offset(int row, int k) {
  return (size_t)(row / 8) * 256 + (size_t)(k / 4) * 128 + (size_t)(row % 8) * 16 + (size_t)(k % 4) * 4;
}

// spheres {cx, cy, cz, signed r} -> [tiles][4 K steps][2048 B]; padding rows can never pass (w = -big)
void tc_build_sphere_operand(const double4 *exact, uint32_t n, std::vector<float> &out, uint32_t &tiles) {
  tiles = (n + kTcTile - 1) / kTcTile;
  if (tiles == 0) tiles = 1;
  out.assign((size_t)tiles * 4 * kTcSlabFloats, 0.0f);
  for (uint32_t s = 0; s < tiles * kTcTile; ++s) {
    double f[11];
    if (s < n) {
      const double cx = exact[s].x, cy = exact[s].y, cz = exact[s].z, r = exact[s].w;
      const double c2 = cx * cx + cy * cy + cz * cz, w = r * r - c2;
      f[0] = cx * cx; f[1] = cy * cy; f[2] = cz * cz; f[3] = cx * cy; f[4] = cx * cz; f[5] = cy * cz;
      f[6] = cx; f[7] = cy; f[8] = cz;
      f[9] = w + (double)kTcKappa * (2.0 * c2 + std::fabs(w));  // slack: can only ADD candidates
      f[10] = 1.0;
    } else {
      for (int k = 0; k < 11; ++k) f[k] = 0.0;
      f[9] = -1.0e30;
    }
    float slot[kTcK];
    for (int k = 0; k < kTcK; ++k) slot[k] = 0.0f;
    for (int i = 0; i < 9; ++i) {  // psi side holds [hi, lo, hi]
      float hi, lo;
      split_tf32(f[i], hi, lo);
      slot[3 * i] = hi; slot[3 * i + 1] = hi; slot[3 * i + 2] = lo;
    }
    slot[27] = 1.0f; slot[28] = 1.0f;  // x the ray's constant k (hi, lo)
    {
      // the truncating split loses at most 2^-20 |w|: add it first, so that the slack only grows
      float hi, lo;
      split_tf32(f[9] + std::fabs(f[9]) * 0x1p-19, hi, lo);
      slot[29] = hi; slot[30] = lo;    // x 1, 1
    }
    const uint32_t tile = s / kTcTile, row = s % kTcTile;
    for (int k = 0; k < kTcK; ++k) {
      const size_t off = ((size_t)tile * 4 + k / 8) * (kTcSlabFloats * 4) + slab_offset((int)row, k % 8);
      out[off / 4] = slot[k];
    }
  }
}

// ---------------------------------------------------------------- probe kernel
// One warpgroup (128 threads = 128 rays per block of rays) per CTA, one CTA per SM.  For every
// block of 128 rays: features -> TMEM, then per tile of 64 spheres 4 MMA steps, read back, fold
// the sign bits into masks[ray][tile] (bit b set: sphere tile * kTcTile + b passes).  `values`
// (optional) receives the raw e of the first `n_values` rays for the error measurement.
__global__ void __launch_bounds__(128, 1)
tc_filter_probe_kernel(const float4 *__restrict__ ray_o, const float4 *__restrict__ ray_d, uint32_t n_rays,
                       const float *__restrict__ sphere_op, uint32_t tiles, unsigned long long *__restrict__ masks,
                       float *__restrict__ values, uint32_t n_values, uint32_t repeats, unsigned int *__restrict__ status) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ uint32_t s_tmem;
  float *s_op = reinterpret_cast<float *>(smem_raw);
  const uint32_t tid = threadIdx.x, warp = tid >> 5;
  for (uint32_t i = tid; i < tiles * 4u * (uint32_t)kTcSlabFloats; i += 128u) s_op[i] = sphere_op[i];
  if (tid == 0) mbar_init(&s_bar, 1);
  if (warp == 0) tmem_alloc(&s_tmem, 128);
  // generic-proxy writes of the operand must be visible to the tensor core's async proxy
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const uint32_t lane_base = (warp * 32u) << 16;
  const uint32_t a_tmem = tmem, d_tmem = tmem + 32u;
  uint32_t parity = 0;
  bool ok = true;
  const uint32_t n_blocks = (n_rays + 127u) / 128u;
  for (uint32_t rep = 0; rep < repeats; ++rep)
    for (uint32_t blk = blockIdx.x; blk < n_blocks && ok; blk += gridDim.x) {
      const uint32_t ray = blk * 128u + tid;
      float4 o = make_float4(0.f, 0.f, 0.f, -1.f), d = make_float4(0.f, 1.f, 0.f, 0.f);
      if (ray < n_rays) { o = ray_o[ray]; d = ray_d[ray]; }
      uint32_t slot[kTcK];
      tc_ray_slots(o, d, slot);
      TC_ST32(a_tmem + lane_base, slot);
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncthreads();
      for (uint32_t t = 0; t < tiles && ok; ++t) {
        if (tid == 0) {
          tc_fence_after();
#pragma unroll
          for (uint32_t j = 0; j < 4; ++j)
            tc_mma_ts(d_tmem, a_tmem + j * 8u, tc_smem_desc(s_op + ((size_t)t * 4 + j) * kTcSlabFloats), kTcIdesc, j);
          tc_commit(&s_bar);
        }
        ok = mbar_wait_bounded(&s_bar, parity);
        parity ^= 1u;
        tc_fence_after();
        uint32_t v[kTcTile];
        TC_LD_TILE(d_tmem + lane_base, v);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncthreads();  // D is free: the next tile's MMAs may overwrite it while the masks are built
        uint32_t p0, p1;
        tc_pass_bits(v, p0, p1);
        const unsigned long long pass = (unsigned long long)p1 << 32 | p0;
        if (ray < n_rays && rep == 0) {
          masks[(size_t)ray * tiles + t] = pass;
          if (values && ray < n_values)
            for (int b = 0; b < kTcTile; ++b) values[((size_t)ray * tiles + t) * kTcTile + b] = __uint_as_float(v[b]);
        }
      }
    }
  if (!ok && tid == 0) atomicExch(status, 1u);
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

}  // namespace rtb

using namespace rtb;

// Measurement / validation aid: the tensor-core form of the filter on `n` rays
// {o.xyz, |d| > 0} / {unit d} against the scene.  masks: HOST [n][tiles] 64-bit pass masks;
// values: optional HOST [n_values][tiles][64] raw e; ms: device time of `repeats` passes over the rays.
extern "C" int rt_tc_filter_probe(const rt_scene *sc, const float *ray_o, const float *ray_d, uint32_t n,
                                  uint64_t *masks, uint32_t *tiles_out, uint32_t *tile_spheres, float *values,
                                  uint32_t n_values, uint32_t repeats, float *ms) {
  if (!sc || !ray_o || !ray_d || !masks || !tiles_out || !tile_spheres) return rt_fail(RT_E_INVALID, "NULL argument");
  *tile_spheres = (uint32_t)kTcTile;
  if (sc->n == 0 || sc->n > 1024u) return rt_fail(RT_E_INVALID, "the probe keeps the sphere operand in shared memory: 1..1024 spheres");
  CU(cudaSetDevice(sc->device));
  const double4 *e = reinterpret_cast<const double4 *>(sc->stage);  // exact rows sit first in the staging buffer
  std::vector<float> op;
  uint32_t tiles = 0;
  tc_build_sphere_operand(e, sc->n, op, tiles);
  *tiles_out = tiles;
  float4 *d_o = nullptr, *d_d = nullptr;
  float *d_op = nullptr, *d_val = nullptr;
  unsigned long long *d_m = nullptr;
  unsigned int *d_st = nullptr, h_st = 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  const size_t op_bytes = op.size() * sizeof(float);
  cudaError_t ce = cudaMalloc(&d_o, (size_t)n * 16);
  if (ce == cudaSuccess) ce = cudaMalloc(&d_d, (size_t)n * 16);
  if (ce == cudaSuccess) ce = cudaMalloc(&d_op, op_bytes);
  if (ce == cudaSuccess) ce = cudaMalloc(&d_m, (size_t)n * tiles * 8);
  if (ce == cudaSuccess) ce = cudaMalloc(&d_st, 4);
  if (ce == cudaSuccess && values && n_values) ce = cudaMalloc(&d_val, (size_t)n_values * tiles * kTcTile * 4);
  if (ce == cudaSuccess) ce = cudaMemcpy(d_o, ray_o, (size_t)n * 16, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemcpy(d_d, ray_d, (size_t)n * 16, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemcpy(d_op, op.data(), op_bytes, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemset(d_st, 0, 4);
  if (ce == cudaSuccess) ce = cudaEventCreate(&e0);
  if (ce == cudaSuccess) ce = cudaEventCreate(&e1);
  int sms = 0;
  if (ce == cudaSuccess) ce = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, sc->device);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(tc_filter_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)op_bytes);
  if (ce == cudaSuccess) {
    const uint32_t blocks = (n + 127u) / 128u;
    const int grid = (int)(blocks < (uint32_t)sms ? blocks : (uint32_t)sms);
    cudaEventRecord(e0, 0);
    tc_filter_probe_kernel<<<grid, 128, op_bytes>>>(d_o, d_d, n, d_op, tiles, d_m, d_val, d_val ? n_values : 0u,
                                                    repeats ? repeats : 1u, d_st);
    cudaEventRecord(e1, 0);
    ce = cudaGetLastError();
  }
  if (ce == cudaSuccess) ce = cudaEventSynchronize(e1);
  if (ce == cudaSuccess && ms) ce = cudaEventElapsedTime(ms, e0, e1);
  if (ce == cudaSuccess) ce = cudaMemcpy(masks, d_m, (size_t)n * tiles * 8, cudaMemcpyDeviceToHost);
  if (ce == cudaSuccess && d_val) ce = cudaMemcpy(values, d_val, (size_t)n_values * tiles * kTcTile * 4, cudaMemcpyDeviceToHost);
  if (ce == cudaSuccess) ce = cudaMemcpy(&h_st, d_st, 4, cudaMemcpyDeviceToHost);
  cudaFree(d_o); cudaFree(d_d); cudaFree(d_op); cudaFree(d_m); cudaFree(d_st); cudaFree(d_val);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  if (ce != cudaSuccess) return rt_fail(RT_E_CUDA, "tensor-core filter probe failed: %s", cudaGetErrorString(ce));
  if (h_st) return rt_fail(RT_E_CUDA, "tensor-core filter probe: an MMA never completed (timeout)");
  return RT_OK;
}
