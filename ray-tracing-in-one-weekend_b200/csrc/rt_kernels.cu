// csrc/rt_kernels.cu — hand-written sm_100a kernels of the path-tracing hot path.
//
//   (the persistent path-tracing kernel lives in rt_wave.cu)
//   reduce_local      canonical pairwise tree over the slices one GPU rendered.
//   finalize_kernel   upper tree levels over per-GPU partials + write_color (color.h:8-23).
//   primary_f32       production-precision primary-hit table (accuracy gate, f32 leg).
//
//   cost order        hand-out order of the next frame's work queue (pixels by ray count).
//   ffma_peak         FP32 FFMA throughput micro-kernel (roofline denominator of bench.py).
//
// All of these stream once over per-pixel buffers (HBM-bound, 16 B per pixel and input plane);
// together they are < 0.1 % of a frame.  Design notes: DESIGN.md section 4.2.
#include "rt_device.cuh"
#include "rt_kernels.h"

#include <cub/device/device_radix_sort.cuh>

namespace rtb {

// ---------------------------------------------------------------- slice tree + finalise
// Canonical pairwise tree over the RT_SLICES = 64 slices: ((s0+s1)+(s2+s3))+((s4+s5)+(s6+s7)) ...
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

// the same tree over N consecutive slice planes of one accumulation buffer
template <int N>
__device__ __forceinline__ float4 tree_sum_planes(const float4 *base, size_t plane, size_t i) {
  if constexpr (N == 1) {
    return __ldcs(base + i);
  } else {
    return add4(tree_sum_planes<N / 2>(base, plane, i),
                tree_sum_planes<N / 2>(base + (N / 2) * plane, plane, i));
  }
}

template <int N>
__global__ void __launch_bounds__(256) reduce_local_kernel(const float4 *__restrict__ first_plane,
                                                           float4 *__restrict__ dst,
                                                           uint32_t npix) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < npix;
       i += (size_t)gridDim.x * blockDim.x)
    dst[i] = tree_sum_planes<N>(first_plane, npix, i);
}

// write_color, src/cpu/color.h:8-23: x = sqrt(sum/spp); int(256 * clamp(x, 0, 0.999)).  gamma =
// false is the write_color of the book's first diffuse steps (no sqrt yet: archive/listing30/color.h,
// archive-gpu/image7...11/color.h), RT_FLAG_LINEAR_OUTPUT.
__device__ __forceinline__ uint32_t quantise(float sum, float scale, bool gamma) {
  float x = gamma ? sqrtf(scale * sum) : scale * sum;
  x = fminf(fmaxf(x, 0.0f), 0.999f);
  return (uint32_t)(int)(256.0f * x);
}

// One thread = 4 consecutive pixels = 12 output bytes = three 32-bit stores (coalesced: a
// warp writes 384 contiguous bytes).  Ray counts (w channel) are block-reduced to one atomic.
template <int N>
__global__ void __launch_bounds__(256) finalize_kernel(const __grid_constant__ PtrPack src,
                                                       uint32_t npix, float scale, bool gamma,
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
      c[k * 3 + 0] = quantise(v.x, scale, gamma);
      c[k * 3 + 1] = quantise(v.y, scale, gamma);
      c[k * 3 + 2] = quantise(v.z, scale, gamma);
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

// ---------------------------------------------------------------- hand-out order by cost
// Work items of a slice are handed out heaviest pixels first (cost = ray segments per sample of
// the pixel in the frame rendered before on this frame object), so the items still in flight when
// the work queue runs dry are the cheap ones and the end-of-frame tail is short.  The order only
// schedules work: every item's sum is independent of it, and so is the image.
__global__ void __launch_bounds__(256) cost_keys_kernel(const float4 *__restrict__ partial, uint32_t npix,
                                                        float inv_samples, uint32_t heavy_quarter_rays,
                                                        bool heavy_only, uint32_t *__restrict__ keys,
                                                        uint32_t *__restrict__ iota) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
    // quarter-ray resolution, 8-bit key, ascending key = descending cost; the sort is stable, so
    // pixels of equal cost keep their scan order (coherent neighbours stay neighbours).
    // heavy_only: two keys -- the heavy pixels, then everything else in scan order
    const float q = __ldg(&partial[i]).w * inv_samples * 4.0f;
    const uint32_t c = q >= 255.0f ? 255u : (uint32_t)q;
    keys[i] = heavy_only ? (c >= heavy_quarter_rays ? 0u : 1u) : 255u - c;
    iota[i] = i;
  }
}

// number of leading entries of the sorted order whose cost is >= kHeavyRays rays per sample
// (binary search in the sorted keys; ascending key = descending cost)
__global__ void heavy_count_kernel(const uint32_t *__restrict__ sorted_keys, uint32_t npix,
                                   uint32_t max_key, uint32_t *__restrict__ heavy) {
  uint32_t lo = 0, hi = npix;  // first index with key > max_key
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (sorted_keys[mid] <= max_key) lo = mid + 1; else hi = mid;
  }
  *heavy = lo;
}

size_t cost_order_temp_bytes(uint32_t npix) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t *)nullptr, (uint32_t *)nullptr,
                                  (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)npix, 0, 8);
  return bytes;
}

cudaError_t launch_cost_order(const float4 *partial, uint32_t npix, float inv_samples, uint32_t *keys,
                              uint32_t *keys_out, uint32_t *iota, uint32_t *order, uint32_t *heavy,
                              bool heavy_only, void *temp, size_t temp_bytes, int sm_count,
                              cudaStream_t st) {
  int grid = sm_count * 8;
  // pixels whose paths are long on average (glass, facing mirrors) are where the rare 50-bounce
  // paths live: ALL their slices are handed out before anything else
  unsigned quarter_rays = 4u * kHeavyRays;
  if (env_cfg().heavy_quarter_rays >= 0) quarter_rays = (unsigned)env_cfg().heavy_quarter_rays;  // tuning knob
  if (quarter_rays > 255u) quarter_rays = 255u;
  cost_keys_kernel<<<grid, 256, 0, st>>>(partial, npix, inv_samples, quarter_rays, heavy_only, keys, iota);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys, keys_out, iota, order, (int)npix, 0, 8, st);
  if (e != cudaSuccess) return e;
  heavy_count_kernel<<<1, 1, 0, st>>>(keys_out, npix, heavy_only ? 0u : 255u - quarter_rays, heavy);
  return cudaGetLastError();
}

// ---------------------------------------------------------------- launchers
static int grid_for(uint32_t work_items, int sm_count, int per_sm = 8) {
  int need = (int)((work_items + 255) / 256);
  int cap = sm_count * per_sm;
  return need < cap ? (need > 0 ? need : 1) : cap;
}

cudaError_t launch_reduce_local(const float4 *accum, uint32_t npix, uint32_t slice_begin,
                                uint32_t slice_end, float4 *dst, int sm_count, cudaStream_t st) {
  uint32_t cnt = slice_end - slice_begin;
  const float4 *pk = accum + (size_t)slice_begin * npix;
  int grid = grid_for(npix, sm_count);
  switch (cnt) {
    case 1: reduce_local_kernel<1><<<grid, 256, 0, st>>>(pk, dst, npix); break;
    case 2: reduce_local_kernel<2><<<grid, 256, 0, st>>>(pk, dst, npix); break;
    case 4: reduce_local_kernel<4><<<grid, 256, 0, st>>>(pk, dst, npix); break;
    case 8: reduce_local_kernel<8><<<grid, 256, 0, st>>>(pk, dst, npix); break;
    case 16: reduce_local_kernel<16><<<grid, 256, 0, st>>>(pk, dst, npix); break;
    case 32: reduce_local_kernel<32><<<grid, 256, 0, st>>>(pk, dst, npix); break;
    case 64: reduce_local_kernel<64><<<grid, 256, 0, st>>>(pk, dst, npix); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_finalize(const float4 *const *partials, uint32_t count, uint32_t npix,
                            float scale, bool gamma, uint32_t *rgb_words, float *linear,
                            unsigned long long *rays, int sm_count, cudaStream_t st) {
  PtrPack pk{};
  for (uint32_t k = 0; k < count && k < 8; ++k) pk.p[k] = partials[k];
  int grid = grid_for((npix + 3) / 4, sm_count);
  switch (count) {
    case 1: finalize_kernel<1><<<grid, 256, 0, st>>>(pk, npix, scale, gamma, rgb_words, linear, rays); break;
    case 2: finalize_kernel<2><<<grid, 256, 0, st>>>(pk, npix, scale, gamma, rgb_words, linear, rays); break;
    case 4: finalize_kernel<4><<<grid, 256, 0, st>>>(pk, npix, scale, gamma, rgb_words, linear, rays); break;
    case 8: finalize_kernel<8><<<grid, 256, 0, st>>>(pk, npix, scale, gamma, rgb_words, linear, rays); break;
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
