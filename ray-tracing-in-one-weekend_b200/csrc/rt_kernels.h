// csrc/rt_kernels.h — host-visible launch interface of the kernels (internal to the .so).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_device.cuh"

#include <vector>

#define RT_SLICES_DEV 64u  // == RT_SLICES (include/rt_b200.h)
// work item = slice << kItemShift | pixel (0xffffffff = none)
constexpr unsigned kItemShift = 26;
constexpr unsigned kItemPixelMask = (1u << kItemShift) - 1u;

// Scene rows are padded to a multiple of kRowPad (padding rows never pass the filter) and the
// filter-row array carries kRowTail more padding rows for the loop's prefetch.
#ifndef RTW_G
#define RTW_G 4  // sphere rows per decision group of the filter loop (two groups per trip)
#endif
constexpr unsigned kRowPad = 2 * RTW_G, kRowTail = 2 * RTW_G;

namespace rtb {

// Filter rows passed by value as a kernel parameter (constant bank 0): scenes of up to
// kConstRows - 4 padded rows.  16 KB of the 32 KB parameter space.
constexpr int kConstRows = 1024;
struct ConstRows {
  float4 row[kConstRows];
};

// ---------------------------------------------------------------- path kernel (rt_wave.cu)
// A warp owns 64 path slots for the whole frame (2 per lane).  Rays and the closest-hit records of
// a wave live in the warp's shared memory; the rest of a slot's state (throughput, item sums, item)
// lives in a pool in device memory (L2-resident).
#ifndef RTW_LIST_SMEM
#define RTW_LIST_SMEM 176  // with 176 the cover scene's CTA (operand 64 KB + 20 warp states + per-hit rows) stays under the 164 KB shared-memory step: 32 KB more L1
#endif
constexpr unsigned kListSmem = RTW_LIST_SMEM;   // candidate entries of a warp and wave kept in shared memory
constexpr unsigned kListGlobalPerSlot = 64;     // ... and this many more per slot in the pool (x 64 slots per warp)
constexpr unsigned kGroupBits = 26;             // entry = local slot << kGroupBits | decision group (row / G)

struct WaveParams {
  SceneView scene;
  CameraF32 cam;
  uint32_t width, height;
  float inv_width;              // 1 / width (pixel -> row estimate, fixed up exactly)
  uint32_t spp, max_depth;
  uint64_t seed;
  uint32_t sample_offset;
  uint32_t n_orders;            // slices handed out (hand-out order, trailing empty slices trimmed)
  uint32_t model_v4;            // 1 = src/gpu semantics (RT_FLAG_V4_MODEL): strict root interval, fuzz on the sphere
  uint32_t no_filter;           // 1 = RT_FLAG_NO_FILTER: no f32 pre-test, every ray resolves every sphere in f64
  float4 *accum;                // [RT_SLICES][H*W] {sum r, sum g, sum b, ray segments}
  unsigned long long *counter;  // work-queue head (zeroed before the frame)
  unsigned int *status;         // set to 1 if a tensor-core MMA never completed (bounded wait)
  const uint32_t *order;        // pixel hand-out order within a slice (nullptr = scan order)
  const uint32_t *heavy;        // leading pixels of `order` whose slices are all handed out first
  // pool of path slots (SoA, n_slots each)
  uint32_t n_slots;
  float4 *thr;         // {throughput r, g, b, depth (int bits): >0 bounces left, 0 new path, -1 retired}
  float4 *acc;         // {item sum r, g, b, ray segments}
  uint4 *it;           // {slice << kItemShift | pixel, next sample, slice end, -}
  uint32_t *glist;     // [n_slots * kListGlobalPerSlot] candidate entries beyond a warp's shared-memory list
  // per-lane indexed tables (constant bank): slice handed out at position k of the hand-out order,
  // and the sample range [bounds[s], bounds[s + 1]) of canonical slice s
  uint8_t order_slice[64];
  uint32_t slice_bounds[65];
};

// RT_B200_* experiment switches, read ONCE per process.
struct EnvCfg {
  bool no_tc;               // RT_B200_NO_TC: SIMT filter loop even where the tensor-core filter applies
  bool no_tc_stream;        // RT_B200_NO_TC_STREAM: ... for scenes beyond kTcMaxSpheres only
  bool no_const_rows;       // RT_B200_NO_CONST_ROWS
  bool no_cost_order;       // RT_B200_NO_COST_ORDER
  int heavy_quarter_rays;   // RT_B200_HEAVY_QUARTER_RAYS, -1 unset
};
const EnvCfg &env_cfg();

struct WaveRunInfo {
  uint32_t launches, n_slots;
  int rays_per_lane, grid;
  bool rows_in_constant_bank, scene_in_smem, tensor_core_filter;
};

size_t wave_pool_bytes(uint32_t n_slots);
uint32_t wave_default_slots(int sm_count);

// Host-side BVH build over {cx, cy, cz, signed r} (rt_bvh.cu).
void build_bvh(const double4 *spheres, uint32_t n, std::vector<BvhNode> &nodes,
               std::vector<uint32_t> &prim);

// The persistent path kernel: one launch per frame; carves n_slots = grid * block * 2 out of the pool.
// use_bvh: closest hit by BVH traversal (scene.bvh_nodes) instead of the brute-force filter.
// f32_hit: candidate resolution in f32 like src/gpu/sphere.h (RT_FLAG_F32_HIT; brute force only).
// use_tc: the brute-force filter as a tf32 contraction on the tensor cores (scene.tc_op), one CTA of
// 512 threads per SM; else the SIMT filter loop (FFMA2, rows from the constant bank).
cudaError_t run_fused(WaveParams p, void *pool, uint32_t pool_slots, const float4 *host_rows,
                      bool use_bvh, bool f32_hit, bool use_tc, int sm_count, size_t smem_optin_max,
                      cudaStream_t st, WaveRunInfo *info);

// Sphere operand of the tensor-core filter (rt_tc.cu): {cx, cy, cz, signed r} rows -> 3xTF32 split
// feature rows in the K-major shared-memory layout the MMA descriptor names.
void tc_build_sphere_operand(const double4 *exact, uint32_t n, std::vector<float> &out, uint32_t &tiles);
constexpr uint32_t kTcMaxSpheres = 960;  // the operand (128 B per sphere) must fit one CTA's shared memory

// Pixel hand-out order by descending cost (ray segments per sample in `partial`), stable.
// heavy_only: only the heavy pixels move to the front, the others keep their scan order (BVH
// mode: neighbouring lanes then trace neighbouring pixels, and coherent rays share nodes).
size_t cost_order_temp_bytes(uint32_t npix);
constexpr unsigned kHeavyRays = 5;  // rays per sample from which a pixel counts as "heavy"
cudaError_t launch_cost_order(const float4 *partial, uint32_t npix, float inv_samples, uint32_t *keys,
                              uint32_t *keys_out, uint32_t *iota, uint32_t *order, uint32_t *heavy,
                              bool heavy_only, void *temp, size_t temp_bytes, int sm_count,
                              cudaStream_t st);

// dst[i] = canonical sub-tree sum of slices [slice_begin, slice_end) (aligned power of 2).
cudaError_t launch_reduce_local(const float4 *accum, uint32_t npix, uint32_t slice_begin,
                                uint32_t slice_end, float4 *dst, int sm_count, cudaStream_t st);

// partials: `count` device pointers (count in {1,2,4,8}) in rank order.
cudaError_t launch_finalize(const float4 *const *partials, uint32_t count, uint32_t npix,
                            float scale, bool gamma, uint32_t *rgb_words, float *linear,
                            unsigned long long *rays, int sm_count, cudaStream_t st);

cudaError_t launch_primary_f32(const SceneView &sc, const CameraF32 &cam, uint32_t W, uint32_t H,
                               int32_t *ids, double *t, double *normal, uint8_t *ff,
                               cudaStream_t st);

// FFMA throughput micro-kernel: blocks x 256 threads x iters x 64 FFMA.  out: blocks*256 floats.
cudaError_t launch_ffma_peak(float *out, int blocks, int iters, long long *cycles,
                             cudaStream_t st);

// rt_primary_f64.cu (compiled with -fmad=false): op-for-op IEEE double gate kernel.
cudaError_t launch_primary_f64(const double4 *exact, uint32_t n, const CameraF64 &cam, uint32_t W,
                               uint32_t H, int32_t *ids, double *t, double *normal, uint8_t *ff,
                               cudaStream_t st);

}  // namespace rtb
