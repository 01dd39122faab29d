// csrc/rt_capi_internal.h — the handle structs and enqueue-level helpers shared by the host-side
// translation units of the C ABI (rt_capi.cu, rt_multi.cu, rt_output.cu).  Internal to the .so.
#pragma once

#include <mutex>
#include <vector>

#include "../../include/rt_b200.h"
#include "rt_kernels.h"

int rt_fail(int code, const char *fmt, ...);

#define CU(expr)                                                                         \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess)                                                              \
      return rt_fail(RT_E_CUDA, "CUDA error %d (%s) at %s:%d '%s'", (int)e__,            \
                     cudaGetErrorString(e__), __FILE__, __LINE__, #expr);                \
  } while (0)

// CU with a cleanup statement: half-built objects are destroyed before the error is returned
#define CU_OR(cleanup, expr)                                                             \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess) {                                                            \
      cleanup;                                                                           \
      return rt_fail(RT_E_CUDA, "CUDA error %d (%s) at %s:%d '%s'", (int)e__,            \
                     cudaGetErrorString(e__), __FILE__, __LINE__, #expr);                \
    }                                                                                    \
  } while (0)

struct rt_scene {
  int device = 0;
  uint32_t n = 0, n_pad = 0, cap = 0;  // spheres, rows in use (padded), rows allocated
  unsigned char *blob = nullptr;       // ONE device allocation: exact | fast | mat | aux
  unsigned char *stage = nullptr;      // pinned host staging buffer of the same layout
  size_t blob_bytes = 0;
  double4 *exact = nullptr;            // device views into blob
  float4 *fast = nullptr, *mat = nullptr;
  float2 *aux = nullptr;
  std::vector<float4> host_fast;       // host copy of the filter rows (kernel-parameter path)
  cudaEvent_t staged = nullptr;        // completion of the last asynchronous upload
  bool staged_pending = false;
  // optional hierarchy, built on the host at the first RT_FLAG_BVH render
  std::mutex bvh_mu;
  bool bvh_built = false;
  rtb::BvhNode *bvh_nodes = nullptr;
  uint32_t *bvh_prim = nullptr;
  uint32_t bvh_node_count = 0;
  // sphere operand of the tensor-core filter, built on the host at the first render that uses it
  bool tc_built = false;
  float *tc_op = nullptr;
  uint32_t tc_tiles = 0;
};

struct rt_frame {
  int device;
  uint32_t w, h, npix;
  float4 *accum;    // [RT_SLICES][npix]
  float4 *partial;  // [npix]
  uint32_t *rgb_words;
  float *linear;
  unsigned long long *counters;    // device: [0] work queue head, [1] ray total, [2] tensor-core status
  unsigned long long *host_words;  // pinned: [1] ray total of the last finalise, [2] tensor-core status
  cudaStream_t stream;      // stream in use (own_stream unless rt_frame_set_stream)
  cudaStream_t own_stream;
  cudaEvent_t ev[4];        // render begin/end, finalise begin/end
  int sm_count;
  size_t smem_optin;
  uint32_t last_n;          // spheres of the last render
  uint64_t last_samples;
  uint32_t last_launches;
  bool rendered, finalized;
  bool last_tc;             // the last render ran the tensor-core filter
  uint32_t last_filter;     // RT_FILTER_* of the last render (rt_stats.filter_unit)
  void *pool;               // path-slot pool (allocated on first use)
  uint32_t pool_slots;
  // hand-out order of the next frame: pixels by descending cost of the frame rendered before
  uint32_t *order, *sort_keys, *sort_keys_out, *sort_iota, *heavy;
  void *sort_temp;
  size_t sort_temp_bytes;
  bool order_valid;
  bool order_heavy_only;    // the order was made for BVH frames (heavy pixels first, the rest in scan order)
};

// camera::camera (src/cpu/camera.h:8-26) in double
struct CamFrame {
  double origin[3], llc[3], hor[3], ver[3], u[3], v[3], w[3], lens_radius;
};
void rt_derive_camera(const rt_camera *c, CamFrame *f);
rtb::CameraF32 rt_narrow_camera(const CamFrame &f, uint32_t W, uint32_t H, bool pixel_square);
rtb::SceneView rt_scene_view(const rt_scene *sc);
int rt_scene_ensure_bvh(const rt_scene *sc);
int rt_scene_ensure_tc(const rt_scene *sc);

// enqueue-only forms (no host synchronisation in the fused mode): the public entry points add the
// synchronisation rule of rt_frame_set_stream on top
int rt_render_enqueue(const rt_scene *, const rt_camera *, const rt_config *, rt_frame *);
int rt_reduce_enqueue(rt_frame *, const rt_config *);
int rt_finalize_enqueue(rt_frame *, const rt_config *, const void *const *partials, uint32_t count,
                        uint8_t *rgb8_out, float *linear_out);
