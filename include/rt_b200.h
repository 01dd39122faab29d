/* include/rt_b200.h — C ABI of the B200-native path-tracing hot path.
 *
 * The reference (kouei/ray-tracing-in-one-weekend) has no FFI/plugin interface: its "API"
 * is the set of C++ types used by main().  This header is the boundary a
 * maintainer binds instead: plain pointers and sizes, no C++/torch types.  Each entry point
 * names the reference code it replaces (paths relative to the reference root).
 *
 * Conventions: every function returns 0 on success or a negative RT_E_* code; the message
 * of the last failure on the calling thread is rt_last_error().  Nothing ever calls exit()
 * (the reference's checkCudaErrors does, src/gpu/cuda_utility.h:8-18).  Handles are not
 * thread-safe; distinct handles may be used from distinct threads.  There is NO CPU
 * fallback: without a CUDA device every compute entry point fails with RT_E_CUDA.
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_OK 0
#define RT_E_INVALID (-1) /* bad argument */
#define RT_E_CUDA (-2)    /* CUDA runtime error / no device */
#define RT_E_NOMEM (-3)
#define RT_E_IO (-4)

#define RT_LAMBERTIAN 0u /* src/cpu/material.h:14-34 */
#define RT_METAL 1u      /* src/cpu/material.h:36-51 */
#define RT_DIELECTRIC 2u /* src/cpu/material.h:53-87 */
/* the diffuse rules of the book's earlier steps (no material classes yet; albedo = the 0.5 of the listing) */
#define RT_DIFFUSE_BALL 3u /* direction = normal + random_in_unit_sphere(), archive/listing33/main.cc:12-13 */
#define RT_DIFFUSE_HEMI 4u /* direction = random_in_hemisphere(normal), archive/listing40/main.cc:26 */
#define RT_KIND_MAX RT_DIFFUSE_HEMI

#define RT_SLICES 64u /* canonical sample slices per pixel (multi-GPU split) */

/* One sphere with its material, in the reference's insertion order
 * (hittable_list::add(make_shared<sphere>(center, radius, material)),
 *  src/cpu/hittable_list.h:20, src/cpu/sphere.h:11-12).  radius keeps its sign
 *  (negative = hollow bubble, archive/listing69/main.cc:61). */
typedef struct rt_sphere {
  double center[3];
  double radius;
  uint32_t kind;      /* RT_LAMBERTIAN | RT_METAL | RT_DIELECTRIC */
  uint32_t _pad;
  double albedo[3];   /* ignored for dielectric (attenuation is 1,1,1) */
  double fuzz_or_ir;  /* metal: fuzz (clamped to <=1 like material.h:38); dielectric: ir */
} rt_sphere;

/* Exactly the constructor arguments of camera, src/cpu/camera.h:8-10. */
typedef struct rt_camera {
  double lookfrom[3], lookat[3], vup[3];
  double vfov_deg, aspect_ratio, aperture, focus_dist;
} rt_camera;

/* Render parameters: the literals of main(), src/cpu/main.cc:82-86. */
typedef struct rt_config {
  uint32_t width, height;
  uint32_t samples_per_pixel; /* total spp of the frame */
  uint32_t max_depth;         /* 50 in the reference */
  uint64_t seed;              /* frame seed of the Philox stream */
  uint32_t sample_offset;     /* added to the sample index in the RNG key (progressive) */
  uint32_t slice_begin, slice_end; /* canonical slices [begin,end) of RT_SLICES this call renders */
  int32_t device;             /* CUDA device ordinal */
  uint32_t flags;             /* RT_FLAG_* */
} rt_config;

#define RT_FLAG_COUNT_RAYS 1u /* fill rt_stats.rays (one atomic per work item) */
#define RT_FLAG_BVH 8u        /* closest hit through the host-built BVH instead of the brute-force scan
                                 over the reference's sphere order (same hit, same image) */
#define RT_FLAG_V4_MODEL 16u  /* the semantics of the reference's CUDA program (src/gpu/, v4 book) instead
                                 of src/cpu/: pixel-square sampling over W x H pixel footprints
                                 (src/gpu/camera.h:140-167), metal fuzz point ON the unit sphere
                                 (src/gpu/material.h:52), strict root interval so that the EARLIER of two
                                 spheres with equal roots wins (src/gpu/sphere.h:31-36, interval.h:19-21).
                                 rt_camera.aperture = 2 * focus_dist * tan(defocus_angle / 2). */
#define RT_FLAG_F32_HIT 32u   /* compatibility/diagnostic: resolve the filter's candidates with the f32
                                 sphere::hit arithmetic of src/gpu/sphere.h:15-47 (un-normalised direction,
                                 f32 hit point, no self-skip) instead of f64.  Reproduces the cancellation of
                                 |oc|^2 - r^2 on the r = 1000 ground sphere that darkens the reference CUDA
                                 program's ground; brute-force scan only (not with RT_FLAG_BVH). */
#define RT_FLAG_NO_FILTER 64u /* debug/soundness switch: the f32 line-sphere pre-test of the brute-force scan is
                                 switched off and every ray resolves EVERY sphere in f64 (hittable_list::hit as
                                 written, src/cpu/hittable_list.h:28-43).  The image must be byte-identical with
                                 and without it: that is the test of the filter's conservativeness
                                 (tests/test_filter_soundness.py).  ~50x slower; brute-force scan only. */
#define RT_FLAG_LINEAR_OUTPUT 128u /* rt_finalize quantises sum/spp without the gamma sqrt: the write_color of the
                                 book's first diffuse steps (archive/listing30/color.h:8-17, before listing35;
                                 archive-gpu/image7...11/color.h) */
#define RT_FLAG_SIMT_FILTER 256u /* brute-force scan with the FFMA2 filter loop (CUDA cores) even where the
                                 tensor-core form applies.  By default scenes of 32 spheres and more run the
                                 filter as one tf32 contraction on the tensor cores (tcgen05);
                                 both find every true hit, so the image is byte-identical either way. */
#define RT_FLAG_ALL (RT_FLAG_COUNT_RAYS | RT_FLAG_BVH | RT_FLAG_V4_MODEL | RT_FLAG_F32_HIT | RT_FLAG_NO_FILTER | \
                     RT_FLAG_LINEAR_OUTPUT | RT_FLAG_SIMT_FILTER)

typedef struct rt_stats {
  uint64_t samples;      /* samples traced by this call */
  uint64_t rays;         /* ray segments (ray_color invocations, src/cpu/main.cc:12) */
  uint64_t sphere_tests; /* rays * sphere count (brute force) */
  float render_ms;       /* device time of the path kernel (CUDA events) */
  float finalize_ms;     /* device time of reduce+finalise */
  uint32_t launches;     /* kernels launched by this call */
  float frame_ms;        /* device time from the start of rt_render to the end of the finalise
                            kernel on the frame's stream (set by rt_finalize) */
  uint32_t filter_unit;  /* where the brute-force line-sphere filter of the last render ran:
                            RT_FILTER_SIMT (FFMA2 loop), RT_FILTER_TENSOR (tcgen05 3xTF32
                            contraction) or RT_FILTER_NONE (BVH traversal / RT_FLAG_NO_FILTER) */
} rt_stats;
#define RT_FILTER_NONE 0u
#define RT_FILTER_SIMT 1u
#define RT_FILTER_TENSOR 2u

typedef struct rt_scene rt_scene;   /* device-resident flattened hittable_list */
typedef struct rt_frame rt_frame;   /* device-resident accumulation buffers of one frame */

const char *rt_last_error(void);
/* Library/ABI version and the CUDA arch it was built for ("sm_100a"). */
const char *rt_version(void);
int rt_device_count(int *count);

/* --- scene: replaces hittable_list + sphere + material objects (src/cpu/hittable_list.h,
 *     sphere.h, material.h) and the GPU twin's new_world kernel (src/gpu/main.cu:18-75).
 *     One device allocation, one host->device copy from a pinned staging buffer the scene owns.
 *     The optional BVH is built on the host at the first RT_FLAG_BVH render, not here. */
int rt_scene_create(const rt_sphere *spheres, uint32_t n, int32_t device, rt_scene **out);
int rt_scene_destroy(rt_scene *);
int rt_scene_count(const rt_scene *, uint32_t *n);
/* Replace the spheres of an existing scene (hittable_list::clear + add, hittable_list.h:17-20) without
 * re-allocating: n must fit the capacity the scene was created with (its n rounded up to 8).  The
 * rows are flattened into the scene's pinned staging buffer and copied with ONE cudaMemcpyAsync on
 * `cuda_stream` (a cudaStream_t; NULL = the default stream, synchronous); renders ordered after it
 * on the same stream see the new spheres.  This is the per-frame H2D of an animation loop. */
int rt_scene_update(rt_scene *, const rt_sphere *spheres, uint32_t n, void *cuda_stream);

/* --- frame: accumulation buffers [RT_SLICES][H][W] float4 on the device. */
int rt_frame_create(uint32_t width, uint32_t height, int32_t device, rt_frame **out);
int rt_frame_destroy(rt_frame *);
/* Device pointer / byte size of the per-GPU partial sum (the pre-summed sub-tree of the
 * slices this GPU rendered): [H][W] float4.  This is what crosses NVLink. */
int rt_frame_partial_ptr(rt_frame *, void **dev_ptr, size_t *bytes);

/* Run the frame's work on a caller-owned CUDA stream (a cudaStream_t; NULL = the frame's own
 * stream) so that collectives and kernels order on one stream.  While a caller-owned stream is
 * set, rt_render, rt_frame_reduce_local and rt_finalize only ENQUEUE work and return without any
 * host synchronisation: their rt_stats carry the host-known fields only (samples, launches), the
 * output buffers of rt_finalize must be pinned host memory to stay asynchronous, and both the
 * outputs and the timings / ray counts (rt_frame_stats) are valid once the caller has
 * synchronised the stream.  With the frame's own stream every call returns with its work done. */
int rt_frame_set_stream(rt_frame *, void *cuda_stream);
/* Timings (CUDA events) and device counters of the last rt_render / rt_finalize on this frame.
 * Call after the frame's stream has been synchronised (it synchronises on the recorded events). */
int rt_frame_stats(rt_frame *, rt_stats *);

/* --- the hot path: replaces the pixel/sample loops + ray_color + everything below
 *     (src/cpu/main.cc:12-30,111-123; GPU twin: render kernel, src/gpu/camera.h:169-195).
 *     Renders slices [slice_begin, slice_end) of every pixel into the frame. */
int rt_render(const rt_scene *, const rt_camera *, const rt_config *, rt_frame *, rt_stats *);

/* Pre-sum this GPU's slices in the canonical pairwise tree into the partial buffer. */
int rt_frame_reduce_local(rt_frame *, const rt_config *);

/* Combine `count` partial buffers (device pointers on cfg->device, in rank order; count in
 * {1,2,4,8}) with the upper levels of the same tree, then sum/spp -> sqrt -> clamp ->
 * 8-bit (write_color, src/cpu/color.h:8-23).  rgb8_out is HOST memory [H][W][3], rows
 * top->bottom; linear_out (optional, HOST, [H][W][3] float) receives the linear sums. */
int rt_finalize(rt_frame *, const rt_config *, const void *const *partials, uint32_t count,
                uint8_t *rgb8_out, float *linear_out, rt_stats *);

/* Convenience: render all slices on one GPU + reduce + finalise (host buffers out).  The frame
 * object (0.83 GB of slice planes at 1200x675) is cached per calling thread and re-used by the
 * next call with the same size and device; rt_release_cached_frames() frees the calling thread's. */
int rt_render_image(const rt_scene *, const rt_camera *, const rt_config *, uint8_t *rgb8_out,
                    float *linear_out, rt_stats *);
int rt_release_cached_frames(void);

/* --- one process, several GPUs (ngpu in {1,2,4,8}); the reference has no multi-device path
 *     (SURVEY.md §8e).  rt_multi owns one scene and one frame per device, set up once; a frame's
 *     samples are split by canonical slices (device g renders slices [64g/G, 64(g+1)/G)), each GPU
 *     pre-sums its sub-tree, and the finalise kernel on devices[0] reads the other GPUs' partial
 *     buffers in place through NVLink peer pointers (no staging buffer, no copy; without peer
 *     access they are copied to devices[0] first).  Everything is enqueued from the calling thread
 *     and the call returns after ONE host synchronisation.  The image is byte-identical for every
 *     ngpu.  (Multi-process deployments use rt_render + rt_frame_partial_ptr + rt_finalize with
 *     their own transport instead, see INTEGRATION.md.) */
typedef struct rt_multi rt_multi;
int rt_multi_create(const int32_t *devices, uint32_t ngpu, uint32_t width, uint32_t height, rt_multi **out);
/* Upload the scene to every device (first call allocates; later calls with a sphere count that
 * fits re-use the allocations: one asynchronous copy per device, ordered before the next render). */
int rt_multi_set_scene(rt_multi *, const rt_sphere *spheres, uint32_t n);
int rt_multi_render(rt_multi *, const rt_camera *, const rt_config *, uint8_t *rgb8_out,
                    float *linear_out, rt_stats *);
int rt_multi_destroy(rt_multi *);
/* Convenience: rt_multi_create + set_scene + render + destroy around one frame. */
int rt_render_image_multi(const rt_sphere *spheres, uint32_t n, const rt_camera *, const rt_config *,
                          const int32_t *devices, uint32_t ngpu, uint8_t *rgb8_out,
                          float *linear_out, rt_stats *);

/* --- accuracy gate: pixel-centre, zero-lens primary rays, closest hit in reference order
 *     (hittable_list::hit + sphere::hit, src/cpu/hittable_list.h:28-43, sphere.h:24-51).
 *     use_f64 = 1: IEEE double without FMA contraction, op-for-op like the reference
 *     (ids bit-exact, normals <= 2 ulp); use_f64 = 0: the production f32 intersection.
 *     Outputs are HOST arrays, rows top->bottom; t/normal are double in both modes. */
int rt_primary_hits(const rt_scene *, const rt_camera *, uint32_t width, uint32_t height,
                    int32_t device, int use_f64, int32_t *ids, double *t, double *normal,
                    uint8_t *front_face);

/* --- rt_primary_hits for callers that may not allocate (code under CUDA graph capture, serving
 *     loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 *     Same kernels, same outputs (HOST arrays) and same results
 *     as rt_primary_hits, but the device-side result table lives in a static __device__ array of
 *     RT_STATIC_MAX_PIXELS * 37 bytes (4.6 MiB) that every GPU which loads the library holds until
 *     the process exits; the call makes no cudaMalloc / cudaFree / cudaMallocHost and creates no
 *     stream or event.  width * height > RT_STATIC_MAX_PIXELS is rejected with RT_E_INVALID.
 *     Work is enqueued on `cuda_stream` (a cudaStream_t, NULL = default stream) and the call
 *     returns after that stream has drained.  All calls share the one array per GPU: the library
 *     serialises them with a process-wide lock; calls that overlapped on the GPU would race. */
#define RT_STATIC_MAX_PIXELS 131072u
int rt_primary_hits_static(const rt_scene *, const rt_camera *, uint32_t width, uint32_t height,
                           int32_t device, int use_f64, void *cuda_stream, int32_t *ids, double *t,
                           double *normal, uint8_t *front_face);

/* --- measurement aid: sustained FP32 FFMA throughput of the device (TFLOP/s, FMA = 2),
 *     measured with a register-resident micro-kernel; the roofline denominator of bench.py. */
int rt_measure_fp32_peak(int32_t device, float *tflops, float *sm_clock_mhz_est);

/* --- measurement / validation aid: the brute-force filter written as ONE tf32 contraction on the
 *     tensor cores (tcgen05.mma, accumulators in tensor memory) on `n` rays
 *     {ox, oy, oz, |d| > 0}, {unit d, -} against a scene of up to 960 spheres.  masks: HOST
 *     [n][tiles] 64-bit words, bit b of word t set = sphere tile_spheres * t + b passes; values
 *     (optional): HOST [n_values][tiles][tile_spheres] raw filter values; ms: device time of
 *     `repeats` passes over the rays.  Not part of the render path. */
int rt_tc_filter_probe(const rt_scene *, const float *ray_o, const float *ray_d, uint32_t n, uint64_t *masks,
                       uint32_t *tiles, uint32_t *tile_spheres, float *values, uint32_t n_values,
                       uint32_t repeats, float *ms);

/* --- output stage (the step after the path): "P3\nW H\n255\n" + "r g b\n" per pixel, rows
 *     top->bottom (src/cpu/main.cc:109, color.h:20-22; the multi-threaded program's loop,
 *     src/cpu-multi-threading/main.cc:148-154), or binary P6.  path "-" = stdout. */
#define RT_PNM_P3 3 /* text, the reference's format */
#define RT_PNM_P6 6 /* binary */
int rt_write_ppm(const char *path, const uint8_t *rgb8, uint32_t width, uint32_t height); /* P3 */
int rt_write_pnm(const char *path, const uint8_t *rgb8, uint32_t width, uint32_t height, int format);
/* Finalise in bands of rows: the finalise kernel runs once, then the 8-bit rows cross to the host
 * through two pinned buffers and `fn` consumes band k (rows [row_begin, row_end), top->bottom,
 * tightly packed RGB8) while the copy of band k+1 is in flight.  fn returns 0 to continue.
 * Synchronous on every stream.  band_rows = 0 picks 64. */
typedef int (*rt_band_fn)(void *user, const uint8_t *rgb8_rows, uint32_t row_begin, uint32_t row_end);
int rt_finalize_bands(rt_frame *, const rt_config *, const void *const *partials, uint32_t count,
                      uint32_t band_rows, rt_band_fn fn, void *user, rt_stats *);
/* rt_finalize_bands with a PNM writer as the consumer: text formatting / fwrite of band k overlaps
 * the device->host copy of band k+1; no whole-image host buffer. */
int rt_finalize_to_file(rt_frame *, const rt_config *, const void *const *partials, uint32_t count,
                        const char *path, int format, rt_stats *);

#ifdef __cplusplus
}
#endif
#endif
