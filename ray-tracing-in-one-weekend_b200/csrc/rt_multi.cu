// csrc/rt_multi.cu — one process, several GPUs, without torch (the reference
// has no multi-device path).  A persistent object owns one scene and one frame per device; a
// frame's samples are split by canonical slices, every GPU pre-sums its sub-tree, and device 0's
// finalise kernel reads the other GPUs' partial buffers DIRECTLY through peer pointers over NVLink
// (no staging allocation, no copy kernel: the gather is fused into the finalise).  Where the
// topology has no peer access the partials are copied to device 0 first.  Everything is enqueued
// from the calling thread; the only host synchronisation is the one at the end of a frame.
#include <cstring>
#include <new>

#include "rt_capi_internal.h"

struct rt_multi {
  uint32_t ngpu = 0, w = 0, h = 0;
  int devices[8] = {};
  rt_scene *scenes[8] = {};
  rt_frame *frames[8] = {};
  bool peer[8] = {};         // device 0 reads device g's partial in place
  float4 *staged[8] = {};    // copy target on device 0 when it cannot
  cudaEvent_t done[8] = {};  // device g's partial is complete
};

extern "C" int rt_multi_destroy(rt_multi *m) {
  if (!m) return RT_OK;
  for (uint32_t g = 0; g < m->ngpu; ++g) {
    if (m->staged[g]) { cudaSetDevice(m->devices[0]); cudaFree(m->staged[g]); }
    if (m->done[g]) { cudaSetDevice(m->devices[g]); cudaEventDestroy(m->done[g]); }
    rt_frame_destroy(m->frames[g]);
    rt_scene_destroy(m->scenes[g]);
  }
  delete m;
  return RT_OK;
}

extern "C" int rt_multi_create(const int32_t *devices, uint32_t ngpu, uint32_t width, uint32_t height,
                               rt_multi **out) {
  if (!out) return rt_fail(RT_E_INVALID, "out is NULL");
  *out = nullptr;
  if (!devices) return rt_fail(RT_E_INVALID, "devices is NULL");
  if (!(ngpu == 1 || ngpu == 2 || ngpu == 4 || ngpu == 8))
    return rt_fail(RT_E_INVALID, "ngpu must be 1, 2, 4 or 8 (got %u)", ngpu);
  for (uint32_t a = 0; a < ngpu; ++a)
    for (uint32_t b = a + 1; b < ngpu; ++b)
      if (devices[a] == devices[b]) return rt_fail(RT_E_INVALID, "device %d listed twice", devices[a]);
  rt_multi *m = new (std::nothrow) rt_multi();
  if (!m) return rt_fail(RT_E_NOMEM, "out of host memory");
  m->ngpu = ngpu; m->w = width; m->h = height;
  for (uint32_t g = 0; g < ngpu; ++g) m->devices[g] = devices[g];
  for (uint32_t g = 0; g < ngpu; ++g) {
    int rc = rt_frame_create(width, height, devices[g], &m->frames[g]);
    if (rc) { rt_multi_destroy(m); return rc; }
    CU_OR(rt_multi_destroy(m), cudaEventCreateWithFlags(&m->done[g], cudaEventDisableTiming));
  }
  CU_OR(rt_multi_destroy(m), cudaSetDevice(devices[0]));
  const size_t bytes = (size_t)width * height * sizeof(float4);
  for (uint32_t g = 1; g < ngpu; ++g) {
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, devices[0], devices[g]) == cudaSuccess && can) {
      cudaError_t pe = cudaDeviceEnablePeerAccess(devices[g], 0);
      if (pe == cudaSuccess || pe == cudaErrorPeerAccessAlreadyEnabled) m->peer[g] = true;
      (void)cudaGetLastError();
    }
    if (!m->peer[g]) CU_OR(rt_multi_destroy(m), cudaMalloc(&m->staged[g], bytes));
  }
  *out = m;
  return RT_OK;
}

extern "C" int rt_multi_set_scene(rt_multi *m, const rt_sphere *spheres, uint32_t n) {
  if (!m) return rt_fail(RT_E_INVALID, "NULL argument");
  for (uint32_t g = 0; g < m->ngpu; ++g) {
    int rc;
    if (m->scenes[g] && (n + kRowPad - 1) / kRowPad * kRowPad <= m->scenes[g]->cap) {
      // same capacity: one asynchronous copy on the frame's stream, ordered before the next render
      rc = rt_scene_update(m->scenes[g], spheres, n, m->frames[g]->stream);
    } else {
      rt_scene_destroy(m->scenes[g]);
      m->scenes[g] = nullptr;
      rc = rt_scene_create(spheres, n, m->devices[g], &m->scenes[g]);
    }
    if (rc) return rc;
  }
  return RT_OK;
}

extern "C" int rt_multi_render(rt_multi *m, const rt_camera *cam, const rt_config *cfg_in,
                               uint8_t *rgb8_out, float *linear_out, rt_stats *stats) {
  if (!m || !cam || !cfg_in) return rt_fail(RT_E_INVALID, "NULL argument");
  if (!m->scenes[0]) return rt_fail(RT_E_INVALID, "rt_multi_set_scene has not been called");
  if (cfg_in->width != m->w || cfg_in->height != m->h)
    return rt_fail(RT_E_INVALID, "config %ux%u does not match the %ux%u frames", cfg_in->width, cfg_in->height, m->w, m->h);
  const uint32_t G = m->ngpu, per = RT_SLICES / G;
  rt_config c[8];
  for (uint32_t g = 0; g < G; ++g) {
    c[g] = *cfg_in;
    c[g].device = m->devices[g];
    c[g].slice_begin = g * per; c[g].slice_end = (g + 1) * per;
  }
  // path kernels first, on every device, so that the GPUs start within a few launch latencies
  int rc = RT_OK;
  for (uint32_t g = 0; g < G && !rc; ++g) rc = rt_render_enqueue(m->scenes[g], cam, &c[g], m->frames[g]);
  for (uint32_t g = 0; g < G && !rc; ++g) {
    rc = rt_reduce_enqueue(m->frames[g], &c[g]);
    if (!rc) CU(cudaEventRecord(m->done[g], m->frames[g]->stream));
  }
  if (rc) {
    for (uint32_t g = 0; g < G; ++g) { cudaSetDevice(m->devices[g]); cudaStreamSynchronize(m->frames[g]->stream); }
    return rc;
  }
  // device 0 waits for the partials and finishes the tree reading them where they lie
  rt_frame *f0 = m->frames[0];
  CU(cudaSetDevice(m->devices[0]));
  const void *parts[8];
  const size_t bytes = (size_t)f0->npix * sizeof(float4);
  parts[0] = f0->partial;
  for (uint32_t g = 1; g < G; ++g) {
    CU(cudaStreamWaitEvent(f0->stream, m->done[g], 0));
    if (m->peer[g]) {
      parts[g] = m->frames[g]->partial;
    } else {
      CU(cudaMemcpyPeerAsync(m->staged[g], m->devices[0], m->frames[g]->partial, m->devices[g], bytes, f0->stream));
      parts[g] = m->staged[g];
    }
  }
  rt_config call = *cfg_in;
  call.device = m->devices[0]; call.slice_begin = 0; call.slice_end = RT_SLICES;
  rc = rt_finalize_enqueue(f0, &call, parts, G, rgb8_out, linear_out);
  if (rc) return rc;
  CU(cudaStreamSynchronize(f0->stream));
  if (stats) {
    rt_stats fin;
    rc = rt_frame_stats(f0, &fin);
    if (rc) return rc;
    *stats = fin;
    stats->samples = 0; stats->render_ms = 0.f; stats->launches = 1;
    for (uint32_t g = 0; g < G; ++g) {
      rt_stats sg;
      rt_frame *fg = m->frames[g];
      CU(cudaSetDevice(m->devices[g]));
      CU(cudaEventSynchronize(fg->ev[1]));
      memset(&sg, 0, sizeof sg);
      CU(cudaEventElapsedTime(&sg.render_ms, fg->ev[0], fg->ev[1]));
      stats->samples += fg->last_samples;
      stats->filter_unit = fg->last_filter;
      stats->launches += fg->last_launches + 1;
      if (sg.render_ms > stats->render_ms) stats->render_ms = sg.render_ms;
    }
    stats->frame_ms = stats->render_ms + fin.finalize_ms;
  }
  return RT_OK;
}

// convenience: one frame, objects created and destroyed around it
extern "C" int rt_render_image_multi(const rt_sphere *spheres, uint32_t n, const rt_camera *cam,
                                     const rt_config *cfg_in, const int32_t *devices, uint32_t ngpu,
                                     uint8_t *rgb8_out, float *linear_out, rt_stats *stats) {
  if (!cam || !cfg_in || !devices) return rt_fail(RT_E_INVALID, "NULL argument");
  rt_multi *m = nullptr;
  int rc = rt_multi_create(devices, ngpu, cfg_in->width, cfg_in->height, &m);
  if (!rc) rc = rt_multi_set_scene(m, spheres, n);
  if (!rc) rc = rt_multi_render(m, cam, cfg_in, rgb8_out, linear_out, stats);
  rt_multi_destroy(m);
  return rc;
}
