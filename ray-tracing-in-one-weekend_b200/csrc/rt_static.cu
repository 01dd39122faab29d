// csrc/rt_static.cu — rt_primary_hits_static: the primary-hit table (rt_primary_hits, rt_capi.cu)
// for callers that may not allocate.  No counterpart in the reference; added
// (INTEGRATION.md section 6).
//
// rt_primary_hits stages its four result arrays in device buffers it cudaMallocs and frees per
// call.  Here the same kernels (primary_f64_kernel / primary_f32_kernel, unchanged) write into a
// file-scope __device__ array instead: RT_STATIC_MAX_PIXELS * 37 bytes that every GPU which loads
// the library holds until the process exits.  The array's name is a host-side stand-in in host
// code, so its device address is taken ONCE per device with cudaGetSymbolAddress (with that device
// current) and that pointer is what the kernels get.
//
// The entry point calls no cudaMalloc / cudaFree / cudaMallocHost and creates no stream or event.
// Every call shares the array: a process-wide lock is held for the whole (synchronous) call, so two
// host threads cannot overlap on the GPU; anything that did overlap would race on the array.
#include <mutex>

#include "rt_capi_internal.h"

using namespace rtb;

namespace {

constexpr size_t kMaxPx = RT_STATIC_MAX_PIXELS;
// layout (all offsets multiples of 256): t[P] f64 | normal[3P] f64 | ids[P] i32 | front_face[P] u8
constexpr size_t kOffT = 0, kOffN = kOffT + kMaxPx * 8, kOffIds = kOffN + kMaxPx * 24,
                 kOffFf = kOffIds + kMaxPx * 4, kScratchBytes = kOffFf + kMaxPx;
static_assert(kMaxPx % 256 == 0, "section offsets must stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
unsigned char *g_static_ptr[kMaxDevices];  // device address of g_static_scratch, per device

// device address of the array on `device` (which must be current); looked up once per device
int static_scratch(int device, unsigned char **out) {
  if (device < 0 || device >= kMaxDevices) return rt_fail(RT_E_INVALID, "device %d out of range", device);
  if (!g_static_ptr[device]) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p)
      return rt_fail(RT_E_CUDA, "cudaGetSymbolAddress(static scratch) on device %d failed: %s", device,
                     cudaGetErrorString(e));
    g_static_ptr[device] = static_cast<unsigned char *>(p);
  }
  *out = g_static_ptr[device];
  return RT_OK;
}

}  // namespace

extern "C" int rt_primary_hits_static(const rt_scene *sc, const rt_camera *cam, uint32_t W, uint32_t H,
                                      int32_t device, int use_f64, void *cuda_stream, int32_t *ids,
                                      double *t, double *normal, uint8_t *front_face) {
  if (!sc || !cam || !ids || !t || !normal || !front_face) return rt_fail(RT_E_INVALID, "NULL argument");
  if (W < 2 || H < 2) return rt_fail(RT_E_INVALID, "image size %ux%u out of range (min 2x2)", W, H);
  const size_t npix = (size_t)W * H;
  if (npix > kMaxPx)
    return rt_fail(RT_E_INVALID,
                   "rt_primary_hits_static: %ux%u = %zu pixels exceed the static scratch "
                   "(RT_STATIC_MAX_PIXELS = %zu); use rt_primary_hits",
                   W, H, npix, kMaxPx);
  if (device != sc->device) return rt_fail(RT_E_INVALID, "scene lives on another device");
  std::lock_guard<std::mutex> lk(g_static_mu);
  CU(cudaSetDevice(device));
  unsigned char *base = nullptr;
  int rc = static_scratch(device, &base);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  // an rt_scene_update on another stream may still be in flight
  if (sc->staged_pending) CU(cudaEventSynchronize(sc->staged));
  double *d_t = reinterpret_cast<double *>(base + kOffT), *d_n = reinterpret_cast<double *>(base + kOffN);
  int32_t *d_ids = reinterpret_cast<int32_t *>(base + kOffIds);
  uint8_t *d_ff = base + kOffFf;
  // The kernels write every element of [0, npix); the ids of the last call are wiped anyway so that
  // a launch that did not run cannot hand back the previous table (-2 is never a valid id).
  CU(cudaMemsetAsync(d_ids, 0xfe, npix * 4, st));
  CamFrame cf;
  rt_derive_camera(cam, &cf);
  if (use_f64) {
    CameraF64 c64;
    for (int k = 0; k < 3; ++k) {
      c64.origin[k] = cf.origin[k]; c64.lower_left[k] = cf.llc[k];
      c64.horizontal[k] = cf.hor[k]; c64.vertical[k] = cf.ver[k];
    }
    CU(launch_primary_f64(sc->exact, sc->n, c64, W, H, d_ids, d_t, d_n, d_ff, st));
  } else {
    CU(launch_primary_f32(rt_scene_view(sc), rt_narrow_camera(cf, W, H, false), W, H, d_ids, d_t, d_n, d_ff, st));
  }
  CU(cudaMemcpyAsync(ids, d_ids, npix * 4, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(t, d_t, npix * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(normal, d_n, npix * 24, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(front_face, d_ff, npix, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return RT_OK;
}
