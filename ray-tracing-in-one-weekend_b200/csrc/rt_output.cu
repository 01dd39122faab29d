// csrc/rt_output.cu — the output stage after the hot path: PPM writers
// and the banded finalise that overlaps the device->host copy of band k+1 with the formatting of
// band k.  Replaces the iostream loops of src/cpu/main.cc:109, color.h:20-22 and
// src/cpu-multi-threading/main.cc:148-154 (three operator<< per pixel: 8.3 M lines at 4K).
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

#include "rt_capi_internal.h"

namespace {

// "0" .. "255" without snprintf in the pixel loop
struct DigitTable {
  char txt[256][4];
  unsigned char len[256];
};
const DigitTable &digits() {
  static DigitTable t;
  static std::once_flag once;
  std::call_once(once, [] {
    for (int v = 0; v < 256; ++v) t.len[v] = (unsigned char)snprintf(t.txt[v], 4, "%d", v);
  });
  return t;
}

// "r g b\n" per pixel (src/cpu/color.h:20-22) for `npx` pixels; returns the end of the text
char *format_p3(const uint8_t *p, size_t npx, char *o) {
  const DigitTable &t = digits();
  for (size_t x = 0; x < npx; ++x, p += 3)
    for (int c = 0; c < 3; ++c) {
      memcpy(o, t.txt[p[c]], 4);  // 4-byte copy, advance by the real length
      o += t.len[p[c]];
      *o++ = c == 2 ? '\n' : ' ';
    }
  return o;
}

struct PnmFile {
  FILE *fp = nullptr;
  bool close_it = false;
  int format = RT_PNM_P3;
  uint32_t W = 0;
  std::vector<char> line;
  const char *path = "";
};

int pnm_open(PnmFile &f, const char *path, uint32_t W, uint32_t H, int format) {
  if (format != RT_PNM_P3 && format != RT_PNM_P6) return rt_fail(RT_E_INVALID, "format must be RT_PNM_P3 or RT_PNM_P6");
  const bool to_stdout = strcmp(path, "-") == 0;
  f.fp = to_stdout ? stdout : fopen(path, format == RT_PNM_P6 ? "wb" : "w");
  if (!f.fp) return rt_fail(RT_E_IO, "cannot open %s", path);
  f.close_it = !to_stdout; f.format = format; f.W = W; f.path = path;
  fprintf(f.fp, "P%d\n%u %u\n255\n", format, W, H);  // src/cpu/main.cc:109
  return RT_OK;
}

int pnm_rows(PnmFile &f, const uint8_t *rgb8, uint32_t rows) {
  const size_t npx = (size_t)rows * f.W;
  size_t want, got;
  if (f.format == RT_PNM_P6) {
    want = npx * 3;
    got = fwrite(rgb8, 1, want, f.fp);
  } else {
    if (f.line.size() < npx * 12 + 4) f.line.resize(npx * 12 + 4);
    char *end = format_p3(rgb8, npx, f.line.data());
    want = (size_t)(end - f.line.data());
    got = fwrite(f.line.data(), 1, want, f.fp);
  }
  if (got != want) return rt_fail(RT_E_IO, "short write to %s", f.path);
  return RT_OK;
}

int pnm_close(PnmFile &f, int rc) {
  if (f.fp) {
    if (f.close_it) {
      if (fclose(f.fp) != 0 && rc == RT_OK) rc = rt_fail(RT_E_IO, "closing %s failed", f.path);
    } else {
      fflush(f.fp);
    }
    f.fp = nullptr;
  }
  return rc;
}

}  // namespace

extern "C" int rt_write_pnm(const char *path, const uint8_t *rgb8, uint32_t W, uint32_t H, int format) {
  if (!path || !rgb8) return rt_fail(RT_E_INVALID, "NULL argument");
  PnmFile f;
  int rc = pnm_open(f, path, W, H, format);
  if (rc) return rc;
  const uint32_t band = 64;  // rows per fwrite
  for (uint32_t y = 0; y < H && rc == RT_OK; y += band)
    rc = pnm_rows(f, rgb8 + (size_t)y * W * 3, H - y < band ? H - y : band);
  return pnm_close(f, rc);
}

extern "C" int rt_write_ppm(const char *path, const uint8_t *rgb8, uint32_t W, uint32_t H) {
  return rt_write_pnm(path, rgb8, W, H, RT_PNM_P3);
}

// ---------------------------------------------------------------- banded finalise
extern "C" int rt_finalize_bands(rt_frame *f, const rt_config *cfg, const void *const *partials,
                                 uint32_t count, uint32_t band_rows, rt_band_fn fn, void *user,
                                 rt_stats *stats) {
  if (!f || !fn) return rt_fail(RT_E_INVALID, "NULL argument");
  if (band_rows == 0) band_rows = 64;
  if (band_rows > f->h) band_rows = f->h;
  // finalise kernel only: the packed RGB8 image stays on the device
  int rc = rt_finalize_enqueue(f, cfg, partials, count, nullptr, nullptr);
  if (rc) return rc;
  const size_t row_bytes = (size_t)f->w * 3, band_bytes = row_bytes * band_rows;
  uint8_t *buf[2] = {nullptr, nullptr};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  cudaError_t e = cudaMallocHost(&buf[0], band_bytes);
  if (e == cudaSuccess) e = cudaMallocHost(&buf[1], band_bytes);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming);
  const uint32_t nb = (f->h + band_rows - 1) / band_rows;
  const uint8_t *dev = reinterpret_cast<const uint8_t *>(f->rgb_words);
  auto copy = [&](uint32_t k) -> cudaError_t {
    const uint32_t r0 = k * band_rows, r1 = r0 + band_rows < f->h ? r0 + band_rows : f->h;
    cudaError_t ce = cudaMemcpyAsync(buf[k & 1], dev + (size_t)r0 * row_bytes, (size_t)(r1 - r0) * row_bytes,
                                     cudaMemcpyDeviceToHost, f->stream);
    return ce == cudaSuccess ? cudaEventRecord(ev[k & 1], f->stream) : ce;
  };
  if (e == cudaSuccess) e = copy(0);
  for (uint32_t k = 0; k < nb && e == cudaSuccess && rc == RT_OK; ++k) {
    if (k + 1 < nb) e = copy(k + 1);  // in flight while band k is consumed on the host
    if (e == cudaSuccess) e = cudaEventSynchronize(ev[k & 1]);
    if (e != cudaSuccess) break;
    const uint32_t r0 = k * band_rows, r1 = r0 + band_rows < f->h ? r0 + band_rows : f->h;
    rc = fn(user, buf[k & 1], r0, r1);
    if (rc > 0) rc = rt_fail(RT_E_IO, "band callback returned %d at rows [%u,%u)", rc, r0, r1);
  }
  cudaStreamSynchronize(f->stream);
  if (buf[0]) cudaFreeHost(buf[0]);
  if (buf[1]) cudaFreeHost(buf[1]);
  if (ev[0]) cudaEventDestroy(ev[0]);
  if (ev[1]) cudaEventDestroy(ev[1]);
  if (e != cudaSuccess) return rt_fail(RT_E_CUDA, "banded read-back failed: %s", cudaGetErrorString(e));
  if (rc == RT_OK && stats) rc = rt_frame_stats(f, stats);
  return rc;
}

extern "C" int rt_finalize_to_file(rt_frame *f, const rt_config *cfg, const void *const *partials,
                                   uint32_t count, const char *path, int format, rt_stats *stats) {
  if (!f || !path) return rt_fail(RT_E_INVALID, "NULL argument");
  PnmFile out;
  int rc = pnm_open(out, path, f->w, f->h, format);
  if (rc) return rc;
  auto band = [](void *user, const uint8_t *rows, uint32_t r0, uint32_t r1) -> int {
    return pnm_rows(*static_cast<PnmFile *>(user), rows, r1 - r0);
  };
  rc = rt_finalize_bands(f, cfg, partials, count, 64, band, &out, stats);
  return pnm_close(out, rc);
}
