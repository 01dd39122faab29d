// csrc/rt_primary_f64.cu — accuracy-gate kernel: primary hits in IEEE double, op-for-op.
//
// COMPILED WITH -fmad=false: every * and + below is individually rounded, like the
// reference built by g++ for baseline x86-64 (no FMA instructions).  CUDA's double
// division and sqrt are IEEE-correct, so sphere ids match the reference exactly and
// t / normals are expected to match to 0 ulp (the gate allows 2).
//
// Restates, per pixel: the pixel-centre ray of src/cpu/main.cc:116-118 with xi = 0.5 and a
// zero lens offset (camera::get_ray, src/cpu/camera.h:28-35), hittable_list::hit
// (src/cpu/hittable_list.h:28-43), sphere::hit (src/cpu/sphere.h:24-51) and
// hit_record::set_face_normal (src/cpu/hittable.h:16-19).
#include "rt_kernels.h"

namespace rtb {

__global__ void __launch_bounds__(128)
primary_f64_kernel(const double4 *__restrict__ exact, uint32_t n,
                   const __grid_constant__ CameraF64 cam, uint32_t W, uint32_t H, int32_t *ids,
                   double *tt, double *normal, uint8_t *ffo) {
  uint32_t px = blockIdx.x * blockDim.x + threadIdx.x;
  if (px >= W * H) return;
  uint32_t row = px / W, i = px - row * W, j = H - 1 - row;
  double u = ((double)i + 0.5) / (double)(W - 1);
  double v = ((double)j + 0.5) / (double)(H - 1);
  // origin + offset, offset = 0
  double ox = cam.origin[0] + 0.0, oy = cam.origin[1] + 0.0, oz = cam.origin[2] + 0.0;
  // lower_left_corner + s*horizontal + t*vertical - origin - offset
  double dx = cam.lower_left[0] + u * cam.horizontal[0] + v * cam.vertical[0] - cam.origin[0] - 0.0;
  double dy = cam.lower_left[1] + u * cam.horizontal[1] + v * cam.vertical[1] - cam.origin[1] - 0.0;
  double dz = cam.lower_left[2] + u * cam.horizontal[2] + v * cam.vertical[2] - cam.origin[2] - 0.0;

  const double t_min = 0.001;
  double closest = __longlong_as_double(0x7ff0000000000000ll);
  int id = -1;
  double a = dx * dx + dy * dy + dz * dz;  // direction().length_squared()
  for (uint32_t s = 0; s < n; ++s) {
    double4 S = exact[s];
    double ocx = ox - S.x, ocy = oy - S.y, ocz = oz - S.z;
    double half_b = ocx * dx + ocy * dy + ocz * dz;
    double c = ocx * ocx + ocy * ocy + ocz * ocz - S.w * S.w;
    double disc = half_b * half_b - a * c;
    if (disc < 0) continue;
    double sq = sqrt(disc);
    double root = (-half_b - sq) / a;
    if (root < t_min || closest < root) {
      root = (-half_b + sq) / a;
      if (root < t_min || closest < root) continue;
    }
    closest = root;
    id = (int)s;
  }
  ids[px] = id;
  if (id < 0) {
    tt[px] = 0; ffo[px] = 0;
    normal[(size_t)px * 3] = normal[(size_t)px * 3 + 1] = normal[(size_t)px * 3 + 2] = 0;
    return;
  }
  double4 S = exact[id];
  double t = closest;
  double qx = ox + t * dx, qy = oy + t * dy, qz = oz + t * dz;  // ray::at
  double ir = 1 / S.w;                                           // vec3 operator/ : (1/t) * v
  double onx = ir * (qx - S.x), ony = ir * (qy - S.y), onz = ir * (qz - S.z);
  bool front = (dx * onx + dy * ony + dz * onz) < 0;
  tt[px] = t;
  ffo[px] = front;
  normal[(size_t)px * 3 + 0] = front ? onx : -onx;
  normal[(size_t)px * 3 + 1] = front ? ony : -ony;
  normal[(size_t)px * 3 + 2] = front ? onz : -onz;
}

cudaError_t launch_primary_f64(const double4 *exact, uint32_t n, const CameraF64 &cam, uint32_t W,
                               uint32_t H, int32_t *ids, double *t, double *normal, uint8_t *ff,
                               cudaStream_t st) {
  uint32_t npix = W * H;
  primary_f64_kernel<<<(npix + 127) / 128, 128, 0, st>>>(exact, n, cam, W, H, ids, t, normal, ff);
  return cudaGetLastError();
}

}  // namespace rtb
