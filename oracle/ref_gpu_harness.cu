// oracle/ref_gpu_harness.cu — TEST INFRASTRUCTURE (a checker), never shipped, never on the product
// path.  Drives the reference's OWN CUDA program (src/gpu/main.cu and the headers it includes,
// compiled unmodified from where they lie) with a fixed seed instead of time(nullptr)
// (main.cu:88), dumps the world its new_world kernel built (main.cu:18-75) and writes the linear
// per-pixel sums its render kernel produced (camera.h:169-195).  tests/test_v4_model.py renders
// the dumped world through RT_FLAG_V4_MODEL and compares the two images statistically — the two
// programs cannot share a random stream (curand XORWOW per pixel there, Philox per sample here).
//
//   ref_gpu_harness <out-prefix> [spp] [seed]
//     <prefix>.scene.f64   N x 9 doubles  cx cy cz r kind r g b fuzz|ir   (the oracle's row layout)
//     <prefix>.linear.f32  H x W x 3 floats, rows top->bottom, SUMS over spp samples
//     <prefix>.meta.txt    W H spp seed N + the camera literals of new_camera (camera.h:53-70)
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <iostream>
#include <vector>
#include <cuda_runtime.h>
#include <curand_kernel.h>

#define main ref_main      /* keep the reference's main() out of the way */
#define private public     /* sphere::center/radius/mat, material parameters */
#include "src/gpu/main.cu" /* -I$(RT_REFERENCE_DIR) */
#undef private
#undef main

// Materials are polymorphic device objects; their dynamic type is recovered from the vtable
// pointer of objects whose kind is known from new_world: object 0 is the lambertian ground, the
// last three are dielectric, lambertian, metal (main.cu:28-31, 63-74).
__global__ void dump_world(hittable_list *world, double *rows, int cap, int *n_out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const int n = (int)world->objects_size;
  *n_out = n;
  if (n > cap || n < 4) return;
  auto vptr = [](const material *m) { return *reinterpret_cast<const void *const *>(m); };
  auto sph = [&](int i) { return static_cast<sphere *>(world->objects[i]); };
  const void *v_lamb = vptr(sph(0)->mat), *v_diel = vptr(sph(n - 3)->mat), *v_metal = vptr(sph(n - 1)->mat);
  for (int i = 0; i < n; ++i) {
    const sphere *s = sph(i);
    double *r = rows + (size_t)i * 9;
    r[0] = s->center.x(); r[1] = s->center.y(); r[2] = s->center.z(); r[3] = s->radius;
    const void *v = vptr(s->mat);
    if (v == v_lamb) {
      const lambertian *m = static_cast<const lambertian *>(s->mat);
      r[4] = 0; r[5] = m->albedo.x(); r[6] = m->albedo.y(); r[7] = m->albedo.z(); r[8] = 0;
    } else if (v == v_metal) {
      const metal *m = static_cast<const metal *>(s->mat);
      r[4] = 1; r[5] = m->albedo.x(); r[6] = m->albedo.y(); r[7] = m->albedo.z(); r[8] = m->fuzz;
    } else if (v == v_diel) {
      const dielectric *m = static_cast<const dielectric *>(s->mat);
      r[4] = 2; r[5] = 1; r[6] = 1; r[7] = 1; r[8] = m->ir;
    } else {
      r[4] = -1;
    }
  }
}

int main(int argc, char **argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s <out-prefix> [spp] [seed]\n", argv[0]); return 2; }
  const std::string prefix = argv[1];
  const int spp = argc > 2 ? atoi(argv[2]) : 500;
  const unsigned long long seed = argc > 3 ? strtoull(argv[3], nullptr, 10) : 20240611ull;

  // exactly the call sequence of the reference's main() (main.cu:86-153), minus the PPM text
  hittable_list *world;
  checkCudaErrors(cudaMalloc(&world, sizeof(*world)));
  material_list *materials;
  checkCudaErrors(cudaMalloc(&materials, sizeof(*materials)));
  new_world<<<1, 1>>>(world, materials, seed);
  checkCudaErrors(cudaGetLastError());
  checkCudaErrors(cudaDeviceSynchronize());

  camera *cam;
  checkCudaErrors(cudaMallocManaged(&cam, sizeof(*cam)));
  new_camera<<<1, 1>>>(cam);
  checkCudaErrors(cudaGetLastError());
  checkCudaErrors(cudaDeviceSynchronize());
  cam->samples_per_pixel = spp;  // not used by any derived camera field (camera.h:58-110)

  const int cap = 1024;
  double *rows; int *n_dev;
  checkCudaErrors(cudaMallocManaged(&rows, sizeof(double) * 9 * cap));
  checkCudaErrors(cudaMallocManaged(&n_dev, sizeof(int)));
  dump_world<<<1, 1>>>(world, rows, cap, n_dev);
  checkCudaErrors(cudaGetLastError());
  checkCudaErrors(cudaDeviceSynchronize());
  const int n = *n_dev;
  if (n > cap) { fprintf(stderr, "world too large: %d\n", n); return 3; }
  for (int i = 0; i < n; ++i) if (rows[i * 9 + 4] < 0) { fprintf(stderr, "unknown material at %d\n", i); return 3; }

  color *frame_buffer;
  const int W = cam->image_width, H = cam->image_height;
  checkCudaErrors(cudaMallocManaged(&frame_buffer, (size_t)W * H * sizeof(*frame_buffer)));
  const int tx = 16, ty = 16;
  dim3 blocks((W + tx - 1) / tx, (H + ty - 1) / ty), threads(tx, ty);
  auto t0 = std::chrono::high_resolution_clock::now();
  render<<<blocks, threads>>>(frame_buffer, cam, world, seed);
  checkCudaErrors(cudaGetLastError());
  checkCudaErrors(cudaDeviceSynchronize());
  auto t1 = std::chrono::high_resolution_clock::now();
  const double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();

  FILE *f = fopen((prefix + ".scene.f64").c_str(), "wb");
  if (!f) { perror("scene"); return 4; }
  fwrite(rows, sizeof(double), (size_t)n * 9, f); fclose(f);
  std::vector<float> lin((size_t)W * H * 3);
  for (size_t p = 0; p < (size_t)W * H; ++p) {
    lin[p * 3 + 0] = frame_buffer[p].x(); lin[p * 3 + 1] = frame_buffer[p].y(); lin[p * 3 + 2] = frame_buffer[p].z();
  }
  f = fopen((prefix + ".linear.f32").c_str(), "wb");
  if (!f) { perror("linear"); return 4; }
  fwrite(lin.data(), sizeof(float), lin.size(), f); fclose(f);
  f = fopen((prefix + ".meta.txt").c_str(), "w");
  if (!f) { perror("meta"); return 4; }
  fprintf(f, "W %d\nH %d\nspp %d\nseed %llu\nN %d\nmax_depth %d\nvfov %.9g\nlookfrom %.9g %.9g %.9g\n"
             "lookat %.9g %.9g %.9g\nvup %.9g %.9g %.9g\ndefocus_angle %.9g\nfocus_dist %.9g\nrender_ms %.3f\n",
          W, H, spp, seed, n, cam->max_depth, cam->vfov, cam->lookfrom.x(), cam->lookfrom.y(), cam->lookfrom.z(),
          cam->lookat.x(), cam->lookat.y(), cam->lookat.z(), cam->vup.x(), cam->vup.y(), cam->vup.z(),
          cam->defocus_angle, cam->focus_dist, ms);
  fclose(f);
  fprintf(stderr, "ref_gpu_harness: %dx%d x %d spp, %d spheres, render %.1f ms\n", W, H, spp, n, ms);

  checkCudaErrors(cudaFree(frame_buffer));
  checkCudaErrors(cudaFree(cam));
  delete_world<<<1, 1>>>(world, materials);
  checkCudaErrors(cudaGetLastError());
  checkCudaErrors(cudaDeviceSynchronize());
  checkCudaErrors(cudaFree(world));
  checkCudaErrors(cudaFree(materials));
  return 0;
}
