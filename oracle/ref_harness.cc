// oracle/ref_harness.cc — TEST INFRASTRUCTURE, not product code.
//
// A parameterised re-host of the reference's own CPU path tracer.  It #includes the
// reference translation unit UNMODIFIED from $RT_REFERENCE_DIR (default /root/reference;
// passed as -I by oracle/Makefile) and drives its functions with run-time W/H/spp/scene.
// No reference source is copied into this repository: the only reference-derived lines
// here are the scene *data* of the three-sphere fixture (archive/listing69/main.cc:43-71),
// expressed through the reference's own classes.
//
// Built by oracle/Makefile into oracle/_ref/ref_harness (git-ignored, travels to the GPU
// box).  Used (a) to pin oracle/rt_oracle.c bit-for-bit, (b) to generate tests/golden/*,
// (c) as the timed CPU baseline ("kind": "reference") of bench.py.
//
// Reference entry points used, all unmodified:
//   random_scene()            src/cpu/main.cc:32-76
//   ray_color()               src/cpu/main.cc:12-30
//   camera::camera/get_ray    src/cpu/camera.h:8-35
//   hittable_list::hit        src/cpu/hittable_list.h:28-43
//   sphere::hit               src/cpu/sphere.h:24-51
//   write_color()             src/cpu/color.h:8-23
// The render loop below restates src/cpu/main.cc:111-123 (same RNG draw order).

#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <limits>
#include <memory>
#include <random>
#include <string>
#include <vector>

// Reach camera's frame vectors for the zero-lens primary-ray table.
#define private public
#define main rt_reference_main
#include "src/cpu/main.cc"
#undef main
#undef private

namespace {

struct options {
  std::string scene = "cover";
  int w = 1200, h = 800, spp = 500, depth = 50;
  int row0 = 0, row1 = -1; // output rows (0 = top), half-open
  long long skip = 0;      // extra random_double() draws discarded after scene build
  std::string dump_scene, primary, linear, ppm;
  bool timing = false;
};

// Three-sphere + ground fixture: data from archive/listing69/main.cc:43-71.
hittable_list three_sphere_scene() {
  hittable_list world;
  auto ground = make_shared<lambertian>(color(0.8, 0.8, 0.0));
  auto center = make_shared<lambertian>(color(0.1, 0.2, 0.5));
  auto left = make_shared<dielectric>(1.5);
  auto right = make_shared<metal>(color(0.8, 0.6, 0.2), 0.0);
  world.add(make_shared<sphere>(point3(0.0, -100.5, -1.0), 100.0, ground));
  world.add(make_shared<sphere>(point3(0.0, 0.0, -1.0), 0.5, center));
  world.add(make_shared<sphere>(point3(-1.0, 0.0, -1.0), 0.5, left));
  world.add(make_shared<sphere>(point3(-1.0, 0.0, -1.0), -0.45, left));
  world.add(make_shared<sphere>(point3(1.0, 0.0, -1.0), 0.5, right));
  return world;
}

// Scene dump: f64 [n][9] = cx cy cz radius(signed) kind(0 lamb,1 metal,2 glass) r g b fuzz|ir
void dump_scene(const hittable_list &world, const std::string &path) {
  FILE *f = fopen(path.c_str(), "wb");
  if (!f) { perror(path.c_str()); exit(2); }
  for (const auto &obj : world.objects) {
    const sphere *s = dynamic_cast<const sphere *>(obj.get());
    double row[9] = {s->center.x(), s->center.y(), s->center.z(), s->radius, 0, 0, 0, 0, 0};
    if (auto *l = dynamic_cast<const lambertian *>(s->mat_ptr.get())) {
      row[4] = 0; row[5] = l->albedo.x(); row[6] = l->albedo.y(); row[7] = l->albedo.z();
    } else if (auto *m = dynamic_cast<const metal *>(s->mat_ptr.get())) {
      row[4] = 1; row[5] = m->albedo.x(); row[6] = m->albedo.y(); row[7] = m->albedo.z();
      row[8] = m->fuzz;
    } else if (auto *d = dynamic_cast<const dielectric *>(s->mat_ptr.get())) {
      row[4] = 2; row[5] = row[6] = row[7] = 1.0; row[8] = d->ir;
    }
    fwrite(row, sizeof(double), 9, f);
  }
  fclose(f);
}

template <class T> void write_raw(FILE *f, const std::vector<T> &v) {
  fwrite(v.data(), sizeof(T), v.size(), f);
}

} // namespace

int main(int argc, char **argv) {
  options o;
  for (int k = 1; k < argc; ++k) {
    std::string a = argv[k];
    auto next = [&]() -> const char * {
      if (k + 1 >= argc) { fprintf(stderr, "missing value for %s\n", a.c_str()); exit(2); }
      return argv[++k];
    };
    if (a == "--scene") o.scene = next();
    else if (a == "--w") o.w = atoi(next());
    else if (a == "--h") o.h = atoi(next());
    else if (a == "--spp") o.spp = atoi(next());
    else if (a == "--depth") o.depth = atoi(next());
    else if (a == "--rows") { const char *v = next(); sscanf(v, "%d:%d", &o.row0, &o.row1); }
    else if (a == "--skip") o.skip = atoll(next());
    else if (a == "--dump-scene") o.dump_scene = next();
    else if (a == "--primary") o.primary = next();
    else if (a == "--linear") o.linear = next();
    else if (a == "--ppm") o.ppm = next();
    else if (a == "--time") o.timing = true;
    else { fprintf(stderr, "unknown option %s\n", a.c_str()); return 2; }
  }
  if (o.row1 < 0) o.row1 = o.h;
  const int W = o.w, H = o.h;
  const double aspect = double(W) / double(H);

  // World + camera, in the reference's order (scene first: it consumes the shared RNG).
  hittable_list world;
  point3 lookfrom, lookat;
  vec3 vup(0, 1, 0);
  double vfov = 20, aperture, focus;
  if (o.scene == "cover") {
    world = random_scene();
    lookfrom = point3(13, 2, 3); lookat = point3(0, 0, 0);
    aperture = 0.1; focus = 10.0; // src/cpu/main.cc:93-99
  } else if (o.scene == "three") {
    world = three_sphere_scene();
    lookfrom = point3(3, 3, 2); lookat = point3(0, 0, -1);
    aperture = 2.0; focus = (lookfrom - lookat).length();
  } else { fprintf(stderr, "unknown scene %s\n", o.scene.c_str()); return 2; }
  camera cam(lookfrom, lookat, vup, vfov, aspect, aperture, focus);

  if (!o.dump_scene.empty()) dump_scene(world, o.dump_scene);

  // Primary-hit table: pixel-centre (xi = 0.5), zero-lens rays; output rows top->bottom.
  if (!o.primary.empty()) {
    std::vector<int32_t> ids(size_t(W) * H);
    std::vector<double> ts(size_t(W) * H), ns(size_t(W) * H * 3);
    std::vector<uint8_t> ff(size_t(W) * H);
    for (int row = 0; row < H; ++row) {
      int j = H - 1 - row;
      for (int i = 0; i < W; ++i) {
        double u = (i + 0.5) / (W - 1), v = (j + 0.5) / (H - 1);
        vec3 offset(0, 0, 0);
        ray r(cam.origin + offset, cam.lower_left_corner + u * cam.horizontal +
                                       v * cam.vertical - cam.origin - offset);
        hit_record rec;
        size_t px = size_t(row) * W + i;
        if (world.hit(r, 0.001, infinity, rec)) {
          int id = -1;
          for (size_t k = 0; k < world.objects.size(); ++k) {
            hit_record tmp;
            if (world.objects[k]->hit(r, 0.001, infinity, tmp) && tmp.t == rec.t) id = int(k);
          }
          ids[px] = id; ts[px] = rec.t; ff[px] = rec.front_face;
          ns[px * 3 + 0] = rec.normal.x(); ns[px * 3 + 1] = rec.normal.y();
          ns[px * 3 + 2] = rec.normal.z();
        } else {
          ids[px] = -1; ts[px] = 0; ff[px] = 0;
          ns[px * 3 + 0] = ns[px * 3 + 1] = ns[px * 3 + 2] = 0;
        }
      }
    }
    FILE *f = fopen(o.primary.c_str(), "wb");
    if (!f) { perror(o.primary.c_str()); return 2; }
    write_raw(f, ids); write_raw(f, ts); write_raw(f, ns); write_raw(f, ff);
    fclose(f);
  }

  if (o.linear.empty() && o.ppm.empty() && !o.timing) return 0;

  for (long long k = 0; k < o.skip; ++k) (void)random_double();

  const int rows = o.row1 - o.row0;
  std::vector<double> lin(size_t(rows) * W * 3);
  std::ofstream ppm;
  if (!o.ppm.empty()) {
    ppm.open(o.ppm);
    ppm << "P3\n" << W << " " << rows << "\n255\n";
  }
  auto t0 = std::chrono::high_resolution_clock::now();
  for (int row = o.row0; row < o.row1; ++row) { // src/cpu/main.cc:111-123
    int j = H - 1 - row;
    for (int i = 0; i < W; ++i) {
      color pixel_color(0, 0, 0);
      for (int s = 0; s < o.spp; ++s) {
        auto u = (i + random_double()) / (W - 1);
        auto v = (j + random_double()) / (H - 1);
        ray r = cam.get_ray(u, v);
        pixel_color += ray_color(r, world, o.depth);
      }
      size_t px = (size_t(row - o.row0) * W + i) * 3;
      lin[px] = pixel_color.x(); lin[px + 1] = pixel_color.y(); lin[px + 2] = pixel_color.z();
      if (ppm.is_open()) write_color(ppm, pixel_color, o.spp);
    }
  }
  auto t1 = std::chrono::high_resolution_clock::now();
  if (!o.linear.empty()) {
    FILE *f = fopen(o.linear.c_str(), "wb");
    if (!f) { perror(o.linear.c_str()); return 2; }
    write_raw(f, lin);
    fclose(f);
  }
  if (o.timing) {
    double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    printf("{\"samples\": %lld, \"ms\": %.3f, \"spheres\": %zu}\n",
           (long long)rows * W * o.spp, ms, world.objects.size());
  }
  return 0;
}
