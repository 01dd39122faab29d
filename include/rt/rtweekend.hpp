// include/rt/rtweekend.hpp — header-only C++ façade over the C ABI (include/rt_b200.h).
//
// The reference has no library boundary: its API is the set of C++ types main() uses
// (SURVEY.md §8b).  This header offers those names — vec3/point3/color, ray, camera(lookfrom,
// lookat, vup, vfov, aspect_ratio, aperture, focus_dist), material + lambertian/metal/
// dielectric, hittable + sphere + hittable_list{add, clear}, random_scene() — so that a main()
// shaped like src/cpu/main.cc:78-133 compiles against it.  Nothing here traces a ray on the
// CPU: objects only DESCRIBE the scene; hittable_list::flatten() turns them into the rt_sphere
// array of the C ABI (reference insertion order) and rt::render() runs the sm_100a kernels.
// There is no CPU fallback.
//
// Reference counterparts (paths relative to the reference root):
//   vec3, point3, color      src/cpu/vec3.h:9-103         ray            src/cpu/ray.h:6-19
//   camera                   src/cpu/camera.h:8-26        sphere         src/cpu/sphere.h:7-22
//   hittable_list            src/cpu/hittable_list.h:12-26
//   lambertian/metal/dielectric  src/cpu/material.h:14-17, 36-38, 53-55
//   random_double            src/cpu/rtweekend.h:27-36    random_scene   src/cpu/main.cc:32-76
//   write_color / P3 output  src/cpu/color.h:8-23, main.cc:109
#ifndef RT_RTWEEKEND_HPP
#define RT_RTWEEKEND_HPP

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <memory>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include "../rt_b200.h"

namespace rt {

using std::make_shared;
using std::shared_ptr;

// ------------------------------------------------------------------ vec3
class vec3 {
 public:
  double e[3];
  vec3() : e{0, 0, 0} {}
  vec3(double x, double y, double z) : e{x, y, z} {}
  double x() const { return e[0]; }
  double y() const { return e[1]; }
  double z() const { return e[2]; }
  double operator[](int i) const { return e[i]; }
  double &operator[](int i) { return e[i]; }
  vec3 operator-() const { return vec3(-e[0], -e[1], -e[2]); }
  vec3 &operator+=(const vec3 &o) { for (int k = 0; k < 3; ++k) e[k] += o.e[k]; return *this; }
  vec3 &operator*=(double t) { for (int k = 0; k < 3; ++k) e[k] *= t; return *this; }
  vec3 &operator/=(double t) { return *this *= 1 / t; }
  double length_squared() const { return e[0] * e[0] + e[1] * e[1] + e[2] * e[2]; }
  double length() const { return std::sqrt(length_squared()); }
  bool near_zero() const {
    const double s = 1e-8;
    return std::fabs(e[0]) < s && std::fabs(e[1]) < s && std::fabs(e[2]) < s;
  }
};
using point3 = vec3;
using color = vec3;

inline vec3 operator+(const vec3 &a, const vec3 &b) { return vec3(a.e[0] + b.e[0], a.e[1] + b.e[1], a.e[2] + b.e[2]); }
inline vec3 operator-(const vec3 &a, const vec3 &b) { return vec3(a.e[0] - b.e[0], a.e[1] - b.e[1], a.e[2] - b.e[2]); }
inline vec3 operator*(const vec3 &a, const vec3 &b) { return vec3(a.e[0] * b.e[0], a.e[1] * b.e[1], a.e[2] * b.e[2]); }
inline vec3 operator*(double t, const vec3 &a) { return vec3(t * a.e[0], t * a.e[1], t * a.e[2]); }
inline vec3 operator*(const vec3 &a, double t) { return t * a; }
inline vec3 operator/(const vec3 &a, double t) { return (1 / t) * a; }
inline double dot(const vec3 &a, const vec3 &b) { return a.e[0] * b.e[0] + a.e[1] * b.e[1] + a.e[2] * b.e[2]; }
inline vec3 cross(const vec3 &a, const vec3 &b) {
  return vec3(a.e[1] * b.e[2] - a.e[2] * b.e[1], a.e[2] * b.e[0] - a.e[0] * b.e[2],
              a.e[0] * b.e[1] - a.e[1] * b.e[0]);
}
inline vec3 unit_vector(const vec3 &a) { return a / a.length(); }

class ray {
 public:
  point3 orig;
  vec3 dir;
  ray() {}
  ray(const point3 &o, const vec3 &d) : orig(o), dir(d) {}
  point3 origin() const { return orig; }
  vec3 direction() const { return dir; }
  point3 at(double t) const { return orig + t * dir; }
};

// ------------------------------------------------------------------ host RNG (scene building only)
// The reference draws scene parameters from one default-seeded std::mt19937 through
// uniform_real_distribution<double>(0,1) (rtweekend.h:27-31).  mt19937 is standard-exact; the
// distribution is implementation-defined, so the libstdc++ formula is spelled out here
// (two 32-bit draws: (g1 + g2 * 2^32) / 2^64) to keep scenes identical on any toolchain.
class scene_rng {
 public:
  explicit scene_rng(uint32_t seed = 5489u) : gen_(seed) {}
  double random_double() {
    double lo = (double)gen_(), hi = (double)gen_();
    double v = (lo + hi * 4294967296.0) / 18446744073709551616.0;
    return v >= 1.0 ? std::nextafter(1.0, 0.0) : v;
  }
  double random_double(double min, double max) { return min + (max - min) * random_double(); }
  // vec3::random(): the reference writes vec3(rd(), rd(), rd()); g++ on x86-64 evaluates the
  // arguments right to left, which is the order the golden scene dump was produced with.
  vec3 random_vec3() { double z = random_double(), y = random_double(), x = random_double(); return vec3(x, y, z); }
  vec3 random_vec3(double min, double max) {
    double z = random_double(min, max), y = random_double(min, max), x = random_double(min, max);
    return vec3(x, y, z);
  }
 private:
  std::mt19937 gen_;
};

// ------------------------------------------------------------------ materials
class material {
 public:
  virtual ~material() {}
  virtual void describe(rt_sphere &out) const = 0;  // fills kind / albedo / fuzz_or_ir
};

class lambertian : public material {
 public:
  color albedo;
  explicit lambertian(const color &a) : albedo(a) {}
  void describe(rt_sphere &o) const override {
    o.kind = RT_LAMBERTIAN;
    for (int k = 0; k < 3; ++k) o.albedo[k] = albedo[k];
    o.fuzz_or_ir = 0;
  }
};

class metal : public material {
 public:
  color albedo;
  double fuzz;
  metal(const color &a, double f) : albedo(a), fuzz(f < 1 ? f : 1) {}
  void describe(rt_sphere &o) const override {
    o.kind = RT_METAL;
    for (int k = 0; k < 3; ++k) o.albedo[k] = albedo[k];
    o.fuzz_or_ir = fuzz;
  }
};

class dielectric : public material {
 public:
  double ir;
  explicit dielectric(double index_of_refraction) : ir(index_of_refraction) {}
  void describe(rt_sphere &o) const override {
    o.kind = RT_DIELECTRIC;
    o.albedo[0] = o.albedo[1] = o.albedo[2] = 1.0;
    o.fuzz_or_ir = ir;
  }
};

// ------------------------------------------------------------------ hittables
class hittable {
 public:
  virtual ~hittable() {}
  virtual void flatten(std::vector<rt_sphere> &out) const = 0;
};

class sphere : public hittable {
 public:
  point3 center;
  double radius;  // signed: a negative radius is a hollow bubble (normals point inward)
  shared_ptr<material> mat_ptr;
  sphere() : radius(0) {}
  sphere(point3 cen, double r, shared_ptr<material> m) : center(cen), radius(r), mat_ptr(std::move(m)) {}
  void flatten(std::vector<rt_sphere> &out) const override {
    rt_sphere s{};
    for (int k = 0; k < 3; ++k) s.center[k] = center[k];
    s.radius = radius;
    if (!mat_ptr) throw std::invalid_argument("sphere without material");
    mat_ptr->describe(s);
    out.push_back(s);
  }
};

class hittable_list : public hittable {
 public:
  std::vector<shared_ptr<hittable>> objects;
  hittable_list() {}
  explicit hittable_list(shared_ptr<hittable> object) { add(std::move(object)); }
  void clear() { objects.clear(); }
  void add(shared_ptr<hittable> object) { objects.push_back(std::move(object)); }
  // insertion order is the closest-hit tie order of hittable_list::hit
  void flatten(std::vector<rt_sphere> &out) const override {
    for (const auto &o : objects) o->flatten(out);
  }
  std::vector<rt_sphere> flatten() const {
    std::vector<rt_sphere> v;
    flatten(v);
    return v;
  }
};

// ------------------------------------------------------------------ camera
class camera {
 public:
  camera(point3 lookfrom, point3 lookat, vec3 vup, double vfov, double aspect_ratio,
         double aperture, double focus_dist) {
    for (int k = 0; k < 3; ++k) {
      c_.lookfrom[k] = lookfrom[k]; c_.lookat[k] = lookat[k]; c_.vup[k] = vup[k];
    }
    c_.vfov_deg = vfov; c_.aspect_ratio = aspect_ratio; c_.aperture = aperture;
    c_.focus_dist = focus_dist;
  }
  const rt_camera &abi() const { return c_; }
 private:
  rt_camera c_;
};

// ------------------------------------------------------------------ scenes
// Book-cover scene: ground + 22x22 grid of small spheres (kept when farther than 0.9 from
// (4, 0.2, 0)) + three big ones.  Draw order per grid cell: material choice, centre z-jitter,
// centre x-jitter, then material parameters (lambertian: right factor before left factor).
// `half` = half-width of the grid of small spheres: 11 is the book cover (22 x 22 cells, 486
// spheres with the default seed); 158 gives the ~100 k-sphere stress scene of BASELINE.json
// configs[3] (same recipe, same draw order, larger grid).
inline hittable_list random_scene(scene_rng &rng, int half = 11) {
  hittable_list world;
  world.add(make_shared<sphere>(point3(0, -1000, 0), 1000, make_shared<lambertian>(color(0.5, 0.5, 0.5))));
  const point3 keep_clear(4, 0.2, 0);
  for (int a = -half; a < half; ++a)
    for (int b = -half; b < half; ++b) {
      const double pick = rng.random_double();
      const double jz = rng.random_double(), jx = rng.random_double();
      const point3 c(a + 0.9 * jx, 0.2, b + 0.9 * jz);
      if (!((c - keep_clear).length() > 0.9)) continue;
      shared_ptr<material> m;
      if (pick < 0.8) {
        const color right = rng.random_vec3(), left = rng.random_vec3();
        m = make_shared<lambertian>(left * right);
      } else if (pick < 0.95) {
        const color albedo = rng.random_vec3(0.5, 1);
        m = make_shared<metal>(albedo, rng.random_double(0, 0.5));
      } else {
        m = make_shared<dielectric>(1.5);
      }
      world.add(make_shared<sphere>(c, 0.2, m));
    }
  world.add(make_shared<sphere>(point3(0, 1, 0), 1.0, make_shared<dielectric>(1.5)));
  world.add(make_shared<sphere>(point3(-4, 1, 0), 1.0, make_shared<lambertian>(color(0.4, 0.2, 0.1))));
  world.add(make_shared<sphere>(point3(4, 1, 0), 1.0, make_shared<metal>(color(0.7, 0.6, 0.5), 0.0)));
  return world;
}
inline hittable_list random_scene() {
  scene_rng rng;
  return random_scene(rng);
}
inline hittable_list stress_scene(int half = 158, uint32_t seed = 1u) {
  scene_rng rng(seed);
  return random_scene(rng, half);
}

// Three spheres on a ground ball with a hollow glass bubble (archive/listing69/main.cc:43-71).
inline hittable_list three_sphere_scene() {
  hittable_list world;
  auto glass = make_shared<dielectric>(1.5);
  world.add(make_shared<sphere>(point3(0.0, -100.5, -1.0), 100.0, make_shared<lambertian>(color(0.8, 0.8, 0.0))));
  world.add(make_shared<sphere>(point3(0.0, 0.0, -1.0), 0.5, make_shared<lambertian>(color(0.1, 0.2, 0.5))));
  world.add(make_shared<sphere>(point3(-1.0, 0.0, -1.0), 0.5, glass));
  world.add(make_shared<sphere>(point3(-1.0, 0.0, -1.0), -0.45, glass));
  world.add(make_shared<sphere>(point3(1.0, 0.0, -1.0), 0.5, make_shared<metal>(color(0.8, 0.6, 0.2), 0.0)));
  return world;
}

// ------------------------------------------------------------------ rendering through the C ABI
class error : public std::runtime_error {
 public:
  error(int code, const std::string &what) : std::runtime_error(what), code(code) {}
  int code;
};
inline void check(int rc) {
  if (rc != RT_OK) throw error(rc, std::string("rt_b200 error ") + std::to_string(rc) + ": " + rt_last_error());
}

struct image {
  uint32_t width = 0, height = 0;
  std::vector<uint8_t> rgb;  // [H][W][3], rows top -> bottom, write_color's 0..255 values
  rt_stats stats{};
  void write_ppm(const char *path /* "-" = stdout */) const { check(rt_write_ppm(path, rgb.data(), width, height)); }
};

// The pixel/sample loops of main() (src/cpu/main.cc:111-123) as one call.
inline image render(const hittable_list &world, const camera &cam, uint32_t image_width,
                    uint32_t image_height, uint32_t samples_per_pixel, uint32_t max_depth,
                    uint64_t seed = 0, int device = 0, bool use_bvh = false, uint32_t flags = 0) {
  std::vector<rt_sphere> flat = world.flatten();
  rt_scene *sc = nullptr;
  check(rt_scene_create(flat.data(), (uint32_t)flat.size(), device, &sc));
  rt_config cfg{};
  cfg.width = image_width; cfg.height = image_height;
  cfg.samples_per_pixel = samples_per_pixel; cfg.max_depth = max_depth;
  cfg.seed = seed; cfg.slice_begin = 0; cfg.slice_end = RT_SLICES; cfg.device = device;
  cfg.flags = (use_bvh ? RT_FLAG_BVH : 0u) | flags;
  image img;
  img.width = image_width; img.height = image_height;
  img.rgb.resize((size_t)image_width * image_height * 3);
  int rc = rt_render_image(sc, &cam.abi(), &cfg, img.rgb.data(), nullptr, &img.stats);
  rt_scene_destroy(sc);
  check(rc);
  return img;
}

// The same frame with its samples split over several GPUs of this process; the image is
// byte-identical to the single-GPU one.
inline image render_multi(const hittable_list &world, const camera &cam, uint32_t image_width,
                          uint32_t image_height, uint32_t samples_per_pixel, uint32_t max_depth,
                          const std::vector<int32_t> &devices, uint64_t seed = 0, bool use_bvh = false) {
  std::vector<rt_sphere> flat = world.flatten();
  rt_config cfg{};
  cfg.width = image_width; cfg.height = image_height;
  cfg.samples_per_pixel = samples_per_pixel; cfg.max_depth = max_depth;
  cfg.seed = seed; cfg.slice_begin = 0; cfg.slice_end = RT_SLICES;
  cfg.flags = use_bvh ? RT_FLAG_BVH : 0u;
  image img;
  img.width = image_width; img.height = image_height;
  img.rgb.resize((size_t)image_width * image_height * 3);
  check(rt_render_image_multi(flat.data(), (uint32_t)flat.size(), &cam.abi(), &cfg, devices.data(),
                              (uint32_t)devices.size(), img.rgb.data(), nullptr, &img.stats));
  return img;
}

}  // namespace rt
#endif
