// include/rt/gpu_api.hpp — the API surface of the reference's CUDA program (src/gpu/, v4 book)
// over the same C ABI and the same sm_100a kernels.
//
// src/gpu/main.cu builds its world with device-side `new` in a <<<1,1>>> kernel, fills a camera
// struct in another one and launches render<<<blocks, threads>>>(frame_buffer, cam, world, seed).
// The names below are those, as host functions: a main() shaped like src/gpu/main.cu:86-155
// compiles against this header (examples/gpu_cover.cc).  Objects only DESCRIBE the scene;
// render() flattens them (insertion order) and runs the fused path kernel with
// RT_FLAG_V4_MODEL, i.e. with the v4 rules that differ from src/cpu/ (SURVEY.md §2.3):
// pixel-square sampling (camera.h:140-157), defocus_angle lens (camera.h:106-109), metal fuzz
// on the unit sphere (material.h:52), strict root interval (sphere.h:31-36).  There is no CPU
// fallback.
//
// Reference counterparts:
//   camera fields + new_camera    src/gpu/camera.h:14-110      render        src/gpu/camera.h:169-195
//   output_image / write_color    src/gpu/camera.h:197-210, color.h:17-38
//   sphere(center, radius, mat*)  src/gpu/sphere.h:12-13       hittable_list src/gpu/hittable_list.h:8-47
//   material_list                 src/gpu/material_list.h:6-45 new_world     src/gpu/main.cu:18-75
//   lambertian/metal/dielectric   src/gpu/material.h:20-22, 42-44, 68
//   checkCudaErrors -> exit(99)   src/gpu/cuda_utility.h:8-18  (here: rt::gpu::check_rt)
#ifndef RT_GPU_API_HPP
#define RT_GPU_API_HPP

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "rtweekend.hpp"

namespace rt {
namespace gpu {

using rt::color;
using rt::point3;
using rt::vec3;
using rt::material;
using rt::lambertian;
using rt::metal;
using rt::dielectric;
typedef material *material_ptr;

// errors end the program like checkCudaErrors (src/gpu/cuda_utility.h:8-18): message, exit(99)
inline void check_rt(int rc, const char *expr, const char *file, int line) {
  if (rc != RT_OK) {
    std::fprintf(stderr, "rt_b200 error = %d at %s:%d '%s': %s\n", rc, file, line, expr, rt_last_error());
    std::exit(99);
  }
}
#define RT_GPU_CHECK(x) ::rt::gpu::check_rt((x), #x, __FILE__, __LINE__)

class hittable {
 public:
  virtual ~hittable() {}
  virtual void flatten(std::vector<rt_sphere> &out) const = 0;
};
typedef hittable *hittable_ptr;

// sphere(center, radius, material*): the material is NOT owned (src/gpu/sphere.h:12-13)
class sphere : public hittable {
 public:
  sphere(point3 _center, float _radius, material *_material) : center(_center), radius(_radius), mat(_material) {}
  void flatten(std::vector<rt_sphere> &out) const override {
    rt_sphere s{};
    for (int k = 0; k < 3; ++k) s.center[k] = center[k];
    s.radius = radius;
    mat->describe(s);
    out.push_back(s);
  }
  point3 center;
  float radius;
  material *mat;
};

// owns its objects like the reference's list (deleted in the destructor, hittable_list.h:38-47)
class hittable_list : public hittable {
 public:
  std::vector<hittable_ptr> objects;
  hittable_list() {}
  hittable_list(const hittable_list &) = delete;
  hittable_list &operator=(const hittable_list &) = delete;
  ~hittable_list() override { for (hittable_ptr o : objects) delete o; }
  size_t objects_size() const { return objects.size(); }
  void add(hittable_ptr object) { objects.push_back(object); }
  void flatten(std::vector<rt_sphere> &out) const override { for (hittable_ptr o : objects) o->flatten(out); }
};

class material_list {
 public:
  std::vector<material_ptr> materials;
  material_list() {}
  material_list(const material_list &) = delete;
  material_list &operator=(const material_list &) = delete;
  ~material_list() { for (material_ptr m : materials) delete m; }
  void add(material_ptr mat) { materials.push_back(mat); }
};

// the public fields of the reference's camera (src/gpu/camera.h:14-31); the derived fields
// (pixel00_loc, pixel_delta_u/v, defocus_disk_u/v) are computed inside the library in f64
struct camera {
  int max_depth, samples_per_pixel;
  float aspect_ratio;
  int image_width, image_height;
  float vfov;
  point3 lookfrom, lookat;
  vec3 vup;
  float defocus_angle, focus_dist;

  // image_height as new_camera derives it (camera.h:62-63)
  void set_image_width(int w) {
    image_width = w;
    image_height = static_cast<int>(image_width / aspect_ratio);
    image_height = image_height < 1 ? 1 : image_height;
  }
  rt_camera abi() const {
    rt_camera c{};
    for (int k = 0; k < 3; ++k) { c.lookfrom[k] = lookfrom[k]; c.lookat[k] = lookat[k]; c.vup[k] = vup[k]; }
    c.vfov_deg = vfov;
    c.aspect_ratio = (double)image_width / image_height;  // viewport_width, camera.h:79-80
    const double pi = 3.1415926535897932385;
    // defocus_radius = focus_dist * tan(radians(defocus_angle / 2)), camera.h:106-107; a
    // non-positive angle is the pinhole (camera.h:162-164)
    c.aperture = defocus_angle <= 0.0f ? 0.0 : 2.0 * focus_dist * std::tan((double)defocus_angle / 2.0 * pi / 180.0);
    c.focus_dist = focus_dist;
    return c;
  }
};

// the literals of new_camera, src/gpu/camera.h:58-70
inline void new_camera(camera *cam) {
  cam->max_depth = 50;
  cam->samples_per_pixel = 500;
  cam->aspect_ratio = 16.0f / 9.0f;
  cam->set_image_width(1920);
  cam->vfov = 20.0f;
  cam->lookfrom = point3(13.0, 2.0, 3.0);
  cam->lookat = point3(0.0, 0.0, 0.0);
  cam->vup = vec3(0.0, 1.0, 0.0);
  cam->defocus_angle = 0.6f;
  cam->focus_dist = 10.0f;
}

// new_world, src/gpu/main.cu:18-75, with the draws of one host stream seeded by `seed` (the
// reference draws from curand(seed, 0, 0); its world is never reproducible, main.cu:88).  Order
// of draws: material choice, centre x, centre z, then the material parameters.  Every material
// is handed to `materials` (the reference leaks the small spheres' ones).
inline void new_world(hittable_list *world, material_list *materials, unsigned long long seed) {
  scene_rng rng((uint32_t)(seed ^ (seed >> 32)));
  auto rf = [&]() { return (float)rng.random_double(); };
  auto rf2 = [&](float lo, float hi) { return lo + (hi - lo) * rf(); };
  material *ground_material = new lambertian(color(0.5, 0.5, 0.5));
  materials->add(ground_material);
  world->add(new sphere(point3(0.0, -1000.0, 0.0), 1000.0f, ground_material));
  for (int a = -11; a < 11; a++) {
    for (int b = -11; b < 11; b++) {
      float choose_mat = rf();
      float cx = a + 0.9f * rf();
      float cz = b + 0.9f * rf();
      point3 center(cx, 0.2f, cz);
      if ((center - point3(4.0, 0.2f, 0.0)).length() > 0.9) {
        material *sphere_material;
        if (choose_mat < 0.8f) {
          float x0 = rf(), y0 = rf(), z0 = rf(), x1 = rf(), y1 = rf(), z1 = rf();
          sphere_material = new lambertian(color(x0 * x1, y0 * y1, z0 * z1));
        } else if (choose_mat < 0.95f) {
          float x = rf2(0.5f, 1.0f), y = rf2(0.5f, 1.0f), z = rf2(0.5f, 1.0f);
          float fuzz = rf2(0.0f, 0.5f);
          sphere_material = new metal(color(x, y, z), fuzz);
        } else {
          sphere_material = new dielectric(1.5f);
        }
        materials->add(sphere_material);
        world->add(new sphere(center, 0.2f, sphere_material));
      }
    }
  }
  material *material1 = new dielectric(1.5f);
  materials->add(material1);
  world->add(new sphere(point3(0.0, 1.0, 0.0), 1.0f, material1));
  material *material2 = new lambertian(color(0.4f, 0.2f, 0.1f));
  materials->add(material2);
  world->add(new sphere(point3(-4.0, 1.0, 0.0), 1.0f, material2));
  material *material3 = new metal(color(0.7f, 0.6f, 0.5f), 0.0f);
  materials->add(material3);
  world->add(new sphere(point3(4.0, 1.0, 0.0), 1.0f, material3));
}

// render<<<blocks, threads>>>(frame_buffer, cam, world, seed), src/gpu/camera.h:169-195:
// frame_buffer[y * image_width + x] = SUM of the pixel's samples (linear, y from the top).
// `flags` may add RT_FLAG_BVH or RT_FLAG_F32_HIT; stats is optional.
inline void render(color *frame_buffer, const camera *cam, const hittable *world,
                   unsigned long long seed, int device = 0, uint32_t flags = 0, rt_stats *stats = nullptr) {
  std::vector<rt_sphere> flat;
  world->flatten(flat);
  rt_scene *sc = nullptr;
  RT_GPU_CHECK(rt_scene_create(flat.data(), (uint32_t)flat.size(), device, &sc));
  rt_config cfg{};
  cfg.width = (uint32_t)cam->image_width; cfg.height = (uint32_t)cam->image_height;
  cfg.samples_per_pixel = (uint32_t)cam->samples_per_pixel; cfg.max_depth = (uint32_t)cam->max_depth;
  cfg.seed = seed; cfg.slice_begin = 0; cfg.slice_end = RT_SLICES; cfg.device = device;
  cfg.flags = RT_FLAG_V4_MODEL | flags;
  const size_t npix = (size_t)cfg.width * cfg.height;
  std::vector<uint8_t> rgb(npix * 3);
  std::vector<float> lin(npix * 3);
  const rt_camera c = cam->abi();
  RT_GPU_CHECK(rt_render_image(sc, &c, &cfg, rgb.data(), lin.data(), stats));
  RT_GPU_CHECK(rt_scene_destroy(sc));
  for (size_t p = 0; p < npix; ++p) frame_buffer[p] = color(lin[p * 3], lin[p * 3 + 1], lin[p * 3 + 2]);
}

// output_image + write_color, src/gpu/camera.h:197-210 and color.h:17-38 (f32: scale, sqrt,
// clamp to [0, 0.999], 256x), P3 text on stdout — through the library's table-driven writer
// (gamma = false: the colour write of archive-gpu/image7...11, before linear_to_gamma)
inline void output_image(const camera *cam, const color *frame_buffer, const char *path = "-", bool gamma = true) {
  const size_t npix = (size_t)cam->image_width * cam->image_height;
  std::vector<uint8_t> rgb(npix * 3);
  const float scale = 1.0f / cam->samples_per_pixel;
  for (size_t p = 0; p < npix; ++p)
    for (int k = 0; k < 3; ++k) {
      float x = gamma ? std::sqrt((float)frame_buffer[p][k] * scale) : (float)frame_buffer[p][k] * scale;
      x = x < 0.000f ? 0.000f : (0.999f < x ? 0.999f : x);
      rgb[p * 3 + k] = (uint8_t)static_cast<int>(256.0f * x);
    }
  RT_GPU_CHECK(rt_write_ppm(path, rgb.data(), (uint32_t)cam->image_width, (uint32_t)cam->image_height));
}

}  // namespace gpu
}  // namespace rt
#endif
