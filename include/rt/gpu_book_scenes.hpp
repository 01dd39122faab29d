// include/rt/gpu_book_scenes.hpp — the intermediate steps of the reference's CUDA program as
// archive-gpu/ keeps them (image7 ... image23), through the rt::gpu API.
// Every step is the world its new_world kernel builds (insertion order) and the literals of its
// new_camera; render() runs them with RT_FLAG_V4_MODEL like the final program.  The diffuse-only
// steps (image7-12: no material classes yet) use the rules their ray_color spells out:
//   image7-9   direction = random_on_hemisphere(normal)      archive-gpu/image7/camera.h:79
//              (a unit vector flipped into the normal's hemisphere: the same direction distribution
//              as archive/listing40's random_in_hemisphere -> RT_DIFFUSE_HEMI)
//   image10-12 direction = normal + random_unit_vector()     archive-gpu/image10/camera.h:80 -> RT_LAMBERTIAN
// with attenuation 0.5 per bounce; image7-11 write colours without gamma (RT_FLAG_LINEAR_OUTPUT),
// image7 stops at 10 bounces.  Not reproduced: the t_min = 0 of image7/8 (shadow acne).
#ifndef RT_GPU_BOOK_SCENES_HPP
#define RT_GPU_BOOK_SCENES_HPP

#include <string>
#include <vector>

#include "book_scenes.hpp"
#include "gpu_api.hpp"

namespace rt {
namespace gpu {
namespace book {

struct step {
  std::string name, listing;
  void (*build)(hittable_list *world, material_list *materials);
  void (*setup)(camera *cam);
  uint32_t flags;  // besides RT_FLAG_V4_MODEL
};

namespace detail {
// camera.h of image7...18: centre at the origin, focal length 1, viewport height 2 (vfov 90 along -z)
inline void early_camera(camera *cam, int max_depth) {
  cam->max_depth = max_depth;
  cam->samples_per_pixel = 10;
  cam->aspect_ratio = 16.0f / 9.0f;
  cam->set_image_width(1920);
  cam->vfov = 90.0f;
  cam->lookfrom = point3(0.0, 0.0, 0.0);
  cam->lookat = point3(0.0, 0.0, -1.0);
  cam->vup = vec3(0.0, 1.0, 0.0);
  cam->defocus_angle = 0.0f;
  cam->focus_dist = 1.0f;
}
inline void positioned(camera *cam, float vfov, float defocus_angle, float focus_dist) {
  early_camera(cam, 50);
  cam->vfov = vfov;
  cam->lookfrom = point3(-2.0, 2.0, 1.0);
  cam->lookat = point3(0.0, 0.0, -1.0);
  cam->defocus_angle = defocus_angle;
  cam->focus_dist = focus_dist;
}
inline void two_spheres(hittable_list *world, material_list *materials, rt::book::diffuse::rule_t rule) {
  material *m = new rt::book::diffuse(rule);  // archive-gpu/image7/main.cu:23-24
  materials->add(m);
  world->add(new sphere(point3(0.0, 0.0, -1.0), 0.5f, m));
  world->add(new sphere(point3(0.0, -100.5, -1.0), 100.0f, m));
}
// ground / centre / left (/ bubble) / right, archive-gpu/image13/main.cu:22-32
inline void trio(hittable_list *world, material_list *materials, material *ground, material *center,
                 material *left, material *right, float bubble, material *bubble_mat) {
  for (material *m : {ground, center, left, right}) materials->add(m);
  if (bubble_mat && bubble_mat != left) materials->add(bubble_mat);
  world->add(new sphere(point3(0.0, -100.5, -1.0), 100.0f, ground));
  world->add(new sphere(point3(0.0, 0.0, -1.0), 0.5f, center));
  world->add(new sphere(point3(-1.0, 0.0, -1.0), 0.5f, left));
  if (bubble != 0.0f) world->add(new sphere(point3(-1.0, 0.0, -1.0), bubble, bubble_mat ? bubble_mat : left));
  world->add(new sphere(point3(1.0, 0.0, -1.0), 0.5f, right));
}
inline void glass_trio(hittable_list *w, material_list *m, float bubble, bool own_bubble_material) {
  trio(w, m, new lambertian(color(0.8f, 0.8f, 0.0f)), new lambertian(color(0.1f, 0.2f, 0.5f)), new dielectric(1.5f),
       new metal(color(0.8f, 0.6f, 0.2f), 0.0f), bubble, own_bubble_material ? new dielectric(1.5f) : nullptr);
}
}  // namespace detail

inline std::vector<step> all() {
  using namespace detail;
  typedef rt::book::diffuse D;
  return {
      {"gpu_hemisphere_depth10", "archive-gpu/image7/main.cu:23-24 + camera.h:74-79",
       [](hittable_list *w, material_list *m) { two_spheres(w, m, D::in_hemisphere); },
       [](camera *c) { early_camera(c, 10); }, RT_FLAG_LINEAR_OUTPUT},
      {"gpu_hemisphere", "archive-gpu/image9/main.cu:23-24 + camera.h:79-80",
       [](hittable_list *w, material_list *m) { two_spheres(w, m, D::in_hemisphere); },
       [](camera *c) { early_camera(c, 50); }, RT_FLAG_LINEAR_OUTPUT},
      {"gpu_lambertian", "archive-gpu/image10/main.cu:23-24 + camera.h:80",
       [](hittable_list *w, material_list *m) { two_spheres(w, m, D::unit_vector); },
       [](camera *c) { early_camera(c, 50); }, RT_FLAG_LINEAR_OUTPUT},
      {"gpu_lambertian_gamma", "archive-gpu/image12/main.cu:23-24 + color.h:12-13",
       [](hittable_list *w, material_list *m) { two_spheres(w, m, D::unit_vector); },
       [](camera *c) { early_camera(c, 50); }, 0u},
      {"gpu_shiny_metal", "archive-gpu/image13/main.cu:22-32",
       [](hittable_list *w, material_list *m) {
         trio(w, m, new lambertian(color(0.8f, 0.8f, 0.0f)), new lambertian(color(0.7f, 0.3f, 0.3f)),
              new metal(color(0.8f, 0.8f, 0.8f), 0.0f), new metal(color(0.8f, 0.6f, 0.2f), 0.0f), 0.0f, nullptr);
       },
       [](camera *c) { early_camera(c, 50); }, 0u},
      {"gpu_fuzzed_metal", "archive-gpu/image14/main.cu:22-32",
       [](hittable_list *w, material_list *m) {
         trio(w, m, new lambertian(color(0.8f, 0.8f, 0.0f)), new lambertian(color(0.7f, 0.3f, 0.3f)),
              new metal(color(0.8f, 0.8f, 0.8f), 0.3f), new metal(color(0.8f, 0.6f, 0.2f), 1.0f), 0.0f, nullptr);
       },
       [](camera *c) { early_camera(c, 50); }, 0u},
      {"gpu_glass_pair", "archive-gpu/image16/main.cu:22-32",
       [](hittable_list *w, material_list *m) {
         trio(w, m, new lambertian(color(0.8f, 0.8f, 0.0f)), new dielectric(1.5f), new dielectric(1.5f),
              new metal(color(0.8f, 0.6f, 0.2f), 1.0f), 0.0f, nullptr);
       },
       [](camera *c) { early_camera(c, 50); }, 0u},
      {"gpu_glass_single", "archive-gpu/image17/main.cu:22-32",
       [](hittable_list *w, material_list *m) { glass_trio(w, m, 0.0f, false); },
       [](camera *c) { early_camera(c, 50); }, 0u},
      {"gpu_hollow_glass", "archive-gpu/image18/main.cu:22-34",
       [](hittable_list *w, material_list *m) { glass_trio(w, m, -0.4f, true); },
       [](camera *c) { early_camera(c, 50); }, 0u},
      {"gpu_wide_angle", "archive-gpu/image19/main.cu:22-30 + camera.h:52",
       [](hittable_list *w, material_list *m) {
         const float R = std::cos(3.1415926535897932385f / 4.0f);
         material *l = new lambertian(color(0.0f, 0.0f, 1.0f)), *r = new lambertian(color(1.0f, 0.0f, 0.0f));
         m->add(l); m->add(r);
         w->add(new sphere(point3(-R, 0.0, -1.0), R, l));
         w->add(new sphere(point3(R, 0.0, -1.0), R, r));
       },
       [](camera *c) { early_camera(c, 50); }, 0u},
      {"gpu_distant_view", "archive-gpu/image20/main.cu:22-33 + camera.h:56-58",
       [](hittable_list *w, material_list *m) { glass_trio(w, m, -0.4f, true); },
       [](camera *c) { positioned(c, 90.0f, 0.0f, 1.0f); }, 0u},
      {"gpu_zoomed_view", "archive-gpu/image21/main.cu:22-33 + camera.h:56-58",
       [](hittable_list *w, material_list *m) { glass_trio(w, m, -0.4f, true); },
       [](camera *c) { positioned(c, 20.0f, 0.0f, 1.0f); }, 0u},
      {"gpu_defocus_blur", "archive-gpu/image22/main.cu:24-38 + camera.h:65-71",
       [](hittable_list *w, material_list *m) { glass_trio(w, m, -0.4f, false); },
       [](camera *c) { positioned(c, 20.0f, 10.0f, 3.4f); }, 0u},
      {"gpu_final_cover", "archive-gpu/image23/main.cu:28-73 == src/gpu/main.cu:18-75",
       [](hittable_list *w, material_list *m) { new_world(w, m, 1984ull); },
       [](camera *c) { new_camera(c); }, 0u},
  };
}

}  // namespace book
}  // namespace gpu
}  // namespace rt
#endif
