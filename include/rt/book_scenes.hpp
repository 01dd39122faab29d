// include/rt/book_scenes.hpp — the scenes and cameras of the book's intermediate steps, as the
// reference's archive/ keeps them: regression scenes for the step
// BEFORE the hot path (scene/camera construction) through the rt:: facade.  Every world below
// is the literal list of `world.add(...)` calls of the cited listing; every camera is the cited
// listing's camera expressed through the final 7-argument constructor (src/cpu/camera.h:8-10):
//   camera()                       archive/listing27/camera.h:8-18   origin 0, viewport height 2,
//                                  focal length 1  ==  (0,0,0)->(0,0,-1), vfov 90, aperture 0, focus 1
//   camera(vfov, aspect)           archive/listing62/camera.h        same with vfov
//   camera(from, at, vup, vfov, aspect)  archive/listing64/camera.h:8-25   aperture 0, focus 1
// The material steps (listing50 on) run the FINAL material rules of src/cpu/material.h (the
// archive's material.h snapshots of earlier steps — e.g. the always-refracting glass of
// listing55 — are not separate code paths).  The diffuse-only steps before them (listing33-40:
// no material classes yet, the rule is written inside ray_color) have their own rules in the
// kernels, carried by book::diffuse: RT_DIFFUSE_BALL = normal + random_in_unit_sphere()
// (listing33/main.cc:12-13), RT_LAMBERTIAN = normal + random_unit_vector() (listing38/main.cc:26),
// RT_DIFFUSE_HEMI = random_in_hemisphere(normal) (listing40/main.cc:26); listing34's write_color
// has no gamma yet (RT_FLAG_LINEAR_OUTPUT).  Not reproduced: the t_min = 0 of listing33/34
// (shadow acne, fixed by the book in listing36) and listing33's missing bounce limit.
#ifndef RT_BOOK_SCENES_HPP
#define RT_BOOK_SCENES_HPP

#include <cmath>
#include <string>
#include <vector>

#include "rtweekend.hpp"

namespace rt {
namespace book {

struct step {
  std::string name;     // short id used by examples/rt_book and the tests
  std::string listing;  // where the reference keeps it
  hittable_list world;
  camera cam;
  int image_width;      // the listing's literals
  double aspect_ratio;
  int samples_per_pixel, max_depth;
  uint32_t flags = 0;   // RT_FLAG_* the step needs (RT_FLAG_LINEAR_OUTPUT before listing35)
};

// The diffuse rule of a step that has no material classes yet: ray_color itself says
// `0.5 * ray_color(ray(rec.p, target - rec.p), ...)` with one of three targets.
class diffuse : public material {
 public:
  enum rule_t { in_unit_sphere, unit_vector, in_hemisphere };
  rule_t rule;
  double reflectance;
  diffuse(rule_t r, double refl = 0.5) : rule(r), reflectance(refl) {}
  void describe(rt_sphere &o) const override {
    o.kind = rule == in_unit_sphere ? RT_DIFFUSE_BALL : rule == unit_vector ? RT_LAMBERTIAN : RT_DIFFUSE_HEMI;
    o.albedo[0] = o.albedo[1] = o.albedo[2] = reflectance;
    o.fuzz_or_ir = 0.0;
  }
};

namespace detail {
const double kAspect = 16.0 / 9.0;
inline camera default_camera(double vfov = 90.0) {
  return camera(point3(0, 0, 0), point3(0, 0, -1), vec3(0, 1, 0), vfov, kAspect, 0.0, 1.0);
}
inline camera positioned_camera(point3 from, point3 at, double vfov) {
  return camera(from, at, vec3(0, 1, 0), vfov, kAspect, 0.0, 1.0);
}
// ground + centre + left + right, the layout every three-material step shares
inline hittable_list trio(shared_ptr<material> ground, shared_ptr<material> center,
                          shared_ptr<material> left, shared_ptr<material> right, double bubble = 0.0) {
  hittable_list world;
  world.add(make_shared<sphere>(point3(0.0, -100.5, -1.0), 100.0, ground));
  world.add(make_shared<sphere>(point3(0.0, 0.0, -1.0), 0.5, center));
  world.add(make_shared<sphere>(point3(-1.0, 0.0, -1.0), 0.5, left));
  if (bubble != 0.0) world.add(make_shared<sphere>(point3(-1.0, 0.0, -1.0), bubble, left));
  world.add(make_shared<sphere>(point3(1.0, 0.0, -1.0), 0.5, right));
  return world;
}
}  // namespace detail

// The two-sphere world of the diffuse-only steps: archive/listing33/main.cc:34-36 (small sphere
// first, then the ground), default camera, 400 px, 100 spp.
inline step diffuse_step(const char *name, const char *listing, diffuse::rule_t rule, uint32_t flags) {
  auto m = make_shared<diffuse>(rule);
  hittable_list world;
  world.add(make_shared<sphere>(point3(0, 0, -1), 0.5, m));
  world.add(make_shared<sphere>(point3(0, -100.5, -1), 100, m));
  step s{name, listing, world, detail::default_camera(), 400, detail::kAspect, 100, 50};
  s.flags = flags;
  return s;
}
// "First render of a diffuse sphere" with the bounce limit: archive/listing34/main.cc:9-24, linear output
inline step diffuse_first() {
  return diffuse_step("diffuse_first", "archive/listing34/main.cc:9-24", diffuse::in_unit_sphere, RT_FLAG_LINEAR_OUTPUT);
}
// gamma 2 (listing35/color.h) and t_min = 0.001: archive/listing36/main.cc:19-32
inline step diffuse_gamma() {
  return diffuse_step("diffuse_gamma", "archive/listing36/main.cc:19-32", diffuse::in_unit_sphere, 0);
}
// "True Lambertian": target = p + normal + random_unit_vector(), archive/listing38/main.cc:26
inline step diffuse_lambertian() {
  return diffuse_step("diffuse_lambertian", "archive/listing38/main.cc:26", diffuse::unit_vector, 0);
}
// hemispherical scattering: target = p + random_in_hemisphere(normal), archive/listing40/main.cc:26
inline step diffuse_hemisphere() {
  return diffuse_step("diffuse_hemisphere", "archive/listing40/main.cc:26", diffuse::in_hemisphere, 0);
}

// "Shiny metal": two mirror-like metals.  archive/listing50/main.cc:53-64 (metal(color) of that
// step == fuzz 0 of the final class, src/cpu/material.h:38)
inline step shiny_metal() {
  return {"shiny_metal", "archive/listing50/main.cc:53-64",
          detail::trio(make_shared<lambertian>(color(0.8, 0.8, 0.0)), make_shared<lambertian>(color(0.7, 0.3, 0.3)),
                       make_shared<metal>(color(0.8, 0.8, 0.8), 0.0), make_shared<metal>(color(0.8, 0.6, 0.2), 0.0)),
          detail::default_camera(), 400, detail::kAspect, 100, 50};
}
// "Fuzzed metal": fuzz 0.3 and 1.0.  archive/listing52/main.cc:53-64
inline step fuzzed_metal() {
  return {"fuzzed_metal", "archive/listing52/main.cc:53-64",
          detail::trio(make_shared<lambertian>(color(0.8, 0.8, 0.0)), make_shared<lambertian>(color(0.7, 0.3, 0.3)),
                       make_shared<metal>(color(0.8, 0.8, 0.8), 0.3), make_shared<metal>(color(0.8, 0.6, 0.2), 1.0)),
          detail::default_camera(), 400, detail::kAspect, 100, 50};
}
// two glass spheres and a fully fuzzed metal.  archive/listing55/main.cc:52-63
inline step glass_pair() {
  return {"glass_pair", "archive/listing55/main.cc:52-63",
          detail::trio(make_shared<lambertian>(color(0.8, 0.8, 0.0)), make_shared<dielectric>(1.5),
                       make_shared<dielectric>(1.5), make_shared<metal>(color(0.8, 0.6, 0.2), 1.0)),
          detail::default_camera(), 400, detail::kAspect, 100, 50};
}
// one glass sphere.  archive/listing59/main.cc:52-63
inline step glass_single() {
  return {"glass_single", "archive/listing59/main.cc:52-63",
          detail::trio(make_shared<lambertian>(color(0.8, 0.8, 0.0)), make_shared<lambertian>(color(0.1, 0.2, 0.5)),
                       make_shared<dielectric>(1.5), make_shared<metal>(color(0.8, 0.6, 0.2), 0.0)),
          detail::default_camera(), 1920, detail::kAspect, 100, 50};
}
// "Hollow glass": a radius -0.4 bubble inside the glass sphere.  archive/listing61/main.cc:53-65
inline step hollow_glass() {
  return {"hollow_glass", "archive/listing61/main.cc:53-65",
          detail::trio(make_shared<lambertian>(color(0.8, 0.8, 0.0)), make_shared<lambertian>(color(0.1, 0.2, 0.5)),
                       make_shared<dielectric>(1.5), make_shared<metal>(color(0.8, 0.6, 0.2), 0.0), -0.4),
          detail::default_camera(), 400, detail::kAspect, 100, 50};
}
// "Wide-angle view": two touching spheres of radius cos(pi/4), vfov 90.  archive/listing63/main.cc:51-61
inline step wide_angle() {
  const double R = std::cos(3.1415926535897932385 / 4);
  hittable_list world;
  world.add(make_shared<sphere>(point3(-R, 0, -1), R, make_shared<lambertian>(color(0, 0, 1))));
  world.add(make_shared<sphere>(point3(R, 0, -1), R, make_shared<lambertian>(color(1, 0, 0))));
  return {"wide_angle", "archive/listing63/main.cc:51-61", world, detail::default_camera(90.0), 1920,
          detail::kAspect, 100, 50};
}
// "A distant view": camera (-2,2,1) -> (0,0,-1), vfov 90, bubble -0.45.  archive/listing65/main.cc:53-65
inline step distant_view() {
  return {"distant_view", "archive/listing65/main.cc:53-65",
          detail::trio(make_shared<lambertian>(color(0.8, 0.8, 0.0)), make_shared<lambertian>(color(0.1, 0.2, 0.5)),
                       make_shared<dielectric>(1.5), make_shared<metal>(color(0.8, 0.6, 0.2), 0.0), -0.45),
          detail::positioned_camera(point3(-2, 2, 1), point3(0, 0, -1), 90.0), 1920, detail::kAspect, 100, 50};
}
// "Zooming in": the same with vfov 20.  archive/listing66/main.cc:53-65
inline step zoomed_view() {
  step s = distant_view();
  s.name = "zoomed_view"; s.listing = "archive/listing66/main.cc:53-65";
  s.cam = detail::positioned_camera(point3(-2, 2, 1), point3(0, 0, -1), 20.0);
  s.image_width = 400;
  return s;
}
// "Depth of field": aperture 2, focus on the centre sphere.  archive/listing69/main.cc:43-71
// (= three_sphere_scene(), BASELINE.json configs[0])
inline step defocus_blur() {
  const point3 lookfrom(3, 3, 2), lookat(0, 0, -1);
  return {"defocus_blur", "archive/listing69/main.cc:43-71", three_sphere_scene(),
          camera(lookfrom, lookat, vec3(0, 1, 0), 20.0, detail::kAspect, 2.0, (lookfrom - lookat).length()),
          400, detail::kAspect, 100, 50};
}
// "Final scene": the book cover.  archive/listing70/main.cc:38-106 == src/cpu/main.cc:32-99
inline step final_cover() {
  return {"final_cover", "archive/listing70/main.cc:38-106", random_scene(),
          camera(point3(13, 2, 3), point3(0, 0, 0), vec3(0, 1, 0), 20.0, 3.0 / 2.0, 0.1, 10.0),
          1200, 3.0 / 2.0, 500, 50};
}

inline std::vector<step> all() {
  return {diffuse_first(), diffuse_gamma(), diffuse_lambertian(), diffuse_hemisphere(), shiny_metal(), fuzzed_metal(), glass_pair(), glass_single(), hollow_glass(), wide_angle(),
          distant_view(), zoomed_view(), defocus_blur(), final_cover()};
}

}  // namespace book
}  // namespace rt
#endif
