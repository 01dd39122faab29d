// examples/book.cc — renders the book's intermediate steps: include/rt/book_scenes.hpp (the
// reference's archive/listing34…70, rt:: facade) and include/rt/gpu_book_scenes.hpp (its
// archive-gpu/image7…23, rt::gpu facade, names starting with gpu_).
//   rt_book --list
//   rt_book <step> [--width W] [--spp N] [--seed S] [--out file.ppm] [--dump file.f64]
// --dump writes the world (N x 9 doubles: cx cy cz r kind r g b fuzz|ir) followed by 12 doubles:
// the camera (lookfrom, lookat, vfov, aspect, aperture, focus_dist), the RT_FLAG_* bits the step
// renders with and its bounce limit; then exits.
#include <cstdlib>
#include <cstring>
#include <iostream>

#include "rt/book_scenes.hpp"
#include "rt/gpu_book_scenes.hpp"

static void dump_scene(const std::string &path, const std::vector<rt_sphere> &flat, const rt_camera &c,
                       uint32_t flags, int max_depth) {
  FILE *f = std::fopen(path.c_str(), "wb");
  if (!f) { std::perror(path.c_str()); std::exit(2); }
  for (const rt_sphere &s : flat) {
    double row[9] = {s.center[0], s.center[1], s.center[2], s.radius, (double)s.kind,
                     s.albedo[0], s.albedo[1], s.albedo[2], s.fuzz_or_ir};
    std::fwrite(row, sizeof(double), 9, f);
  }
  double cam[12] = {c.lookfrom[0], c.lookfrom[1], c.lookfrom[2], c.lookat[0], c.lookat[1], c.lookat[2],
                    c.vfov_deg, c.aspect_ratio, c.aperture, c.focus_dist, (double)flags, (double)max_depth};
  std::fwrite(cam, sizeof(double), 12, f);
  std::fclose(f);
}

// a step of the CUDA program's history: world and camera through the rt::gpu API
static int gpu_step(const rt::gpu::book::step &gs, int argc, char **argv) {
  namespace g = rt::gpu;
  g::hittable_list world;
  g::material_list materials;
  g::camera cam;
  gs.build(&world, &materials);
  gs.setup(&cam);
  unsigned long long seed = 0;
  std::string out = "-", dump;
  for (int k = 2; k < argc; ++k) {
    auto val = [&]() -> const char * {
      if (k + 1 >= argc) { std::cerr << "missing value for " << argv[k] << "\n"; std::exit(2); }
      return argv[++k];
    };
    if (!strcmp(argv[k], "--width")) cam.set_image_width(std::atoi(val()));
    else if (!strcmp(argv[k], "--spp")) cam.samples_per_pixel = std::atoi(val());
    else if (!strcmp(argv[k], "--seed")) seed = std::strtoull(val(), nullptr, 10);
    else if (!strcmp(argv[k], "--out")) out = val();
    else if (!strcmp(argv[k], "--dump")) dump = val();
    else { std::cerr << "unknown option " << argv[k] << "\n"; return 2; }
  }
  if (!dump.empty()) {
    std::vector<rt_sphere> flat;
    world.flatten(flat);
    dump_scene(dump, flat, cam.abi(), RT_FLAG_V4_MODEL | gs.flags, cam.max_depth);
    return 0;
  }
  std::cerr << gs.name << " (" << gs.listing << "): " << cam.image_width << " x " << cam.image_height << ", "
            << cam.samples_per_pixel << " spp\n";
  std::vector<g::color> fb((size_t)cam.image_width * cam.image_height);
  rt_stats stats{};
  g::render(fb.data(), &cam, &world, seed, 0, gs.flags, &stats);
  g::output_image(&cam, fb.data(), out.c_str(), !(gs.flags & RT_FLAG_LINEAR_OUTPUT));
  std::cerr << "Time Cost (ms): " << stats.frame_ms << "\n";
  return 0;
}

int main(int argc, char **argv) {
  std::vector<rt::book::step> steps = rt::book::all();
  std::vector<rt::gpu::book::step> gsteps = rt::gpu::book::all();
  if (argc < 2 || !strcmp(argv[1], "--list")) {
    for (const auto &s : steps) std::cout << s.name << "\t" << s.listing << "\t" << s.world.objects.size() << " spheres\t"
                                     << s.image_width << " px\t" << s.samples_per_pixel << " spp\tdepth " << s.max_depth << "\n";
    for (const auto &s : gsteps) {
      rt::gpu::hittable_list w; rt::gpu::material_list m; rt::gpu::camera c;
      s.build(&w, &m); s.setup(&c);
      std::cout << s.name << "\t" << s.listing << "\t" << w.objects_size() << " spheres\t" << c.image_width
                << " px\t" << c.samples_per_pixel << " spp\tdepth " << c.max_depth << "\n";
    }
    return argc < 2 ? 2 : 0;
  }
  for (const auto &s : gsteps) if (s.name == argv[1]) return gpu_step(s, argc, argv);
  const rt::book::step *st = nullptr;
  for (const auto &s : steps) if (s.name == argv[1]) st = &s;
  if (!st) { std::cerr << "unknown step " << argv[1] << " (try --list)\n"; return 2; }
  int width = st->image_width, spp = st->samples_per_pixel;
  unsigned long long seed = 0;
  std::string out = "-", dump;
  for (int k = 2; k < argc; ++k) {
    auto val = [&]() -> const char * {
      if (k + 1 >= argc) { std::cerr << "missing value for " << argv[k] << "\n"; std::exit(2); }
      return argv[++k];
    };
    if (!strcmp(argv[k], "--width")) width = std::atoi(val());
    else if (!strcmp(argv[k], "--spp")) spp = std::atoi(val());
    else if (!strcmp(argv[k], "--seed")) seed = std::strtoull(val(), nullptr, 10);
    else if (!strcmp(argv[k], "--out")) out = val();
    else if (!strcmp(argv[k], "--dump")) dump = val();
    else { std::cerr << "unknown option " << argv[k] << "\n"; return 2; }
  }
  if (!dump.empty()) {
    dump_scene(dump, st->world.flatten(), st->cam.abi(), st->flags, st->max_depth);
    return 0;
  }
  const int height = static_cast<int>(width / st->aspect_ratio);
  std::cerr << st->name << " (" << st->listing << "): " << width << " x " << height << ", " << spp << " spp\n";
  try {
    rt::image img = rt::render(st->world, st->cam, width, height, spp, st->max_depth, seed, 0, false, st->flags);
    img.write_ppm(out.c_str());
    std::cerr << "Time Cost (ms): " << img.stats.frame_ms << "\n";
  } catch (const rt::error &e) {
    std::cerr << e.what() << "\n";
    return 99;
  }
  return 0;
}
