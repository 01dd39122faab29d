// examples/book.cc — renders the book's intermediate steps (include/rt/book_scenes.hpp, the
// scenes of the reference's archive/listing50…70) through the rt:: facade.
//   rt_book --list
//   rt_book <step> [--width W] [--spp N] [--seed S] [--out file.ppm] [--dump file.f64]
// --dump writes the world (N x 9 doubles: cx cy cz r kind r g b fuzz|ir) followed by the 10
// camera doubles (lookfrom, lookat, vfov, aspect, aperture, focus_dist) and exits.
#include <cstdlib>
#include <cstring>
#include <iostream>

#include "rt/book_scenes.hpp"

int main(int argc, char **argv) {
  std::vector<rt::book::step> steps = rt::book::all();
  if (argc < 2 || !strcmp(argv[1], "--list")) {
    for (const auto &s : steps) std::cout << s.name << "\t" << s.listing << "\t" << s.world.objects.size() << " spheres\t"
                                     << s.image_width << " px\t" << s.samples_per_pixel << " spp\tdepth " << s.max_depth << "\n";
    return argc < 2 ? 2 : 0;
  }
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
    FILE *f = std::fopen(dump.c_str(), "wb");
    if (!f) { std::perror(dump.c_str()); return 2; }
    for (const rt_sphere &s : st->world.flatten()) {
      double row[9] = {s.center[0], s.center[1], s.center[2], s.radius, (double)s.kind,
                       s.albedo[0], s.albedo[1], s.albedo[2], s.fuzz_or_ir};
      std::fwrite(row, sizeof(double), 9, f);
    }
    const rt_camera &c = st->cam.abi();
    double cam[10] = {c.lookfrom[0], c.lookfrom[1], c.lookfrom[2], c.lookat[0], c.lookat[1], c.lookat[2],
                      c.vfov_deg, c.aspect_ratio, c.aperture, c.focus_dist};
    std::fwrite(cam, sizeof(double), 10, f);
    std::fclose(f);
    return 0;
  }
  const int height = static_cast<int>(width / st->aspect_ratio);
  std::cerr << st->name << " (" << st->listing << "): " << width << " x " << height << ", " << spp << " spp\n";
  try {
    rt::image img = rt::render(st->world, st->cam, width, height, spp, st->max_depth, seed);
    img.write_ppm(out.c_str());
    std::cerr << "Time Cost (ms): " << img.stats.frame_ms << "\n";
  } catch (const rt::error &e) {
    std::cerr << e.what() << "\n";
    return 99;
  }
  return 0;
}
