// examples/cover.cc — the reference's cpu_ray_tracer main() (src/cpu/main.cc:78-133) on top of
// the B200 library: same scene/camera vocabulary, same defaults (1200 px wide, 3:2, 500 spp,
// depth 50), P3 image on stdout and a human log on stderr.  The pixel loops are one call into
// the sm_100a kernels (rt::render); nothing is traced on the CPU.
//
//   ./rt_cover [--scene cover|three|stress [--half H]] [--bvh] [--gpus N] [--width W] [--aspect 16:9] [--spp N] [--depth D]
//              [--seed S] [--out file.ppm] [--dump-scene file.f64] > image.ppm
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>

#include "rt/rtweekend.hpp"

using namespace rt;

int main(int argc, char **argv) {
  std::string scene = "cover", out = "-", dump;
  double aspect_ratio = 3.0 / 2.0;
  int image_width = 1200, samples_per_pixel = 500, max_depth = 50;
  unsigned long long seed = 0;
  int half = 158;
  bool use_bvh = false;
  int gpus = 1;
  for (int k = 1; k < argc; ++k) {
    std::string a = argv[k];
    auto val = [&]() -> const char * {
      if (k + 1 >= argc) { std::cerr << "missing value for " << a << "\n"; std::exit(2); }
      return argv[++k];
    };
    if (a == "--scene") scene = val();
    else if (a == "--width") image_width = std::atoi(val());
    else if (a == "--aspect") {
      const char *v = val();
      const char *c = std::strchr(v, ':');
      aspect_ratio = c ? std::atof(v) / std::atof(c + 1) : std::atof(v);
    }
    else if (a == "--spp") samples_per_pixel = std::atoi(val());
    else if (a == "--depth") max_depth = std::atoi(val());
    else if (a == "--seed") seed = std::strtoull(val(), nullptr, 10);
    else if (a == "--out") out = val();
    else if (a == "--half") half = std::atoi(val());
    else if (a == "--bvh") use_bvh = true;
    else if (a == "--gpus") gpus = std::atoi(val());
    else if (a == "--dump-scene") dump = val();
    else { std::cerr << "unknown option " << a << "\n"; return 2; }
  }
  const int image_height = static_cast<int>(image_width / aspect_ratio);

  // World
  hittable_list world = scene == "three" ? three_sphere_scene()
                        : scene == "stress" ? stress_scene(half) : random_scene();
  if (!dump.empty()) {  // f64 [n][9]: cx cy cz radius kind r g b fuzz|ir (the oracle's dump layout)
    FILE *f = std::fopen(dump.c_str(), "wb");
    if (!f) { std::perror(dump.c_str()); return 2; }
    for (const rt_sphere &s : world.flatten()) {
      double row[9] = {s.center[0], s.center[1], s.center[2], s.radius, (double)s.kind,
                       s.albedo[0], s.albedo[1], s.albedo[2], s.fuzz_or_ir};
      std::fwrite(row, sizeof(double), 9, f);
    }
    std::fclose(f);
    return 0;
  }

  // Camera
  point3 lookfrom(13, 2, 3), lookat(0, 0, 0);
  vec3 vup(0, 1, 0);
  double dist_to_focus = 10.0, aperture = 0.1;
  if (scene == "three") {
    lookfrom = point3(3, 3, 2); lookat = point3(0, 0, -1);
    dist_to_focus = (lookfrom - lookat).length(); aperture = 2.0;
  }
  camera cam(lookfrom, lookat, vup, 20, aspect_ratio, aperture, dist_to_focus);

  // Render
  std::cerr << "Image Size: " << image_width << " x " << image_height << "\n"
            << "Samples Per Pixel: " << samples_per_pixel << "\n"
            << "Spheres: " << world.objects.size() << "\n";
  try {
    auto t0 = std::chrono::high_resolution_clock::now();
    std::vector<int32_t> devices;
    for (int g = 0; g < gpus; ++g) devices.push_back(g);
    image img = gpus > 1 ? render_multi(world, cam, image_width, image_height, samples_per_pixel,
                                        max_depth, devices, seed, use_bvh)
                         : render(world, cam, image_width, image_height, samples_per_pixel, max_depth,
                                  seed, 0, use_bvh);
    auto t1 = std::chrono::high_resolution_clock::now();
    img.write_ppm(out.c_str());
    double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    std::cerr << "Time Cost (ms): " << ms << " (device " << img.stats.frame_ms << " ms, "
              << img.stats.samples / (img.stats.frame_ms * 1e3) << " Msamples/s, "
              << img.stats.rays / (img.stats.frame_ms * 1e3) << " Mrays/s)\n";
  } catch (const rt::error &e) {
    std::cerr << e.what() << "\n";
    return 99;  // the reference's checkCudaErrors exit code (src/gpu/cuda_utility.h:8-18)
  }
  return 0;
}
