// examples/gpu_cover.cc — a main() shaped like the reference's CUDA program
// (src/gpu/main.cu:86-155) on the rt::gpu facade (include/rt/gpu_api.hpp): new_world,
// new_camera, render(frame_buffer, cam, world, seed), output_image; same log lines on stderr,
// P3 image on stdout.  Optional arguments override the stock literals:
//   rt_gpu_cover [--width W] [--spp N] [--seed S] [--out file.ppm] [--bvh] [--f32-hit]
//                [--dump-scene file]   (world as N x 9 doubles: cx cy cz r kind r g b fuzz|ir; no render)
#include <chrono>
#include <cstring>
#include <ctime>
#include <iostream>
#include <string>

#include "rt/gpu_api.hpp"

using namespace rt::gpu;

int main(int argc, char **argv) {
  unsigned long long seed = (unsigned long long)time(nullptr);  // main.cu:88
  int width = 0, spp = 0;
  uint32_t flags = 0;
  std::string out = "-", dump;
  for (int i = 1; i < argc; ++i) {
    auto val = [&]() -> const char * {
      if (i + 1 >= argc) { std::cerr << "missing value for " << argv[i] << "\n"; std::exit(99); }
      return argv[++i];
    };
    if (!strcmp(argv[i], "--width")) width = atoi(val());
    else if (!strcmp(argv[i], "--spp")) spp = atoi(val());
    else if (!strcmp(argv[i], "--seed")) seed = strtoull(val(), nullptr, 10);
    else if (!strcmp(argv[i], "--out")) out = val();
    else if (!strcmp(argv[i], "--dump-scene")) dump = val();
    else if (!strcmp(argv[i], "--bvh")) flags |= RT_FLAG_BVH;
    else if (!strcmp(argv[i], "--f32-hit")) flags |= RT_FLAG_F32_HIT;
    else { std::cerr << "unknown argument " << argv[i] << "\n"; return 99; }
  }

  // World
  hittable_list world;
  material_list materials;
  new_world(&world, &materials, seed);

  if (!dump.empty()) {
    std::vector<rt_sphere> flat;
    world.flatten(flat);
    FILE *f = fopen(dump.c_str(), "wb");
    if (!f) { std::cerr << "cannot open " << dump << "\n"; return 99; }
    for (const rt_sphere &s : flat) {
      double row[9] = {s.center[0], s.center[1], s.center[2], s.radius, (double)s.kind,
                       s.albedo[0], s.albedo[1], s.albedo[2], s.fuzz_or_ir};
      fwrite(row, sizeof(double), 9, f);
    }
    fclose(f);
    return 0;
  }

  // Camera
  camera cam;
  new_camera(&cam);
  if (width > 0) cam.set_image_width(width);
  if (spp > 0) cam.samples_per_pixel = spp;

  // Frame Buffer
  int n_pixels = cam.image_width * cam.image_height;
  std::vector<color> frame_buffer(n_pixels);

  // Render
  std::clog << "Image Size = " << cam.image_width << " x " << cam.image_height << "\n";
  std::clog << "Samples Per Pixel = " << cam.samples_per_pixel << "\n";
  std::clog << "Random Seed = " << seed << "\n";

  auto start = std::chrono::high_resolution_clock::now();
  rt_stats stats{};
  render(frame_buffer.data(), &cam, &world, seed, 0, flags, &stats);
  auto end = std::chrono::high_resolution_clock::now();
  auto timer_in_ms = std::chrono::duration<float, std::milli>(end - start);
  auto time_cost_in_ms = static_cast<int>(timer_in_ms.count() + 0.999f);
  auto time_cost_in_sec = (time_cost_in_ms + 999) / 1000;
  std::clog << "Time Cost (ms) = " << time_cost_in_ms << " ms\n";
  std::clog << "Time Cost (sec) = " << time_cost_in_sec << " sec\n";
  std::clog << "Kernel Time (ms) = " << stats.render_ms << " ms, " << world.objects_size() << " spheres\n";

  // Output Image
  output_image(&cam, frame_buffer.data(), out.c_str());
  return 0;
}
