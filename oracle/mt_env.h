/* oracle/mt_env.h -- test infrastructure.  Force-included when oracle/Makefile builds the
 * reference's cpu_multi_threading_ray_tracer (src/cpu-multi-threading/main.cc): the Makefile's sed
 * pass replaces the three literals of that main() that fix the workload (image_width,
 * samples_per_pixel, n_threads; main.cc:85,87,105) by calls of rt_env_int(), so one binary can be
 * timed at any width / spp / thread count.  Nothing else of the program changes. */
#ifndef RT_ORACLE_MT_ENV_H
#define RT_ORACLE_MT_ENV_H
#include <atomic>  /* the program uses std::atomic_int without including <atomic> (main.cc:107) */
#include <cstdlib>
static inline int rt_env_int(const char *name, int dflt) {
  const char *v = std::getenv(name);
  return v && *v ? std::atoi(v) : dflt;
}
#endif
