/* oracle/rt_oracle.c — TEST INFRASTRUCTURE: CPU parity oracle (see rt_oracle.h header).
 * Plain C11, double precision, no FMA contraction (-ffp-contract=off in oracle/Makefile):
 * every operation is written in the same order as the reference expression it restates so
 * that, compiled with this image's gcc, results are bit-identical to the reference's own
 * binary (pinned by tests/test_oracle_pinning.py).
 * All file:line citations are relative to /root/reference/.
 */
#include "rt_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ vec3 (src/cpu/vec3.h) */
typedef struct { double x, y, z; } v3;

static inline v3 V(double x, double y, double z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }   /* :68 */
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }   /* :72 */
static inline v3 vmulv(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }  /* :76 */
static inline v3 vscale(double t, v3 a) { return V(t * a.x, t * a.y, t * a.z); }   /* :80 */
static inline v3 vdiv(v3 a, double t) { return vscale(1 / t, a); }                 /* :86 */
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }
static inline double vdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; } /* :88 */
static inline double vlen2(v3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }      /* :43 */
static inline double vlen(v3 a) { return sqrt(vlen2(a)); }
static inline v3 vcross(v3 a, v3 b) {                                               /* :92 */
  return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline v3 vunit(v3 a) { return vdiv(a, vlen(a)); }                           /* :98 */
static inline int vnear_zero(v3 a) {                                                /* :46 */
  const double s = 1e-8;
  return (fabs(a.x) < s) && (fabs(a.y) < s) && (fabs(a.z) < s);
}
static inline v3 vreflect(v3 v, v3 n) {                                             /* :124 */
  return vsub(v, vscale(2 * vdot(v, n), n));
}
static inline v3 vrefract(v3 uv, v3 n, double etai_over_etat) {                     /* :126-131 */
  double cos_theta = fmin(vdot(vneg(uv), n), 1.0);
  v3 perp = vscale(etai_over_etat, vadd(uv, vscale(cos_theta, n)));
  v3 par = vscale(-sqrt(fabs(1.0 - vlen2(perp))), n);
  return vadd(perp, par);
}

/* ------------------------------------------------------------------ mt19937 + libstdc++ */
/* std::mt19937 is fully specified by the C++ standard; generate_canonical<double,53> is
 * libstdc++ 13's (bits/random.tcc): two 32-bit draws, sum = g1 + g2*2^32, / 2^64, and a
 * nextafter(1,0) fix-up.  uniform_real_distribution(0,1) returns it unchanged. */
struct orc_ctx {
  uint32_t mt[624];
  int idx;
};

static void mt_seed(orc_ctx *c, uint32_t seed) {
  c->mt[0] = seed;
  for (int i = 1; i < 624; ++i)
    c->mt[i] = 1812433253u * (c->mt[i - 1] ^ (c->mt[i - 1] >> 30)) + (uint32_t)i;
  c->idx = 624;
}

static uint32_t mt_next(orc_ctx *c) {
  if (c->idx >= 624) {
    for (int i = 0; i < 624; ++i) {
      uint32_t y = (c->mt[i] & 0x80000000u) | (c->mt[(i + 1) % 624] & 0x7fffffffu);
      c->mt[i] = c->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    c->idx = 0;
  }
  uint32_t y = c->mt[c->idx++];
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

orc_ctx *orc_create(void) {
  orc_ctx *c = (orc_ctx *)malloc(sizeof(orc_ctx));
  if (c) mt_seed(c, 5489u);
  return c;
}
void orc_destroy(orc_ctx *c) { free(c); }

double orc_random_double(orc_ctx *c) { /* src/cpu/rtweekend.h:27-31 */
  double sum = 0.0, tmp = 1.0;
  sum += (double)mt_next(c) * tmp; tmp *= 4294967296.0;
  sum += (double)mt_next(c) * tmp; tmp *= 4294967296.0;
  double ret = sum / tmp;
  if (ret >= 1.0) ret = nextafter(1.0, 0.0);
  return ret;
}
static inline double rnd_range(orc_ctx *c, double lo, double hi) { /* rtweekend.h:33-36 */
  return lo + (hi - lo) * orc_random_double(c);
}

/* g++ evaluates the arguments of vec3(a(), b(), c()) RIGHT TO LEFT on x86-64 (unspecified
 * by the language; pinned empirically against oracle/_ref/ref_harness). */
static v3 rnd_vec3(orc_ctx *c) { /* vec3.h:11-13 */
  double z = orc_random_double(c), y = orc_random_double(c), x = orc_random_double(c);
  return V(x, y, z);
}
static v3 rnd_vec3_range(orc_ctx *c, double lo, double hi) { /* vec3.h:15-18 */
  double z = rnd_range(c, lo, hi), y = rnd_range(c, lo, hi), x = rnd_range(c, lo, hi);
  return V(x, y, z);
}
static v3 mt_in_unit_sphere(orc_ctx *c) { /* vec3.h:105-112 */
  for (;;) {
    v3 p = rnd_vec3_range(c, -1, 1);
    if (vlen2(p) >= 1) continue;
    return p;
  }
}
static v3 mt_unit_vector(orc_ctx *c) { return vunit(mt_in_unit_sphere(c)); } /* vec3.h:114 */
static v3 mt_in_unit_disk(orc_ctx *c) { /* vec3.h:133-140 */
  for (;;) {
    double y = rnd_range(c, -1, 1), x = rnd_range(c, -1, 1);
    v3 p = V(x, y, 0);
    if (vlen2(p) >= 1) continue;
    return p;
  }
}

/* ------------------------------------------------------------------ Philox4x32-10 */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* The product's stream contract (mirrored in csrc/rt_kernels.cu):
 *   key = (pixel index in output order, absolute sample index)
 *   ctr = (slot, 0, seed_lo, seed_hi); slot 0 = camera, slot k>=1 = scatter at the k-th hit
 *   uniform_i = (word_i >> 8) * 2^-24  in [0,1)  (exact in both f32 and f64) */
static void philox_uniforms(uint32_t pixel, uint32_t sample, uint32_t slot, uint64_t seed,
                            double xi[4]) {
  uint32_t ctr[4] = {slot, 0u, (uint32_t)seed, (uint32_t)(seed >> 32)};
  uint32_t key[2] = {pixel, sample}, o[4];
  orc_philox4x32_10(ctr, key, o);
  for (int i = 0; i < 4; ++i) xi[i] = (double)(o[i] >> 8) * (1.0 / 16777216.0);
}

#define ORC_TWO_PI 6.283185307179586476925286766559
/* Direct samplers, distribution-equivalent to the reference's rejection loops. */
static v3 direct_unit_vector(double x0, double x1) { /* == vec3.h:114 in distribution */
  double z = 1.0 - 2.0 * x0;
  double rxy = sqrt(fmax(0.0, 1.0 - z * z));
  double phi = ORC_TWO_PI * x1;
  return V(rxy * cos(phi), rxy * sin(phi), z);
}
static v3 direct_in_unit_sphere(double x0, double x1, double x2) { /* == vec3.h:105-112 */
  return vscale(cbrt(x2), direct_unit_vector(x0, x1));
}
static v3 direct_in_unit_disk(double x0, double x1) { /* == vec3.h:133-140 */
  double r = sqrt(x0), phi = ORC_TWO_PI * x1;
  return V(r * cos(phi), r * sin(phi), 0);
}

/* ------------------------------------------------------------------ camera */
static v3 ld3(const double *p) { return V(p[0], p[1], p[2]); }
static void st3(double *p, v3 a) { p[0] = a.x; p[1] = a.y; p[2] = a.z; }

void orc_camera_init(orc_camera *cam, const double lookfrom[3], const double lookat[3],
                     const double vup[3], double vfov_deg, double aspect, double aperture,
                     double focus_dist) { /* src/cpu/camera.h:8-26 */
  const double pi = 3.1415926535897932385;       /* rtweekend.h:18 */
  double theta = vfov_deg * pi / 180.0;          /* rtweekend.h:22-24 */
  double h = tan(theta / 2);
  double viewport_height = 2.0 * h;
  double viewport_width = aspect * viewport_height;
  v3 w = vunit(vsub(ld3(lookfrom), ld3(lookat)));
  v3 u = vunit(vcross(ld3(vup), w));
  v3 v = vcross(w, u);
  v3 origin = ld3(lookfrom);
  v3 horizontal = vscale(focus_dist * viewport_width, u);
  v3 vertical = vscale(focus_dist * viewport_height, v);
  v3 llc = vsub(vsub(vsub(origin, vdiv(horizontal, 2)), vdiv(vertical, 2)),
                vscale(focus_dist, w));
  st3(cam->origin, origin); st3(cam->lower_left, llc);
  st3(cam->horizontal, horizontal); st3(cam->vertical, vertical);
  st3(cam->u, u); st3(cam->v, v); st3(cam->w, w);
  cam->lens_radius = aperture / 2;
}

/* ------------------------------------------------------------------ scenes */
static void put_sphere(double *row, v3 c, double radius, int kind, v3 albedo, double param) {
  row[0] = c.x; row[1] = c.y; row[2] = c.z; row[3] = radius; row[4] = (double)kind;
  row[5] = albedo.x; row[6] = albedo.y; row[7] = albedo.z; row[8] = param;
}

int orc_cover_scene(orc_ctx *c, double *sp, int cap) { /* src/cpu/main.cc:32-76 */
  int n = 0;
#define ORC_ADD(C, R, K, A, P)                                                    \
  do {                                                                            \
    if (n >= cap) return -1;                                                      \
    put_sphere(sp + (size_t)n * ORC_SPHERE_STRIDE, (C), (R), (K), (A), (P));      \
    ++n;                                                                          \
  } while (0)
  ORC_ADD(V(0, -1000, 0), 1000, ORC_LAMBERTIAN, V(0.5, 0.5, 0.5), 0);
  for (int a = -11; a < 11; a++) {
    for (int b = -11; b < 11; b++) {
      double choose_mat = orc_random_double(c);
      /* point3 center(a + 0.9*rd(), 0.2, b + 0.9*rd()): arguments right-to-left (g++) */
      double cz = b + 0.9 * orc_random_double(c);
      double cx = a + 0.9 * orc_random_double(c);
      v3 center = V(cx, 0.2, cz);
      if (vlen(vsub(center, V(4, 0.2, 0))) > 0.9) {
        if (choose_mat < 0.8) {
          /* color::random() * color::random(): right operand evaluated first (g++) */
          v3 rhs = rnd_vec3(c);
          v3 lhs = rnd_vec3(c);
          ORC_ADD(center, 0.2, ORC_LAMBERTIAN, vmulv(lhs, rhs), 0);
        } else if (choose_mat < 0.95) {
          v3 albedo = rnd_vec3_range(c, 0.5, 1);
          double fuzz = rnd_range(c, 0, 0.5);
          ORC_ADD(center, 0.2, ORC_METAL, albedo, fuzz < 1 ? fuzz : 1); /* material.h:38 */
        } else {
          ORC_ADD(center, 0.2, ORC_DIELECTRIC, V(1, 1, 1), 1.5);
        }
      }
    }
  }
  ORC_ADD(V(0, 1, 0), 1.0, ORC_DIELECTRIC, V(1, 1, 1), 1.5);
  ORC_ADD(V(-4, 1, 0), 1.0, ORC_LAMBERTIAN, V(0.4, 0.2, 0.1), 0);
  ORC_ADD(V(4, 1, 0), 1.0, ORC_METAL, V(0.7, 0.6, 0.5), 0.0);
  return n;
}

int orc_three_scene(double *sp, int cap) { /* archive/listing69/main.cc:43-71 */
  int n = 0;
  ORC_ADD(V(0.0, -100.5, -1.0), 100.0, ORC_LAMBERTIAN, V(0.8, 0.8, 0.0), 0);
  ORC_ADD(V(0.0, 0.0, -1.0), 0.5, ORC_LAMBERTIAN, V(0.1, 0.2, 0.5), 0);
  ORC_ADD(V(-1.0, 0.0, -1.0), 0.5, ORC_DIELECTRIC, V(1, 1, 1), 1.5);
  ORC_ADD(V(-1.0, 0.0, -1.0), -0.45, ORC_DIELECTRIC, V(1, 1, 1), 1.5);
  ORC_ADD(V(1.0, 0.0, -1.0), 0.5, ORC_METAL, V(0.8, 0.6, 0.2), 0.0);
  return n;
#undef ORC_ADD
}

/* ------------------------------------------------------------------ model switch
 * ORC_MODEL_CPU  = src/cpu/ (the spec, default).
 * ORC_MODEL_GPU  = the semantics of the reference's CUDA program src/gpu/ (v4 book), restated in
 *                  f64: strict root interval (src/gpu/sphere.h:31-36 + interval.h:19-21), metal
 *                  fuzz point on the unit sphere (src/gpu/material.h:52), W x H pixel footprints
 *                  (src/gpu/camera.h:92-103,140-157).  Philox mode only. */
static int g_model = ORC_MODEL_CPU;
void orc_set_model(int model) { g_model = model; }
int orc_get_model(void) { return g_model; }

/* ------------------------------------------------------------------ intersection */
typedef struct { v3 p, normal; double t; int front_face; int id; } hitrec;

/* src/cpu/sphere.h:24-51 */
static int sphere_hit(const double *s, v3 org, v3 dir, double t_min, double t_max, hitrec *rec) {
  v3 center = V(s[0], s[1], s[2]);
  double radius = s[3];
  v3 oc = vsub(org, center);
  double a = vlen2(dir);
  double half_b = vdot(oc, dir);
  double c = vlen2(oc) - radius * radius;
  double discriminant = half_b * half_b - a * c;
  if (discriminant < 0) return 0;
  double sqrtd = sqrt(discriminant);
  double root = (-half_b - sqrtd) / a;
  if (g_model == ORC_MODEL_GPU) { /* !ray_t.surrounds(root), src/gpu/sphere.h:31-36 */
    if (!(t_min < root && root < t_max)) {
      root = (-half_b + sqrtd) / a;
      if (!(t_min < root && root < t_max)) return 0;
    }
  } else if (root < t_min || t_max < root) {
    root = (-half_b + sqrtd) / a;
    if (root < t_min || t_max < root) return 0;
  }
  rec->t = root;
  rec->p = vadd(org, vscale(root, dir));                /* ray.h:16 */
  v3 outward = vdiv(vsub(rec->p, center), radius);
  rec->front_face = vdot(dir, outward) < 0;             /* hittable.h:16-19 */
  rec->normal = rec->front_face ? outward : vneg(outward);
  return 1;
}

/* src/cpu/hittable_list.h:28-43 */
static int world_hit(const double *sp, int n, v3 org, v3 dir, double t_min, double t_max,
                     hitrec *rec) {
  hitrec tmp;
  int hit_anything = 0;
  double closest = t_max;
  for (int k = 0; k < n; ++k) {
    if (sphere_hit(sp + (size_t)k * ORC_SPHERE_STRIDE, org, dir, t_min, closest, &tmp)) {
      hit_anything = 1;
      closest = tmp.t;
      *rec = tmp;
      rec->id = k;
    }
  }
  return hit_anything;
}

int orc_closest_hit(const double *spheres, int n, const double org[3], const double dir[3],
                    double t_min, double t_max, double *t, double p[3], double normal[3],
                    int *front_face) {
  hitrec rec;
  if (!world_hit(spheres, n, ld3(org), ld3(dir), t_min, t_max, &rec)) return -1;
  if (t) *t = rec.t;
  if (p) st3(p, rec.p);
  if (normal) st3(normal, rec.normal);
  if (front_face) *front_face = rec.front_face;
  return rec.id;
}

void orc_primary_hits(const double *spheres, int n, const orc_camera *cam, int W, int H,
                      int32_t *ids, double *t, double *normal, uint8_t *front_face) {
  orc_primary_hits_rows(spheres, n, cam, W, H, 0, H, ids, t, normal, front_face);
}

/* The same table for image rows [row_begin, row_end) only (outputs hold just those rows), so that
 * full-size tables can be computed by several threads at test time. */
void orc_primary_hits_rows(const double *spheres, int n, const orc_camera *cam, int W, int H,
                           int row_begin, int row_end, int32_t *ids, double *t, double *normal,
                           uint8_t *front_face) {
  v3 origin = ld3(cam->origin), llc = ld3(cam->lower_left);
  v3 hor = ld3(cam->horizontal), ver = ld3(cam->vertical), zero = V(0, 0, 0);
  for (int row = row_begin; row < row_end; ++row) {
    int j = H - 1 - row;
    for (int i = 0; i < W; ++i) {
      double u = (i + 0.5) / (W - 1), v = (j + 0.5) / (H - 1); /* main.cc:116-117, xi=.5 */
      v3 org = vadd(origin, zero);
      v3 dir = vsub(vsub(vadd(vadd(llc, vscale(u, hor)), vscale(v, ver)), origin), zero);
      hitrec rec;
      size_t px = (size_t)(row - row_begin) * W + i;
      if (world_hit(spheres, n, org, dir, 0.001, INFINITY, &rec)) {
        ids[px] = rec.id; t[px] = rec.t; front_face[px] = (uint8_t)rec.front_face;
        st3(normal + px * 3, rec.normal);
      } else {
        ids[px] = -1; t[px] = 0; front_face[px] = 0;
        st3(normal + px * 3, zero);
      }
    }
  }
}

/* ------------------------------------------------------------------ materials */
typedef struct {
  int mode;
  orc_ctx *c;                 /* MT mode */
  uint32_t pixel, sample, slot; /* Philox mode: slot is bumped at every hit */
  uint64_t seed;
} rng_t;

/* src/cpu/material.h:82-87 */
static double reflectance(double cosine, double ref_idx) {
  double r0 = (1 - ref_idx) / (1 + ref_idx);
  r0 = r0 * r0;
  return r0 + (1 - r0) * pow((1 - cosine), 5);
}

/* Returns 1 if the path continues.  src/cpu/material.h:19-32 | 40-46 | 57-75 */
static int scatter(const double *s, v3 dir_in, const hitrec *rec, rng_t *g, v3 *atten,
                   v3 *org_out, v3 *dir_out) {
  int kind = (int)s[4];
  double xi[4] = {0, 0, 0, 0};
  if (g->mode == ORC_RNG_PHILOX) {
    g->slot += 1;
    philox_uniforms(g->pixel, g->sample, g->slot, g->seed, xi);
  }
  *org_out = rec->p;
  if (kind == ORC_LAMBERTIAN) {
    v3 ruv = g->mode == ORC_RNG_MT19937 ? mt_unit_vector(g->c) : direct_unit_vector(xi[0], xi[1]);
    v3 d = vadd(rec->normal, ruv);
    if (vnear_zero(d)) d = rec->normal;
    *dir_out = d;
    *atten = V(s[5], s[6], s[7]);
    return 1;
  }
  if (kind == ORC_DIFFUSE_BALL || kind == ORC_DIFFUSE_HEMI) {
    /* archive/listing33/main.cc:12-13 (listing36: t_min 0.001) and archive/listing40/main.cc:26 with
     * random_in_hemisphere of archive/listing39/vec3.h:110-116; the 0.5 of those listings is the
     * row's albedo */
    v3 ball = g->mode == ORC_RNG_MT19937 ? mt_in_unit_sphere(g->c)
                                         : direct_in_unit_sphere(xi[0], xi[1], xi[2]);
    if (kind == ORC_DIFFUSE_BALL) {
      *dir_out = vadd(rec->normal, ball);
    } else {
      *dir_out = vdot(ball, rec->normal) > 0.0 ? ball : vneg(ball);
    }
    *atten = V(s[5], s[6], s[7]);
    return 1;
  }
  if (kind == ORC_METAL) {
    v3 reflected = vreflect(vunit(dir_in), rec->normal);
    v3 ball = g->mode == ORC_RNG_MT19937 ? mt_in_unit_sphere(g->c)
              : g_model == ORC_MODEL_GPU ? direct_unit_vector(xi[0], xi[1]) /* src/gpu/material.h:52 */
                                         : direct_in_unit_sphere(xi[0], xi[1], xi[2]);
    *dir_out = vadd(reflected, vscale(s[8], ball));
    *atten = V(s[5], s[6], s[7]);
    return vdot(*dir_out, rec->normal) > 0;
  }
  /* dielectric */
  *atten = V(1.0, 1.0, 1.0);
  double ir = s[8];
  double ratio = rec->front_face ? (1.0 / ir) : ir;
  v3 ud = vunit(dir_in);
  double cos_theta = fmin(vdot(vneg(ud), rec->normal), 1.0);
  double sin_theta = sqrt(1.0 - cos_theta * cos_theta);
  int cannot_refract = ratio * sin_theta > 1.0;
  int refl = cannot_refract;
  if (!refl) { /* short-circuit ||: the draw happens only if refraction is possible */
    double x = g->mode == ORC_RNG_MT19937 ? orc_random_double(g->c) : xi[0];
    refl = reflectance(cos_theta, ratio) > x;
  }
  *dir_out = refl ? vreflect(ud, rec->normal) : vrefract(ud, rec->normal, ratio);
  return 1;
}

/* src/cpu/main.cc:12-30, recursion unrolled with an attenuation stack so that the products
 * are formed in the reference's order: a1 * (a2 * (... * L)). */
static v3 ray_color(const double *sp, int n, v3 org, v3 dir, int depth, rng_t *g,
                    uint64_t *rays) {
  v3 stack[64];
  int top = 0;
  v3 L;
  for (;;) {
    if (depth <= 0) { L = V(0, 0, 0); break; }
    hitrec rec;
    ++*rays;
    if (world_hit(sp, n, org, dir, 0.001, INFINITY, &rec)) {
      v3 atten, o2, d2;
      if (scatter(sp + (size_t)rec.id * ORC_SPHERE_STRIDE, dir, &rec, g, &atten, &o2, &d2)) {
        if (top < 64) stack[top++] = atten;
        org = o2; dir = d2; depth -= 1;
        continue;
      }
      L = V(0, 0, 0);
      break;
    }
    v3 ud = vunit(dir);
    double t = 0.5 * (ud.y + 1.0);
    L = vadd(vscale(1.0 - t, V(1.0, 1.0, 1.0)), vscale(t, V(0.5, 0.7, 1.0)));
    break;
  }
  while (top > 0) L = vmulv(stack[--top], L);
  return L;
}

void orc_render(orc_ctx *c, const double *spheres, int n, const orc_camera *cam, int W, int H,
                int spp, int max_depth, int rng_mode, uint64_t seed, uint32_t sample_offset,
                int row0, int row1, double *linear, orc_stats *stats) {
  v3 origin = ld3(cam->origin), llc = ld3(cam->lower_left);
  v3 hor = ld3(cam->horizontal), ver = ld3(cam->vertical);
  v3 cu = ld3(cam->u), cv = ld3(cam->v);
  uint64_t rays = 0, samples = 0;
  if (max_depth > 64) max_depth = 64;
  rng_t g;
  memset(&g, 0, sizeof g);
  g.mode = rng_mode; g.c = c; g.seed = seed;
  for (int row = row0; row < row1; ++row) { /* src/cpu/main.cc:111-123 */
    int j = H - 1 - row;
    for (int i = 0; i < W; ++i) {
      v3 pix = V(0, 0, 0);
      for (int s = 0; s < spp; ++s) {
        double xu, xv;
        v3 disk;
        if (rng_mode == ORC_RNG_MT19937) {
          xu = orc_random_double(c);
          xv = orc_random_double(c);
        } else {
          double xi[4];
          g.pixel = (uint32_t)row * (uint32_t)W + (uint32_t)i;
          g.sample = sample_offset + (uint32_t)s;
          g.slot = 0;
          philox_uniforms(g.pixel, g.sample, 0, seed, xi);
          xu = xi[0]; xv = xi[1];
          disk = direct_in_unit_disk(xi[2], xi[3]);
        }
        double u = (i + xu) / (W - 1);
        double v = (j + xv) / (H - 1);
        if (g_model == ORC_MODEL_GPU) {
          /* pixel00 + i du + row dv + (xi-1/2)(du + dv), src/gpu/camera.h:140-157, written from
           * the lower-left corner: the vertical jitter is 1 - xi there (same distribution) */
          u = (i + xu) / W;
          v = (j + xv) / H;
        }
        if (rng_mode == ORC_RNG_MT19937) disk = mt_in_unit_disk(c);
        /* camera::get_ray, src/cpu/camera.h:28-35 */
        v3 rd = vscale(cam->lens_radius, disk);
        v3 offset = vadd(vscale(rd.x, cu), vscale(rd.y, cv));
        v3 org = vadd(origin, offset);
        v3 dir = vsub(vsub(vadd(vadd(llc, vscale(u, hor)), vscale(v, ver)), origin), offset);
        v3 col = ray_color(spheres, n, org, dir, max_depth, &g, &rays);
        pix = vadd(pix, col);
        ++samples;
      }
      st3(linear + ((size_t)(row - row0) * W + i) * 3, pix);
    }
  }
  if (stats) { stats->samples = samples; stats->rays = rays; }
}

static double clampd(double x, double lo, double hi) { /* rtweekend.h:38-44 */
  if (x < lo) return lo;
  if (x > hi) return hi;
  return x;
}

void orc_write_color(const double sum[3], int spp, int out[3]) { /* src/cpu/color.h:8-23 */
  double scale = 1.0 / spp;
  for (int k = 0; k < 3; ++k) {
    double x = sqrt(scale * sum[k]);
    out[k] = (int)(256 * clampd(x, 0.0, 0.999));
  }
}
