/* orc_common.h — shared definitions of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under oracle/ is linked into, imported by or executed from
 * the product (libpmhot / photon-mapping_b200); only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs use it, as the checker.
 *
 * The oracle restates, on the CPU, what the reference's device code computes. Each function
 * cites the reference file:line it follows (paths relative to the reference root).
 *
 * Canonical arithmetic (see DESIGN.md "Arithmetic contract"): IEEE-754 binary32, one rounding
 * per written operation, NO implicit contraction (compiled with -ffp-contract=off), except
 *   - kNN squared distance = fmaf(dz,dz, fmaf(dx,dx, dy*dy)) with d = node - query, which is what
 *     nvcc emits for cukd's dot(d,d) (cukd/math.h:208-211, 250-251) under its default -fmad=true;
 *   - the sphere-sampling trigonometry (sinf/cosf/acosf in common/cuda/helpers.h:27-34) uses the
 *     polynomial kernels below instead of a libm, so that host and device agree to the bit.
 */
#pragma once
#include <cstdint>
#include <cstring>
#include <cmath>
#include "../include/pmhot.h"

namespace orc {

struct vec3 { float x, y, z; };

static inline vec3 v3(float x, float y, float z) { return vec3{x, y, z}; }
static inline vec3 v3(const float* p) { return vec3{p[0], p[1], p[2]}; }
static inline vec3 operator+(vec3 a, vec3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline vec3 operator-(vec3 a, vec3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline vec3 operator*(vec3 a, vec3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline vec3 operator*(vec3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
static inline vec3 operator*(float s, vec3 a) { return v3(s * a.x, s * a.y, s * a.z); }
static inline vec3 operator/(vec3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
static inline vec3 operator-(vec3 a) { return v3(-a.x, -a.y, -a.z); }
/* owl/common/math/vec.h:346-349: (x*x + y*y) + z*z */
static inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* owl/common/math/vec.h:330-335 */
static inline vec3 cross(vec3 a, vec3 b) {
  return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
/* owl/common/math/vec.h:353-356 with rsqrt = 1.f/sqrtf (owl-common.h:262) */
static inline vec3 normalize(vec3 v) { return v * (1.f / sqrtf(dot(v, v))); }
/* common/cuda/helpers.h:23-25 */
static inline float norm(vec3 v) { return sqrtf(dot(v, v)); }

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* ---- owl::LCG<4>  (owl/common/math/random.h:56-96) ------------------------------------ */
struct LCG4 {
  uint32_t state;
  void init(uint32_t val0, uint32_t val1) {
    uint32_t v0 = val0, v1 = val1, s0 = 0;
    for (int n = 0; n < 4; n++) {       /* four TEA rounds */
      s0 += 0x9e3779b9u;
      v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
      v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    state = v0;
  }
  float next() {
    state = 1664525u * state + 1013904223u;
    return (float)(state & 0x00FFFFFFu) / (float)0x01000000;
  }
};

/* ---- counter RNG (throughput mode; no reference counterpart) -------------------------- *
 * draw j of stream (seed, id): a 64-bit SplitMix-style finaliser of a counter; 24-bit floats
 * like the LCG. Same function in the device code.                                          */
struct CounterRng {
  uint64_t key; uint32_t ctr;
  void init(uint64_t seed, uint64_t id) {
    uint64_t z = seed * 0x9E3779B97F4A7C15ull + id * 0xD1B54A32D192ED03ull + 0x8CB92BA72F3D8DD7ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    key = z ^ (z >> 31); ctr = 0;
  }
  float next() {
    uint64_t z = key + (uint64_t)(++ctr) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (float)((uint32_t)(z >> 40)) / (float)0x01000000;
  }
};

/* either generator behind one interface */
struct Rng {
  int mode; LCG4 lcg; CounterRng cr;
  void init_lcg(uint32_t a, uint32_t b) { mode = PM_RNG_LCG_PARITY; lcg.init(a, b); }
  void init_counter(uint64_t seed, uint64_t id) { mode = PM_RNG_COUNTER; cr.init(seed, id); }
  float next() { return mode == PM_RNG_LCG_PARITY ? lcg.next() : cr.next(); }
};

/* ---- deterministic trigonometry (Cephes single-precision kernels) --------------------- */
static inline void pm_sincosf(float x, float* s_out, float* c_out) {
  const float kf = rintf(x * 0.636619772f);           /* x * 2/pi, ties-to-even */
  const int k = (int)kf;
  float r = fmaf(-kf, 1.57079637050628662109375f, x); /* pi/2 hi */
  r = fmaf(-kf, -4.37113900018624283e-8f, r);         /* pi/2 lo */
  const float z = r * r;
  float sp = fmaf(z, -1.9515295891e-4f, 8.3321608736e-3f);
  sp = fmaf(sp, z, -1.6666654611e-1f);
  const float s = fmaf(sp * z, r, r);
  float cp = fmaf(z, 2.443315711809948e-5f, -1.388731625493765e-3f);
  cp = fmaf(cp, z, 4.166664568298827e-2f);
  const float c = fmaf(cp, z * z, fmaf(-0.5f, z, 1.0f));
  switch (k & 3) {
    case 0: *s_out = s;  *c_out = c;  break;
    case 1: *s_out = c;  *c_out = -s; break;
    case 2: *s_out = -s; *c_out = -c; break;
    default: *s_out = -c; *c_out = s; break;
  }
}
static inline float pm_asin_core(float a) { /* |a| <= 0.5 */
  const float z = a * a;
  float p = fmaf(4.2163199048e-2f, z, 2.4181311049e-2f);
  p = fmaf(p, z, 4.5470025998e-2f);
  p = fmaf(p, z, 7.4953002686e-2f);
  p = fmaf(p, z, 1.6666752422e-1f);
  return fmaf(p * z, a, a);
}
static inline float pm_acosf(float x) {
  if (x > 1.f) x = 1.f;
  if (x < -1.f) x = -1.f;
  if (x > 0.5f) return 2.f * pm_asin_core(sqrtf(0.5f * (1.f - x)));
  if (x < -0.5f) return 3.14159274101257324f - 2.f * pm_asin_core(sqrtf(0.5f * (1.f + x)));
  return 1.57079637050628662f - pm_asin_core(x);
}

/* ---- common/cuda/helpers.h ------------------------------------------------------------ */
#define ORC_EPS 1e-3f
#define ORC_INFTY 1e10f            /* helpers.h:7 (double literal passed to a float parameter) */
#define ORC_PI 3.14159274101257324f /* float(3.141592653), helpers.h:9 */

/* helpers.h:15-17 — note: no fabs, as in the reference */
static inline bool nearZero(vec3 v) { return v.x < ORC_EPS && v.y < ORC_EPS && v.z < ORC_EPS; }

/* helpers.h:27-34 */
template <class R> static inline vec3 randomPointInUnitSphere(R& rng) {
  const float u = rng.next();
  const float v = rng.next();
  const float theta = 2.f * ORC_PI * u;
  const float phi = pm_acosf(2.f * v - 1.f);
  float st, ct, sp, cp;
  pm_sincosf(theta, &st, &ct);
  pm_sincosf(phi, &sp, &cp);
  return v3(sp * ct, sp * st, cp);
}
/* helpers.h:36-43 */
template <class R> static inline vec3 randomUnitVector(R& rng) {
  vec3 v;
  do {
    v.x = 2.f * rng.next() - 1.f;
    v.y = 2.f * rng.next() - 1.f;
    v.z = 2.f * rng.next() - 1.f;
  } while (dot(v, v) >= 1.f);
  return normalize(v);
}
/* helpers.h:45-47: normalize(n + sphere * (1 - EPS)) */
template <class R> static inline vec3 cosineSampleHemisphere(vec3 n, R& rng) {
  return normalize(n + randomPointInUnitSphere(rng) * (1 - ORC_EPS));
}
/* square emitter: LightSource::normal / side_length (common/src/world.h:17-20) are declared by the reference but
 * never read (assetImporter.cxx:122 makes every light a point light), so these semantics are ours: uniform point on
 * the square, cosine-distributed direction about the normal; draw order u, v, then cosineSampleHemisphere's two. */
template <class R> static inline void sampleSquareLight(vec3 pos, vec3 nrm, float side, R& rng, vec3* org, vec3* dir) {
  const vec3 n = normalize(nrm);
  const vec3 a = fabsf(n.x) > 0.9f ? v3(0.f, 1.f, 0.f) : v3(1.f, 0.f, 0.f);
  const vec3 t = normalize(cross(a, n));
  const vec3 b = cross(n, t);
  const float u = rng.next();
  const float v = rng.next();
  *org = pos + ((u - 0.5f) * side) * t + ((v - 0.5f) * side) * b;
  *dir = cosineSampleHemisphere(n, rng);
}
/* helpers.h:49-51: incoming - 2.f*dot(incoming,normal)*normal */
static inline vec3 reflect(vec3 in, vec3 n) { return in - (2.f * dot(in, n)) * n; }
/* helpers.h:57-74 (3-argument refract used by pass 1) */
static inline vec3 refract3(vec3 in, vec3 n, float ior) {
  float cosTheta = -dot(in, n);
  float mu;
  if (cosTheta > 0.f) { mu = 1.f / ior; } else { mu = ior; cosTheta = -cosTheta; }
  const float cosPhi = 1.f - mu * mu * (1.f - cosTheta * cosTheta);
  if (cosPhi >= 0) return mu * in + (mu * cosTheta - sqrtf(cosPhi)) * n;
  return reflect(in, n);
}
/* helpers.h:91-107 (4-argument refract used by pass 2) */
static inline bool refract4(vec3 v, vec3 n, float ni_over_nt, vec3* refracted) {
  const vec3 uv = normalize(v);
  const float dt = dot(uv, n);
  const float disc = 1.0f - ni_over_nt * ni_over_nt * (1 - dt * dt);
  if (disc > 0.f) { *refracted = ni_over_nt * (uv - n * dt) - n * sqrtf(disc); return true; }
  return false;
}
/* helpers.h:109-113 — evaluated in double as the reference's literals force; pow(x,5) is
 * defined as x*x*x*x*x here and in the device code. */
static inline float schlick(float cosv, float ior) {
  /* helpers.h:109-113: `float r0 = (1. - ior) / (1. + ior)` is a double quotient rounded to float, `r0 = r0 * r0` a
   * float product; the return expression is evaluated in double (the literal 1. promotes it). pow(x, 5) is taken as
   * the product x*x*x*x*x (stated deviation: CUDA's pow is not available to the host). */
  float r0 = (float)((1. - (double)ior) / (1. + (double)ior));
  r0 = r0 * r0;
  const double b = 1. - (double)cosv;
  const double b5 = b * b * b * b * b;
  return (float)((double)r0 + (1. - (double)r0) * b5);
}

/* ---- scene description shared by the stage-1/3 oracles -------------------------------- */
struct Scene {
  const float* verts; int nv;
  const int32_t* tris; int nt;
  const int32_t* tri_mesh;
  const pm_material* mats; int n_mesh;
};

/* Closest hit by brute force over all triangles: the ground truth for hit ids.
 * Canonical ray/triangle test (ours — OptiX's is closed source): Möller–Trumbore,
 * both orientations, accept tmin < t < tmax, closest t wins, ties → lowest triangle id. */
struct Hit { int prim; float t; };
Hit closest_hit(const Scene& sc, vec3 org, vec3 dir, float tmin, float tmax);
bool any_hit(const Scene& sc, vec3 org, vec3 dir, float tmin, float tmax);
/* helpers.h:76-85 */
vec3 primitive_normal(const Scene& sc, int prim);

}  // namespace orc
