/* pm_devmath.cuh — device-side vector maths, RNGs and sampling shared by stages 1 and 3.
 *
 * Semantics follow common/cuda/helpers.h:11-111, owl/common/math/vec.h:330-356 and
 * owl/common/math/random.h:56-96 of the reference. The translation unit is compiled with
 * -fmad=false: every written operation rounds once; fused operations appear only as explicit
 * __fmaf_rn. sin/cos/acos are fixed polynomial kernels (not CUDA's libm) so that results are
 * reproducible to the bit on any host (DESIGN.md "Arithmetic contract").
 */
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include "../../include/pmhot.h"

namespace pmd {

struct vec3 { float x, y, z; };

__device__ __forceinline__ vec3 v3(float x, float y, float z) { return vec3{x, y, z}; }
__device__ __forceinline__ vec3 v3(const float* p) { return vec3{p[0], p[1], p[2]}; }
__device__ __forceinline__ vec3 operator+(vec3 a, vec3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ vec3 operator-(vec3 a, vec3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ vec3 operator*(vec3 a, vec3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ vec3 operator*(vec3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ vec3 operator*(float s, vec3 a) { return v3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ vec3 operator/(vec3 a, float s) { return v3(__fdiv_rn(a.x, s), __fdiv_rn(a.y, s), __fdiv_rn(a.z, s)); }
__device__ __forceinline__ vec3 operator-(vec3 a) { return v3(-a.x, -a.y, -a.z); }
/* vec.h:346-349: (x*x + y*y) + z*z */
__device__ __forceinline__ float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* vec.h:330-335 */
__device__ __forceinline__ vec3 cross(vec3 a, vec3 b) {
  return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
/* vec.h:353-356, rsqrt = 1.f/sqrtf (owl-common.h:262) */
__device__ __forceinline__ vec3 normalize(vec3 v) { return v * __fdiv_rn(1.f, __fsqrt_rn(dot(v, v))); }
__device__ __forceinline__ float norm(vec3 v) { return __fsqrt_rn(dot(v, v)); }

/* ---- owl::LCG<4> (random.h:56-96) ------------------------------------------------------ */
struct LCG4 {
  uint32_t state;
  __device__ __forceinline__ void init(uint32_t val0, uint32_t val1) {
    uint32_t v0 = val0, v1 = val1, s0 = 0;
#pragma unroll
    for (int n = 0; n < 4; n++) {
      s0 += 0x9e3779b9u;
      v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
      v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    state = v0;
  }
  __device__ __forceinline__ float next() {
    state = 1664525u * state + 1013904223u;
    return __fdiv_rn((float)(state & 0x00FFFFFFu), (float)0x01000000);
  }
};

/* counter RNG of the throughput mode (same function as the CPU oracle CounterRng) */
struct CounterRng {
  uint64_t key; uint32_t ctr;
  __device__ __forceinline__ void init(uint64_t seed, uint64_t id) {
    uint64_t z = seed * 0x9E3779B97F4A7C15ull + id * 0xD1B54A32D192ED03ull + 0x8CB92BA72F3D8DD7ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    key = z ^ (z >> 31); ctr = 0;
  }
  __device__ __forceinline__ float next() {
    uint64_t z = key + (uint64_t)(++ctr) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return __fdiv_rn((float)((uint32_t)(z >> 40)), (float)0x01000000);
  }
};

struct Rng {
  int mode; LCG4 lcg; CounterRng cr;
  __device__ __forceinline__ void init_lcg(uint32_t a, uint32_t b) { mode = PM_RNG_LCG_PARITY; lcg.init(a, b); }
  __device__ __forceinline__ void init_counter(uint64_t seed, uint64_t id) { mode = PM_RNG_COUNTER; cr.init(seed, id); }
  __device__ __forceinline__ float next() { return mode == PM_RNG_LCG_PARITY ? lcg.next() : cr.next(); }
};

/* ---- deterministic trigonometry (same kernels as the CPU oracle) ------------------ */
__device__ __forceinline__ void pm_sincosf(float x, float* s_out, float* c_out) {
  const float kf = rintf(x * 0.636619772f);
  const int k = (int)kf;
  float r = __fmaf_rn(-kf, 1.57079637050628662109375f, x);
  r = __fmaf_rn(-kf, -4.37113900018624283e-8f, r);
  const float z = r * r;
  float sp = __fmaf_rn(z, -1.9515295891e-4f, 8.3321608736e-3f);
  sp = __fmaf_rn(sp, z, -1.6666654611e-1f);
  const float s = __fmaf_rn(sp * z, r, r);
  float cp = __fmaf_rn(z, 2.443315711809948e-5f, -1.388731625493765e-3f);
  cp = __fmaf_rn(cp, z, 4.166664568298827e-2f);
  const float c = __fmaf_rn(cp, z * z, __fmaf_rn(-0.5f, z, 1.0f));
  switch (k & 3) {
    case 0: *s_out = s;  *c_out = c;  break;
    case 1: *s_out = c;  *c_out = -s; break;
    case 2: *s_out = -s; *c_out = -c; break;
    default: *s_out = -c; *c_out = s; break;
  }
}
__device__ __forceinline__ float pm_asin_core(float a) {
  const float z = a * a;
  float p = __fmaf_rn(4.2163199048e-2f, z, 2.4181311049e-2f);
  p = __fmaf_rn(p, z, 4.5470025998e-2f);
  p = __fmaf_rn(p, z, 7.4953002686e-2f);
  p = __fmaf_rn(p, z, 1.6666752422e-1f);
  return __fmaf_rn(p * z, a, a);
}
__device__ __forceinline__ float pm_acosf(float x) {
  if (x > 1.f) x = 1.f;
  if (x < -1.f) x = -1.f;
  if (x > 0.5f) return 2.f * pm_asin_core(__fsqrt_rn(0.5f * (1.f - x)));
  if (x < -0.5f) return 3.14159274101257324f - 2.f * pm_asin_core(__fsqrt_rn(0.5f * (1.f + x)));
  return 1.57079637050628662f - pm_asin_core(x);
}

/* ---- common/cuda/helpers.h ------------------------------------------------------------- */
#define PMD_EPS 1e-3f
#define PMD_INFTY 1e10f
#define PMD_PI 3.14159274101257324f

/* helpers.h:15-17 (no fabs, as in the reference) */
__device__ __forceinline__ bool nearZero(vec3 v) { return v.x < PMD_EPS && v.y < PMD_EPS && v.z < PMD_EPS; }

/* helpers.h:27-34 */
template <class R> __device__ __forceinline__ vec3 randomPointInUnitSphere(R& rng) {
  const float u = rng.next();
  const float v = rng.next();
  const float theta = 2.f * PMD_PI * u;
  const float phi = pm_acosf(2.f * v - 1.f);
  float st, ct, sp, cp;
  pm_sincosf(theta, &st, &ct);
  pm_sincosf(phi, &sp, &cp);
  return v3(sp * ct, sp * st, cp);
}
/* helpers.h:36-43 */
template <class R> __device__ __forceinline__ vec3 randomUnitVector(R& rng) {
  vec3 v;
  do {
    v.x = 2.f * rng.next() - 1.f;
    v.y = 2.f * rng.next() - 1.f;
    v.z = 2.f * rng.next() - 1.f;
  } while (dot(v, v) >= 1.f);
  return normalize(v);
}
/* helpers.h:45-47 */
template <class R> __device__ __forceinline__ vec3 cosineSampleHemisphere(vec3 n, R& rng) {
  return normalize(n + randomPointInUnitSphere(rng) * (1 - PMD_EPS));
}
/* square emitter (LightSource::normal / side_length, common/src/world.h:17-20 — declared by the reference, never
 * read by it; semantics are ours): a point uniform on the square of side `side` centred at `pos` in the plane
 * orthogonal to `nrm`, direction cosine-distributed about `nrm` with the reference's diffuse sampler. Draw order:
 * u, v (position), then the two draws of cosineSampleHemisphere. */
template <class R> __device__ __forceinline__ void sampleSquareLight(vec3 pos, vec3 nrm, float side, R& rng, vec3* org, vec3* dir) {
  const vec3 n = normalize(nrm);
  const vec3 a = fabsf(n.x) > 0.9f ? v3(0.f, 1.f, 0.f) : v3(1.f, 0.f, 0.f);
  const vec3 t = normalize(cross(a, n));
  const vec3 b = cross(n, t);
  const float u = rng.next();
  const float v = rng.next();
  *org = pos + ((u - 0.5f) * side) * t + ((v - 0.5f) * side) * b;
  *dir = cosineSampleHemisphere(n, rng);
}
/* helpers.h:49-51 */
__device__ __forceinline__ vec3 reflect(vec3 in, vec3 n) { return in - (2.f * dot(in, n)) * n; }
/* helpers.h:57-74 */
__device__ __forceinline__ vec3 refract3(vec3 in, vec3 n, float ior) {
  float cosTheta = -dot(in, n);
  float mu;
  if (cosTheta > 0.f) { mu = __fdiv_rn(1.f, ior); } else { mu = ior; cosTheta = -cosTheta; }
  const float cosPhi = 1.f - mu * mu * (1.f - cosTheta * cosTheta);
  if (cosPhi >= 0) return mu * in + (mu * cosTheta - __fsqrt_rn(cosPhi)) * n;
  return reflect(in, n);
}
/* helpers.h:91-107 */
__device__ __forceinline__ bool refract4(vec3 v, vec3 n, float ni_over_nt, vec3* refracted) {
  const vec3 uv = normalize(v);
  const float dt = dot(uv, n);
  const float disc = 1.0f - ni_over_nt * ni_over_nt * (1 - dt * dt);
  if (disc > 0.f) { *refracted = ni_over_nt * (uv - n * dt) - n * __fsqrt_rn(disc); return true; }
  return false;
}
/* helpers.h:109-113: r0 is a double quotient rounded to float and squared in float; the return expression is
 * evaluated in double (its literals force that); pow(x, 5) as the product the oracle uses */
__device__ __forceinline__ float schlick(float cosv, float ior) {
  float r0 = (float)__ddiv_rn(1. - (double)ior, 1. + (double)ior);
  r0 = __fmul_rn(r0, r0);
  const double b = 1. - (double)cosv;
  const double b5 = __dmul_rn(__dmul_rn(__dmul_rn(__dmul_rn(b, b), b), b), b);
  return (float)__dadd_rn((double)r0, __dmul_rn(1. - (double)r0, b5));
}

}  // namespace pmd
