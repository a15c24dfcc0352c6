/* orc_misc.cpp — small entry points exposing the oracle's primitives to the tests.
 * TEST INFRASTRUCTURE ONLY (see orc_common.h). */
#include "orc_common.h"
using namespace orc;

extern "C" {

/* owl::LCG<4> (owl/common/math/random.h:56-96): init(v0,v1) then n draws; returns the state */
uint32_t orc_lcg_draws(uint32_t v0, uint32_t v1, int n, float* out) {
  LCG4 r; r.init(v0, v1);
  for (int i = 0; i < n; i++) out[i] = r.next();
  return r.state;
}

void orc_counter_draws(uint64_t seed, uint64_t id, int n, float* out) {
  CounterRng r; r.init(seed, id);
  for (int i = 0; i < n; i++) out[i] = r.next();
}

void orc_normalize(float x, float y, float z, float* out) {
  const vec3 v = normalize(v3(x, y, z));
  out[0] = v.x; out[1] = v.y; out[2] = v.z;
}

void orc_sincos(float x, float* s, float* c) { pm_sincosf(x, s, c); }
float orc_acos(float x) { return pm_acosf(x); }

/* randomPointInUnitSphere (common/cuda/helpers.h:27-34) for the LCG stream (v0, v1) */
void orc_sphere_point(uint32_t v0, uint32_t v1, float* out) {
  LCG4 r; r.init(v0, v1);
  const vec3 p = randomPointInUnitSphere(r);
  out[0] = p.x; out[1] = p.y; out[2] = p.z;
}

}  // extern "C"
