/* orc_scene.cpp — CPU oracle for ray/scene intersection: brute force over every triangle.
 * TEST INFRASTRUCTURE ONLY (see orc_common.h).
 *
 * Parity status: UNPINNED against the reference. The reference intersects rays through
 * optixTrace (photon-mapping/cuda/deviceCode.cu:27,40; ray-tracer/cuda/deviceCode.cu:29-41,70-85,
 * 129-141); NVIDIA OptiX 7.x is closed source, has no CPU implementation and its SDK headers are
 * not on this machine, and the reference holds no test or golden vector for hit ids.
 * This file therefore DEFINES the canonical ray/triangle test the CUDA traversal must reproduce
 * bit for bit:
 *
 *   Möller–Trumbore, IEEE binary32, one rounding per operation, no contraction:
 *     e1 = B - A, e2 = C - A, p = cross(dir, e2), det = dot(e1, p); det == 0 (or NaN) -> miss
 *     inv = 1/det, s = org - A, u = dot(s,p)*inv, reject u < 0 or u > 1
 *     q = cross(s, e1), v = dot(dir,q)*inv, reject v < 0 or u + v > 1
 *     t = dot(e2,q)*inv, accept tmin < t < tmax                 (both face orientations hit,
 *     as OptiX does with ray flags 0 / DISABLE_ANYHIT: owl_device.h:142-192)
 *   closest hit = smallest t, ties -> lowest global triangle id.
 */
#include <omp.h>
#include "orc_common.h"

namespace orc {

static inline bool tri_test(const Scene& sc, int prim, vec3 org, vec3 dir, float tmin, float tmax, float* t_out) {
  const int32_t* ix = sc.tris + 3 * (size_t)prim;
  const vec3 A = v3(sc.verts + 3 * (size_t)ix[0]);
  const vec3 B = v3(sc.verts + 3 * (size_t)ix[1]);
  const vec3 C = v3(sc.verts + 3 * (size_t)ix[2]);
  const vec3 e1 = B - A, e2 = C - A;
  const vec3 p = cross(dir, e2);
  const float det = dot(e1, p);
  if (!(det != 0.f)) return false; /* also rejects NaN */
  const float inv = 1.f / det;
  const vec3 s = org - A;
  const float u = dot(s, p) * inv;
  if (!(u >= 0.f && u <= 1.f)) return false;
  const vec3 q = cross(s, e1);
  const float v = dot(dir, q) * inv;
  if (!(v >= 0.f && u + v <= 1.f)) return false;
  const float t = dot(e2, q) * inv;
  if (!(t > tmin && t < tmax)) return false;
  *t_out = t;
  return true;
}

Hit closest_hit(const Scene& sc, vec3 org, vec3 dir, float tmin, float tmax) {
  Hit best{-1, tmax};
  for (int i = 0; i < sc.nt; i++) {
    float t;
    if (!tri_test(sc, i, org, dir, tmin, tmax, &t)) continue;
    if (best.prim < 0 || t < best.t) { best.prim = i; best.t = t; } /* ascending ids: ties keep the lowest */
  }
  return best;
}

bool any_hit(const Scene& sc, vec3 org, vec3 dir, float tmin, float tmax) {
  for (int i = 0; i < sc.nt; i++) {
    float t;
    if (tri_test(sc, i, org, dir, tmin, tmax, &t)) return true;
  }
  return false;
}

/* getPrimitiveNormal (common/cuda/helpers.h:76-85): normalize(cross(B-A, C-A)), never flipped */
vec3 primitive_normal(const Scene& sc, int prim) {
  const int32_t* ix = sc.tris + 3 * (size_t)prim;
  const vec3 A = v3(sc.verts + 3 * (size_t)ix[0]);
  const vec3 B = v3(sc.verts + 3 * (size_t)ix[1]);
  const vec3 C = v3(sc.verts + 3 * (size_t)ix[2]);
  return normalize(cross(B - A, C - A));
}

}  // namespace orc

using namespace orc;

extern "C" {

void orc_trace_closest(const float* verts, int nv, const int32_t* tris, int nt, const pm_ray* rays, int64_t n,
                       pm_hit* hits) {
  Scene sc{verts, nv, tris, nt, nullptr, nullptr, 0};
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t i = 0; i < n; i++) {
    const Hit h = closest_hit(sc, v3(rays[i].org), v3(rays[i].dir), rays[i].tmin, rays[i].tmax);
    hits[i].prim = h.prim;
    hits[i].t = h.prim >= 0 ? h.t : rays[i].tmax;
  }
}

void orc_trace_occluded(const float* verts, int nv, const int32_t* tris, int nt, const pm_ray* rays, int64_t n,
                        uint8_t* occluded) {
  Scene sc{verts, nv, tris, nt, nullptr, nullptr, 0};
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t i = 0; i < n; i++)
    occluded[i] = any_hit(sc, v3(rays[i].org), v3(rays[i].dir), rays[i].tmin, rays[i].tmax) ? 1 : 0;
}

void orc_primitive_normal(const float* verts, const int32_t* tris, int prim, float* out) {
  Scene sc{verts, 0, tris, prim + 1, nullptr, nullptr, 0};
  const vec3 n = primitive_normal(sc, prim);
  out[0] = n.x; out[1] = n.y; out[2] = n.z;
}

}  // extern "C"
