/* orc_photon.cpp — CPU oracle for stage 1: photon emission and the multi-bounce walk.
 * TEST INFRASTRUCTURE ONLY (see orc_common.h).
 *
 * Restates photon-mapping/cuda/deviceCode.cu:10-136 (pointLightRayGen, shootPhoton,
 * shootCausticsPhoton, triangleMeshClosestHit, scatter*), common/cuda/helpers.h:27-89 and the launch
 * sizing of photon-mapping/src/hostCode.cu:72-90,102-110.
 *
 * Parity status: UNPINNED against the reference (its programs need OptiX; no tests, no golden
 * vectors — SURVEY §4, §8c). What IS pinned: owl::LCG<4> and normalize (tests/golden/lcg_kat.json,
 * generated from the reference's own headers). Ray/triangle hits follow orc_scene.cpp; the sphere
 * sampling uses the deterministic sin/cos/acos kernels of orc_common.h in place of CUDA's libm
 * (ULP-level difference from the reference, documented in DESIGN.md).
 *
 * Output order (ours; the reference's is a nondeterministic global atomic): light-major, then
 * launch index, then bounce.
 */
#include <vector>
#include <omp.h>
#include "orc_common.h"

using namespace orc;

namespace {

enum Event { EV_MISS = 0, EV_ABSORBED = 1, EV_DIFFUSE = 2, EV_SPECULAR = 4, EV_REFRACT = 8 };

struct Scattered { vec3 origin, direction, color; };

/* optixTrace + triangleMeshClosestHit / miss (deviceCode.cu:74-136). tmax 1e30 is OWL's Ray default
 * (owl_device.h:142-192). */
int trace_and_scatter(const Scene& sc, vec3 org, vec3 dir, vec3 color, Rng& rng, Scattered* out) {
  const Hit h = closest_hit(sc, org, dir, ORC_EPS, 1e30f);
  if (h.prim < 0) return EV_MISS;
  const pm_material& m = sc.mats[sc.tri_mesh[h.prim]];
  const float diffuseProb = m.diffuse;
  const float specularProb = m.specular + diffuseProb;
  const float transmissionProb = m.transmission + specularProb;
  const float r = rng.next();
  int ev;
  if (r < diffuseProb) ev = EV_DIFFUSE;
  else if (r < specularProb) ev = EV_SPECULAR;
  else if (r < transmissionProb) ev = EV_REFRACT;
  else return EV_ABSORBED;
  const vec3 hitPoint = org + h.t * dir;
  const vec3 n = primitive_normal(sc, h.prim);
  out->origin = hitPoint;
  if (ev == EV_DIFFUSE) out->direction = cosineSampleHemisphere(n, rng);
  else if (ev == EV_SPECULAR) out->direction = reflect(dir, n);
  else out->direction = refract3(dir, n, m.refraction_idx);
  out->color = v3(m.albedo) * color;
  return ev;
}

/* one launch index of one light; appends to `store` */
void walk(const Scene& sc, const pm_light& L, const pm_light_shape* shape, int max_depth, bool caustics, Rng rng,
          std::vector<pm_photon40>& store, int64_t* segments) {
  vec3 color = v3(L.rgb);
  vec3 org = v3(L.pos);
  vec3 dir;
  if (shape && shape->type == PM_LIGHT_SQUARE) sampleSquareLight(v3(L.pos), v3(shape->normal), shape->side_length, rng, &org, &dir);
  else dir = randomPointInUnitSphere(rng);
  for (int i = 0; i < max_depth; i++) {
    Scattered s;
    const int ev = trace_and_scatter(sc, org, dir, color, rng, &s);
    (*segments)++;
    const bool save = caustics ? (i > 0 && ev == EV_DIFFUSE) : (ev == EV_DIFFUSE && i > 0);
    if (save) {
      pm_photon40 p;
      p.pos[0] = s.origin.x; p.pos[1] = s.origin.y; p.pos[2] = s.origin.z;
      p.dir[0] = s.direction.x; p.dir[1] = s.direction.y; p.dir[2] = s.direction.z;
      p.power = 0;
      p.color[0] = color.x; p.color[1] = color.y; p.color[2] = color.z; /* colour BEFORE this hit's albedo */
      store.push_back(p);
    }
    const bool cont = caustics ? ((ev & (EV_SPECULAR | EV_REFRACT)) != 0) : (ev == EV_DIFFUSE);
    if (!cont) break;
    org = s.origin; dir = s.direction; color = s.color;
  }
}

}  // namespace

extern "C" {

/* number of launch indices of light l (hostCode.cu:86,102-110): int(power * int(casted / sum power)) */
int64_t orc_light_paths(const pm_light* lights, int n_lights, int64_t casted, int l) {
  double total = 0;
  for (int i = 0; i < n_lights; i++) total += lights[i].power;
  const int per_watt = (int)((double)(int)casted / total);
  return (int64_t)(int)(lights[l].power * per_watt);
}

/* Same contract as pm_emit_photons. Returns the number of photons (also when it exceeds capacity;
 * only the first `capacity` are written). */
int64_t orc_emit_photons_shaped(const float* verts, int nv, const int32_t* tris, int nt, const int32_t* tri_mesh,
                                const pm_material* mats, int n_mesh, const pm_light* lights, const pm_light_shape* shapes,
                                int n_lights, int64_t casted, int max_depth, int caustics_mode, uint64_t seed, int rng_mode,
                                int64_t first_photon, int64_t photon_stride, pm_photon40* out, int64_t capacity,
                                int64_t* n_paths, int64_t* n_segments) {
  Scene sc{verts, nv, tris, nt, tri_mesh, mats, n_mesh};
  int64_t written = 0, paths = 0, segs = 0;
  for (int l = 0; l < n_lights; l++) {
    const int64_t n = orc_light_paths(lights, n_lights, casted, l);
    const int64_t mine = n > first_photon ? (n - first_photon + photon_stride - 1) / photon_stride : 0;
    paths += mine;
    const int nth = omp_get_max_threads();
    std::vector<std::vector<pm_photon40>> parts(nth);
    std::vector<int64_t> part_segs(nth, 0);
    /* static contiguous chunks keep the concatenation in launch-index order */
#pragma omp parallel num_threads(nth)
    {
      const int t = omp_get_thread_num();
      const int64_t b = mine * t / nth, e = mine * (t + 1) / nth;
      for (int64_t j = b; j < e; j++) {
        const int64_t idx = first_photon + j * photon_stride;
        Rng rng;
        if (rng_mode == PM_RNG_LCG_PARITY) rng.init_lcg((uint32_t)idx, 0u); /* deviceCode.cu:59, id.y == 0 */
        else rng.init_counter(seed, ((uint64_t)((caustics_mode ? n_lights : 0) + l) << 32) | (uint64_t)idx);
        walk(sc, lights[l], shapes ? &shapes[l] : nullptr, max_depth, caustics_mode != 0, rng, parts[t], &part_segs[t]);
      }
    }
    for (int t = 0; t < nth; t++) {
      segs += part_segs[t];
      for (const pm_photon40& p : parts[t]) {
        if (written < capacity) out[written] = p;
        written++;
      }
    }
  }
  if (n_paths) *n_paths = paths;
  if (n_segments) *n_segments = segs;
  return written;
}

int64_t orc_emit_photons(const float* verts, int nv, const int32_t* tris, int nt, const int32_t* tri_mesh,
                         const pm_material* mats, int n_mesh, const pm_light* lights, int n_lights, int64_t casted,
                         int max_depth, int caustics_mode, uint64_t seed, int rng_mode, int64_t first_photon,
                         int64_t photon_stride, pm_photon40* out, int64_t capacity, int64_t* n_paths,
                         int64_t* n_segments) {
  return orc_emit_photons_shaped(verts, nv, tris, nt, tri_mesh, mats, n_mesh, lights, nullptr, n_lights, casted, max_depth,
                                 caustics_mode, seed, rng_mode, first_photon, photon_stride, out, capacity, n_paths, n_segments);
}

/* record conversion of ray-tracer/src/hostCode.cu:65-83 */
void orc_photons_to_maps(const pm_photon40* g, int64_t ng, const pm_photon40* c, int64_t nc, pm_photon44* out_global,
                         pm_photon44* out_caustic) {
  auto conv = [](const pm_photon40& s, float power) {
    pm_photon44 d;
    memset(&d, 0, sizeof d);
    for (int k = 0; k < 3; k++) { d.pos[k] = s.pos[k]; d.dir[k] = s.dir[k]; d.color[k] = s.color[k]; }
    d.power = power;
    return d;
  };
  for (int64_t i = 0; i < ng; i++) out_global[i] = conv(g[i], 1.0f);
  for (int64_t k = 0; k < nc; k++) {
    out_caustic[k] = conv(c[k], 0.5f);
    out_global[ng + k] = conv(c[k], 0.5f);
  }
}

}  // extern "C"
