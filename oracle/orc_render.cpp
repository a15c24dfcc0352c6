/* orc_render.cpp — CPU oracle for stage 3: the camera pass (simpleRayGen / tracePath / ray_colour).
 * TEST INFRASTRUCTURE ONLY (see orc_common.h).
 *
 * Restates ray-tracer/cuda/deviceCode.cu:25-281 and ray-tracer/cuda/shading.h:20-121 (the kNN +
 * gatherPhotons part lives in orc_kdtree.cpp), make_rgba (owl/owl_device.h:81-93) and the camera of
 * ray-tracer/src/hostCode.cu:100-108 (set up by the caller).
 *
 * Parity status: UNPINNED against the reference (OptiX programs, no tests, no golden images —
 * SURVEY §4, §8c). Positions taken on the reference's undefined behaviour:
 *   (i)   nearZero has no fabs — kept literally (helpers.h:15-17);
 *   (ii)  reflect_or_refract_ray / calculate_refracted take the RNG BY VALUE — kept: prd.random does
 *         not advance across path depth (shading.h:24,61);
 *   (iii) after a primary miss tracePath would read a stale hit record — defined: the path ends;
 *         a final-gather ray that misses reads an uninitialised material — defined: contributes 0;
 *   (iv)  framebuffer row = H - py: py = 0 would write one row out of bounds — not written; row 0
 *         is never written and stays 0 (deviceCode.cu:224-230);
 *   (v)   photonID > num_photons (sic) kept;  (vi) jitter draws u then v.
 */
#include <vector>
#include <omp.h>
#include "orc_common.h"

namespace orc {
/* orc_kdtree.cpp */
void gather_one(const pm_photon44* nodes, int n, const float hit[3], float brdf, int k, float cutoff,
                uint64_t* heap_mem, float out[3]);
}

using namespace orc;

namespace {

constexpr int kK = 50;            /* K_NEAREST_NEIGHBOURS, shading.h:7 */
constexpr float kCutoff = 100.f;  /* K_MAX_DISTANCE, shading.h:8 */
constexpr int kDiffuseSamples = 20;

struct HitRecord { vec3 hitpoint, normal; pm_material material; };

struct RenderCtx {
  Scene sc;
  const pm_light* lights; int n_lights;
  const pm_photon44* gmap; int ng;
  const pm_photon44* cmap; int nc;
  vec3 sky;
};

/* closestHit() (deviceCode.cu:233-254): hit record with the normal flipped against the ray */
bool trace_record(const RenderCtx& R, vec3 org, vec3 dir, float tmin, float tmax, HitRecord* rec) {
  const Hit h = closest_hit(R.sc, org, dir, tmin, tmax);
  if (h.prim < 0) return false;
  rec->material = R.sc.mats[R.sc.tri_mesh[h.prim]];
  rec->hitpoint = org + dir * h.t;
  const vec3 n = primitive_normal(R.sc, h.prim);
  rec->normal = normalize((dot(dir, n) < 0.f) ? n : -n);
  return true;
}

vec3 gather(const pm_photon44* map, int n, vec3 p, float brdf, uint64_t* heap) {
  if (n < 1) return v3(0.f, 0.f, 0.f);
  const float hp[3] = {p.x, p.y, p.z};
  float out[3];
  gather_one(map, n, hp, brdf, kK, kCutoff, heap, out);
  return v3(out[0], out[1], out[2]);
}

/* shading.h:80-91 */
float specular_brdf(float spec, vec3 in_light, vec3 out_light, vec3 n) {
  const vec3 refl = reflect(in_light, n);
  if (nearZero(refl - out_light)) return spec;
  return 0.f;
}

/* shading.h:20-57; `rand` is a copy */
vec3 calculate_refracted(const pm_material& m, vec3 ray_dir, vec3 normal, LCG4 rand) {
  vec3 outward, refracted = v3(0.f, 0.f, 0.f);
  float ni_over_nt, coef, cosine;
  if (dot(ray_dir, normal) > 0.f) {
    outward = -normal;
    ni_over_nt = m.refraction_idx;
    cosine = dot(ray_dir, normal);
    cosine = sqrtf(1.f - m.refraction_idx * m.refraction_idx * (1.f - cosine * cosine));
  } else {
    outward = normal;
    ni_over_nt = 1.f / m.refraction_idx;
    cosine = -dot(ray_dir, normal);
  }
  if (refract4(ray_dir, outward, ni_over_nt, &refracted)) coef = schlick(cosine, m.refraction_idx);
  else coef = 1.f;
  if (rand.next() < coef) return reflect(ray_dir, normal);
  return refracted;
}

/* shading.h:59-78; `rand` is a copy of prd.random */
vec3 reflect_or_refract(const pm_material& m, vec3 ray_dir, vec3 normal, LCG4 rand, bool* absorbed, float* coef) {
  *absorbed = false;
  const float r = rand.next();
  if (r < m.specular) { *coef = m.specular; return reflect(ray_dir, normal); }
  if (r < m.specular + m.transmission) { *coef = m.transmission; return calculate_refracted(m, ray_dir, normal, rand); }
  *coef = 0.f;
  *absorbed = true;
  return v3(0.f, 0.f, 0.f);
}

/* ray_colour (deviceCode.cu:25-163). Returns false on a primary miss (colour = sky). */
bool ray_colour(const RenderCtx& R, vec3 org, vec3 dir, LCG4& rng, HitRecord* rec, vec3* colour, uint64_t* heap) {
  if (!trace_record(R, org, dir, ORC_EPS, ORC_INFTY, rec)) { *colour = R.sky; return false; }
  const vec3 albedo = v3(rec->material.albedo);
  const float diffuse_brdf = rec->material.diffuse / ORC_PI;

  vec3 direct = v3(0.f, 0.f, 0.f);
  for (int l = 0; l < R.n_lights; l++) {
    const pm_light& L = R.lights[l];
    vec3 light_dir = v3(L.pos) - rec->hitpoint;
    const float dist = norm(light_dir);
    light_dir = normalize(light_dir);
    const float ldn = dot(light_dir, rec->normal);
    if (ldn < 0.f) continue;
    const float vis = any_hit(R.sc, rec->hitpoint, light_dir, ORC_EPS, dist * (1.f - ORC_EPS)) ? 0.f : 1.f;
    const float sb = specular_brdf(rec->material.specular, light_dir, dir, rec->normal);
    /* light_visibility * float(power) * light_dot_norm * (1.f/(d*d)) * (brdf sum) * rgb, left to right */
    vec3 t = v3(vis, vis, vis) * (float)L.power;
    t = t * ldn;
    t = t * (1.f / (dist * dist));
    t = t * (diffuse_brdf + sb);
    t = t * v3(L.rgb);
    direct = direct + t;
  }
  const vec3 direct_term = albedo * direct;

  const vec3 caustics = gather(R.cmap, R.nc, rec->hitpoint, diffuse_brdf, heap);

  vec3 diffuse_term = v3(0.f, 0.f, 0.f);
  for (int s = 0; s < kDiffuseSamples && diffuse_brdf > 0.f; s++) {
    const vec3 normal = normalize(rec->normal);
    vec3 rdir;
    do {
      rdir = normal + randomUnitVector(rng);
    } while (nearZero(rdir));
    rdir = normalize(rdir);
    HitRecord r2;
    if (!trace_record(R, rec->hitpoint, rdir, 3 * ORC_EPS, ORC_INFTY, &r2)) continue; /* quirk (iii) */
    if (r2.material.diffuse > 0.f) {
      const float brdf2 = r2.material.diffuse / ORC_PI;
      const vec3 c = gather(R.gmap, R.ng, r2.hitpoint, brdf2, heap);
      diffuse_term = diffuse_term + c * v3(r2.material.albedo);
    }
  }
  diffuse_term = diffuse_term / (float)kDiffuseSamples;
  diffuse_term = diffuse_term * albedo;

  colour->x = 0.2f * diffuse_term.x + 0.08f * caustics.x + 0.8f * direct_term.x;
  colour->y = 0.2f * diffuse_term.y + 0.08f * caustics.y + 0.8f * direct_term.y;
  colour->z = 0.2f * diffuse_term.z + 0.08f * caustics.z + 0.8f * direct_term.z;
  return true;
}

/* make_8bit / make_rgba (owl_device.h:81-93); float->int like CUDA's cvt.rzi.s32.f32 (NaN -> 0, saturating) */
uint32_t make_8bit(float f) {
  const float x = f * 256.f;
  int i;
  if (!(x == x)) i = 0;
  else if (x >= 2147483648.f) i = 2147483647;
  else if (x <= -2147483648.f) i = -2147483647 - 1;
  else i = (int)x;
  i = i < 0 ? 0 : i;
  i = i > 255 ? 255 : i;
  return (uint32_t)i;
}

}  // namespace

extern "C" {

uint32_t orc_make_rgba(float r, float g, float b) {
  return (make_8bit(r) << 0) + (make_8bit(g) << 8) + (make_8bit(b) << 16) + (0xffu << 24);
}

/* Same contract as pm_render; additionally writes the linear float colour of every rendered pixel
 * to rgb_out (3 floats per launch index, may be NULL) for tolerance-based comparisons. */
void orc_render(const float* verts, int nv, const int32_t* tris, int nt, const int32_t* tri_mesh,
                const pm_material* mats, int n_mesh, const pm_light* lights, int n_lights,
                const pm_photon44* gmap, int ng, const pm_photon44* cmap, int nc, const pm_camera* cam,
                int W, int H, int spp, int depth, const float* sky, int64_t pixel_begin, int64_t pixel_end,
                uint32_t* fb, float* rgb_out) {
  RenderCtx R{Scene{verts, nv, tris, nt, tri_mesh, mats, n_mesh}, lights, n_lights, gmap, ng, cmap, nc, v3(sky)};
  const vec3 cpos = v3(cam->pos), d00 = v3(cam->dir_00), du = v3(cam->dir_du), dv = v3(cam->dir_dv);
#pragma omp parallel
  {
    std::vector<uint64_t> heap(kK);
#pragma omp for schedule(dynamic, 16)
    for (int64_t p = pixel_begin; p < pixel_end; p++) {
      const int px = (int)(p % W), py = (int)(p / W);
      LCG4 rng;
      rng.init((uint32_t)px, (uint32_t)py);
      vec3 final_colour = v3(0.f, 0.f, 0.f);
      for (int s = 0; s < spp; s++) {
        const float eu = rng.next();
        const float ev = rng.next();
        const float su = ((float)px + eu) / (float)W, sv = ((float)py + ev) / (float)H;
        vec3 org = cpos;
        vec3 dir = normalize(d00 + su * du + sv * dv);
        /* tracePath (deviceCode.cu:165-189) */
        vec3 colour = v3(0.f, 0.f, 0.f), atten = v3(1.f, 1.f, 1.f);
        for (int d = 0; d < depth; d++) {
          HitRecord rec;
          vec3 c;
          const bool hit = ray_colour(R, org, dir, rng, &rec, &c, heap.data());
          colour = colour + c * atten;
          if (!hit) break; /* quirk (iii) */
          bool absorbed; float coef;
          const vec3 out_dir = reflect_or_refract(rec.material, dir, rec.normal, rng, &absorbed, &coef);
          if (absorbed) break;
          atten = atten * (coef * v3(rec.material.albedo));
          org = rec.hitpoint; dir = out_dir;
        }
        final_colour = final_colour + colour;
      }
      final_colour = final_colour * (1.f / (float)spp);
      if (rgb_out) { rgb_out[3 * p] = final_colour.x; rgb_out[3 * p + 1] = final_colour.y; rgb_out[3 * p + 2] = final_colour.z; }
      const int y = H - py;
      if (fb && y < H) fb[(size_t)px + (size_t)W * y] = orc_make_rgba(final_colour.x, final_colour.y, final_colour.z);
    }
  }
}

}  // extern "C"
