/* pm_render.cu — stage 3 driver: the camera pass.
 *
 * Replaces owlRayGenLaunch2D(simpleRayGen) (ray-tracer/src/hostCode.cu:231-237) and the programs of
 * ray-tracer/cuda/deviceCode.cu:25-281 + shading.h:20-91. Same arithmetic and RNG draw order as the
 * CPU restatement in the CPU oracle (render restatement) (which lists the positions taken on the reference's
 * undefined behaviour).
 *
 * The reference runs one monolithic thread per pixel: spp x depth x (1 primary + L shadow rays +
 * 1 caustic kNN + 20 final-gather rays each followed by a global kNN). Here one "wave" = one path
 * vertex of one sample for every pixel of a chunk, split into stages so that each stage is a wide,
 * homogeneous launch and all kNN queries of a wave (21 per pixel) go through the Morton-sorted
 * gather kernel together:
 *   k_vertex   thread / pixel      primary ray, direct light (shadow rays), caustic query,
 *                                  the 20 final-gather directions (serial RNG stream), next ray
 *   k_fg       thread / (pixel,j)  final-gather ray -> global-map query
 *   gather     pm::gather_device   caustic map (1 / pixel) and global map (20 / pixel)
 *   k_resolve  thread / pixel      combine, accumulate, advance the path
 */
#include "pm_scene.cuh"

namespace {

using namespace pmd;

constexpr int kK = 50;            /* K_NEAREST_NEIGHBOURS, shading.h:7 */
constexpr float kCutoff = 100.f;  /* K_MAX_DISTANCE, shading.h:8 */
constexpr int kFG = 20;           /* NUM_DIFFUSE_SAMPLES, deviceCode.cu:15 */
constexpr int kMaxLights = 64;

struct Lights {
  int n;
  float pos[kMaxLights][3];
  float rgb[kMaxLights][3];
  float power[kMaxLights];
};

struct Chunk {
  int64_t begin; int count; /* launch indices [begin, begin + count) */
  int W, H;
  uint32_t* rng;            /* LCG state per pixel */
  float4 *org, *dir, *atten, *colour, *final_c;
  uint8_t* alive;
  /* per-wave vertex record */
  float4 *direct, *albedo, *next_dir, *next_mul; /* next_mul.w: 1 = continue, 0 = absorbed; direct.w: 1 = vertex valid */
  float *cq_hit, *cq_brdf, *cq_out;
  float *gq_hit, *gq_brdf, *gq_alb, *gq_out;
};

__device__ __forceinline__ float4 f4(vec3 v, float w) { return make_float4(v.x, v.y, v.z, w); }
__device__ __forceinline__ vec3 xyz(float4 v) { return v3(v.x, v.y, v.z); }

__global__ void k_init(Chunk c) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.count) return;
  const int64_t p = c.begin + i;
  LCG4 r;
  r.init((uint32_t)(p % c.W), (uint32_t)(p / c.W)); /* deviceCode.cu:197 */
  c.rng[i] = r.state;
  c.final_c[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

__global__ void k_begin_sample(Chunk c, pm_camera cam) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.count) return;
  const int64_t p = c.begin + i;
  const int px = (int)(p % c.W), py = (int)(p / c.W);
  LCG4 r; r.state = c.rng[i];
  const float eu = r.next();
  const float ev = r.next();
  c.rng[i] = r.state;
  const float su = __fdiv_rn((float)px + eu, (float)c.W), sv = __fdiv_rn((float)py + ev, (float)c.H);
  const vec3 d = normalize(v3(cam.dir_00) + su * v3(cam.dir_du) + sv * v3(cam.dir_dv));
  c.org[i] = f4(v3(cam.pos), 0.f);
  c.dir[i] = f4(d, 0.f);
  c.atten[i] = make_float4(1.f, 1.f, 1.f, 0.f);
  c.colour[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  c.alive[i] = 1;
}

/* shading.h:80-91 */
__device__ __forceinline__ float specular_brdf(float spec, vec3 in_light, vec3 out_light, vec3 n) {
  const vec3 refl = reflect(in_light, n);
  if (nearZero(refl - out_light)) return spec;
  return 0.f;
}

/* shading.h:20-57 (RNG by value) */
__device__ __forceinline__ vec3 calculate_refracted(const pm_material& m, vec3 ray_dir, vec3 normal, LCG4 rand) {
  vec3 outward, refracted = v3(0.f, 0.f, 0.f);
  float ni_over_nt, coef, cosine;
  if (dot(ray_dir, normal) > 0.f) {
    outward = -normal;
    ni_over_nt = m.refraction_idx;
    cosine = dot(ray_dir, normal);
    cosine = __fsqrt_rn(1.f - m.refraction_idx * m.refraction_idx * (1.f - cosine * cosine));
  } else {
    outward = normal;
    ni_over_nt = __fdiv_rn(1.f, m.refraction_idx);
    cosine = -dot(ray_dir, normal);
  }
  if (refract4(ray_dir, outward, ni_over_nt, &refracted)) coef = schlick(cosine, m.refraction_idx);
  else coef = 1.f;
  if (rand.next() < coef) return reflect(ray_dir, normal);
  return refracted;
}

__global__ void __launch_bounds__(128)
k_vertex(Chunk c, SceneView sc, Lights L, float sky_r, float sky_g, float sky_b) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.count) return;
  const size_t g0 = (size_t)i * kFG;
  c.direct[i] = make_float4(0.f, 0.f, 0.f, 0.f); /* vertex invalid until proven otherwise */
  c.cq_brdf[i] = -1.f;
  if (!c.alive[i]) {
    for (int s = 0; s < kFG; s++) c.gq_brdf[g0 + s] = -1.f;
    return;
  }
  const vec3 org = xyz(c.org[i]), dir = xyz(c.dir[i]);
  const Hit h = trace<false>(sc, org, dir, PMD_EPS, PMD_INFTY);
  if (h.prim < 0) {
    /* miss program: colour = sky; tracePath adds it and (quirk iii) the path ends */
    const vec3 a = xyz(c.atten[i]);
    const vec3 col = xyz(c.colour[i]) + v3(sky_r, sky_g, sky_b) * a;
    c.colour[i] = f4(col, 0.f);
    c.alive[i] = 0;
    for (int s = 0; s < kFG; s++) c.gq_brdf[g0 + s] = -1.f;
    return;
  }
  /* closestHit (deviceCode.cu:233-254) */
  const pm_material m = sc.mats[slot_mesh(sc, h.slot)];
  const vec3 hp = org + dir * h.t;
  const vec3 gn = slot_normal(sc, h.slot);
  const vec3 n = normalize((dot(dir, gn) < 0.f) ? gn : -gn);
  const vec3 albedo = v3(m.albedo);
  const float diffuse_brdf = __fdiv_rn(m.diffuse, PMD_PI);

  /* direct light (deviceCode.cu:54-99) */
  vec3 direct = v3(0.f, 0.f, 0.f);
  for (int l = 0; l < L.n; l++) {
    vec3 light_dir = v3(L.pos[l]) - hp;
    const float dist = norm(light_dir);
    light_dir = normalize(light_dir);
    const float ldn = dot(light_dir, n);
    if (ldn < 0.f) continue;
    const float vis = trace<true>(sc, hp, light_dir, PMD_EPS, dist * (1.f - PMD_EPS)).prim >= 0 ? 0.f : 1.f;
    const float sb = specular_brdf(m.specular, light_dir, dir, n);
    vec3 t = v3(vis, vis, vis) * L.power[l];
    t = t * ldn;
    t = t * __fdiv_rn(1.f, dist * dist);
    t = t * (diffuse_brdf + sb);
    t = t * v3(L.rgb[l]);
    direct = direct + t;
  }
  c.direct[i] = f4(albedo * direct, 1.f);
  c.albedo[i] = f4(albedo, 0.f);
  c.cq_hit[3 * (size_t)i + 0] = hp.x; c.cq_hit[3 * (size_t)i + 1] = hp.y; c.cq_hit[3 * (size_t)i + 2] = hp.z;
  c.cq_brdf[i] = diffuse_brdf;

  /* final-gather directions (deviceCode.cu:110-121): one serial RNG stream per pixel */
  LCG4 rng; rng.state = c.rng[i];
  const vec3 nn = normalize(n);
  for (int s = 0; s < kFG; s++) {
    if (!(diffuse_brdf > 0.f)) { c.gq_brdf[g0 + s] = -1.f; continue; }
    vec3 rdir;
    do {
      rdir = nn + randomUnitVector(rng);
    } while (nearZero(rdir));
    rdir = normalize(rdir);
    c.gq_hit[3 * (g0 + s) + 0] = rdir.x; c.gq_hit[3 * (g0 + s) + 1] = rdir.y; c.gq_hit[3 * (g0 + s) + 2] = rdir.z;
    c.gq_brdf[g0 + s] = 1.f; /* pending: k_fg replaces it */
  }
  c.rng[i] = rng.state;

  /* reflect_or_refract_ray (shading.h:59-78) on a COPY of the RNG (quirk ii) */
  LCG4 rand = rng;
  const float r = rand.next();
  if (r < m.specular) {
    c.next_dir[i] = f4(reflect(dir, n), 0.f);
    c.next_mul[i] = f4(m.specular * albedo, 1.f);
  } else if (r < m.specular + m.transmission) {
    c.next_dir[i] = f4(calculate_refracted(m, dir, n, rand), 0.f);
    c.next_mul[i] = f4(m.transmission * albedo, 1.f);
  } else {
    c.next_dir[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    c.next_mul[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

/* WARP_VOTE: the final-gather rays of a large scene are traced with the vote-driven traversal (pm_scene.cuh): every
 * lane of the warp stays until the trace is done, lanes without a ray included */
template <bool WARP_VOTE>
__global__ void __launch_bounds__(128)
k_fg(Chunk c, SceneView sc) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = t < (int64_t)c.count * kFG && c.gq_brdf[t] > 0.f;
  if (!WARP_VOTE && !valid) return;
  const int i = valid ? (int)(t / kFG) : 0;
  const vec3 org = valid ? v3(c.cq_hit + 3 * (size_t)i) : v3(0.f, 0.f, 0.f);
  const vec3 dir = valid ? v3(c.gq_hit + 3 * (size_t)t) : v3(0.f, 0.f, 1.f);
  const Hit h = WARP_VOTE ? trace_warp<false>(sc, valid, org, dir, 3 * PMD_EPS, PMD_INFTY) : trace<false>(sc, org, dir, 3 * PMD_EPS, PMD_INFTY);
  if (!valid) return;
  if (h.prim < 0) { c.gq_brdf[t] = -1.f; return; } /* quirk iii: a missed sample contributes nothing */
  const pm_material m = sc.mats[slot_mesh(sc, h.slot)];
  if (!(m.diffuse > 0.f)) { c.gq_brdf[t] = -1.f; return; }
  const vec3 hp = org + dir * h.t;
  c.gq_hit[3 * (size_t)t + 0] = hp.x; c.gq_hit[3 * (size_t)t + 1] = hp.y; c.gq_hit[3 * (size_t)t + 2] = hp.z;
  c.gq_brdf[t] = __fdiv_rn(m.diffuse, PMD_PI);
  c.gq_alb[3 * (size_t)t + 0] = m.albedo[0]; c.gq_alb[3 * (size_t)t + 1] = m.albedo[1]; c.gq_alb[3 * (size_t)t + 2] = m.albedo[2];
}

__global__ void k_resolve(Chunk c) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.count) return;
  const float4 d4 = c.direct[i];
  if (d4.w == 0.f) return; /* dead pixel or primary miss */
  const vec3 albedo = xyz(c.albedo[i]);
  vec3 diffuse_term = v3(0.f, 0.f, 0.f);
  const size_t g0 = (size_t)i * kFG;
  for (int s = 0; s < kFG; s++) {
    if (!(c.gq_brdf[g0 + s] > 0.f)) continue;
    diffuse_term = diffuse_term + v3(c.gq_out + 3 * (g0 + s)) * v3(c.gq_alb + 3 * (g0 + s));
  }
  diffuse_term = diffuse_term / (float)kFG;
  diffuse_term = diffuse_term * albedo;
  const vec3 caustics = v3(c.cq_out + 3 * (size_t)i);
  const vec3 direct_term = xyz(d4);
  vec3 fc;
  fc.x = 0.2f * diffuse_term.x + 0.08f * caustics.x + 0.8f * direct_term.x;
  fc.y = 0.2f * diffuse_term.y + 0.08f * caustics.y + 0.8f * direct_term.y;
  fc.z = 0.2f * diffuse_term.z + 0.08f * caustics.z + 0.8f * direct_term.z;
  const vec3 atten = xyz(c.atten[i]);
  c.colour[i] = f4(xyz(c.colour[i]) + fc * atten, 0.f);
  const float4 mul = c.next_mul[i];
  if (mul.w == 0.f) { c.alive[i] = 0; return; } /* absorbed */
  c.atten[i] = f4(atten * xyz(mul), 0.f);
  c.org[i] = make_float4(c.cq_hit[3 * (size_t)i], c.cq_hit[3 * (size_t)i + 1], c.cq_hit[3 * (size_t)i + 2], 0.f);
  c.dir[i] = c.next_dir[i];
}

__global__ void k_end_sample(Chunk c) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.count) return;
  c.final_c[i] = f4(xyz(c.final_c[i]) + xyz(c.colour[i]), 0.f);
}

/* make_8bit / make_rgba (owl_device.h:81-93): int(f * 256.f) is cvt.rzi (NaN -> 0, saturating) */
__device__ __forceinline__ uint32_t make_8bit(float f) { return (uint32_t)min(255, max(0, __float2int_rz(f * 256.f))); }

__global__ void k_write(Chunk c, int spp, uint32_t* __restrict__ fb, float* __restrict__ rgb) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.count) return;
  const int64_t p = c.begin + i;
  const int px = (int)(p % c.W), py = (int)(p / c.W);
  const vec3 col = xyz(c.final_c[i]) * __fdiv_rn(1.f, (float)spp);
  if (rgb) { rgb[3 * p] = col.x; rgb[3 * p + 1] = col.y; rgb[3 * p + 2] = col.z; } /* by launch index px + W * py */
  const int y = c.H - py; /* deviceCode.cu:224-227; row H (py == 0) is out of bounds and is not written */
  if (y >= c.H) return;
  fb[(size_t)px + (size_t)c.W * y] = (make_8bit(col.x) << 0) + (make_8bit(col.y) << 8) + (make_8bit(col.z) << 16) + (0xffu << 24);
}

}  // namespace

extern "C" {

int pm_render(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights, int n_lights, const pm_map* global_map,
              const pm_map* caustic_map, const pm_camera* camera, int width, int height, int spp, int depth,
              const float sky[3], int64_t pixel_begin, int64_t pixel_end, uint32_t* fb) {
  return pm_render_rgb(ctx, scene, lights, n_lights, global_map, caustic_map, camera, width, height, spp, depth, sky,
                       pixel_begin, pixel_end, fb, nullptr);
}

int pm_render_rgb(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights, int n_lights, const pm_map* global_map,
                  const pm_map* caustic_map, const pm_camera* camera, int width, int height, int spp, int depth,
                  const float sky[3], int64_t pixel_begin, int64_t pixel_end, uint32_t* fb, float* rgb_linear) {
  PM_REQUIRE(ctx && scene && camera && sky && fb, "ctx/scene/camera/sky/fb is NULL");
  PM_REQUIRE((!global_map || !global_map->cells_only) && (!caustic_map || !caustic_map->cells_only),
             "maps built by pm_map_build_cells carry no kd-tree: fixed-radius gathers only");
  PM_REQUIRE(width > 0 && height > 0 && (int64_t)width * height < ((int64_t)1 << 31), "bad framebuffer size");
  PM_REQUIRE(spp >= 1 && depth >= 0, "bad spp/depth");
  PM_REQUIRE(n_lights >= 0 && n_lights <= kMaxLights && (n_lights == 0 || lights), "bad lights");
  PM_REQUIRE(pixel_begin >= 0 && pixel_begin <= pixel_end && pixel_end <= (int64_t)width * height, "bad pixel range");
  PM_REQUIRE(scene->nt == 0 || scene->mats, "scene has triangles but no materials");
  PM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  Lights L;
  L.n = n_lights;
  for (int l = 0; l < n_lights; l++) {
    for (int d = 0; d < 3; d++) { L.pos[l][d] = lights[l].pos[d]; L.rgb[l][d] = lights[l].rgb[d]; }
    L.power[l] = (float)lights[l].power; /* static_cast<float>(current_light.power), deviceCode.cu:92 */
  }
  const size_t fb_bytes = sizeof(uint32_t) * (size_t)width * height;
  const bool fb_dev = pm::is_device_ptr(fb);
  uint32_t* fb_d = fb;
  if (!fb_dev) {
    PM_CHECK(ctx->get_scratch(pm::S_STAGE_OUT0, (size_t)width * height, &fb_d));
    PM_CUDA(cudaMemcpyAsync(fb_d, fb, fb_bytes, cudaMemcpyHostToDevice, st)); /* keep what other tiles wrote */
  }
  const size_t rgb_bytes = sizeof(float) * 3 * (size_t)width * height;
  const bool rgb_dev = !rgb_linear || pm::is_device_ptr(rgb_linear);
  float* rgb_d = rgb_linear;
  if (!rgb_dev) {
    PM_CHECK(ctx->get_scratch(pm::S_STAGE_OUT1, (size_t)3 * width * height, &rgb_d));
    PM_CUDA(cudaMemcpyAsync(rgb_d, rgb_linear, rgb_bytes, cudaMemcpyHostToDevice, st)); /* keep other tiles' pixels */
  }
  const int64_t total = pixel_end - pixel_begin;
  const int64_t chunk_max = (int64_t)1 << 21;
  const int cap = (int)(total < chunk_max ? total : chunk_max);
  float ms_gather = 0.f;
  if (cap > 0) {
    Chunk c;
    c.W = width; c.H = height;
    PM_CHECK(ctx->get_scratch(pm::S_RENDER0, (size_t)cap, &c.rng));
    float4* f4buf;
    PM_CHECK(ctx->get_scratch(pm::S_RENDER1, (size_t)cap * 9, &f4buf));
    c.org = f4buf; c.dir = f4buf + (size_t)cap; c.atten = f4buf + 2 * (size_t)cap; c.colour = f4buf + 3 * (size_t)cap;
    c.final_c = f4buf + 4 * (size_t)cap; c.direct = f4buf + 5 * (size_t)cap; c.albedo = f4buf + 6 * (size_t)cap;
    c.next_dir = f4buf + 7 * (size_t)cap; c.next_mul = f4buf + 8 * (size_t)cap;
    PM_CHECK(ctx->get_scratch(pm::S_RENDER2, (size_t)cap, &c.alive));
    float* cq;
    PM_CHECK(ctx->get_scratch(pm::S_RENDER3, (size_t)cap * 7, &cq));
    c.cq_hit = cq; c.cq_brdf = cq + 3 * (size_t)cap; c.cq_out = cq + 4 * (size_t)cap;
    float* gq;
    PM_CHECK(ctx->get_scratch(pm::S_RENDER4, (size_t)cap * kFG * 10, &gq));
    c.gq_hit = gq; c.gq_brdf = gq + 3 * (size_t)cap * kFG; c.gq_alb = gq + 4 * (size_t)cap * kFG; c.gq_out = gq + 7 * (size_t)cap * kFG;
    const SceneView sv = scene->view();
    for (int64_t b = pixel_begin; b < pixel_end; b += cap) {
      c.begin = b;
      c.count = (int)(pixel_end - b < cap ? pixel_end - b : cap);
      const unsigned nb = (unsigned)pm::div_up(c.count, 128);
      k_init<<<nb, 128, 0, st>>>(c);
      ctx->launches++;
      for (int s = 0; s < spp; s++) {
        k_begin_sample<<<nb, 128, 0, st>>>(c, *camera);
        ctx->launches++;
        for (int d = 0; d < depth; d++) {
          k_vertex<<<nb, 128, 0, st>>>(c, sv, L, sky[0], sky[1], sky[2]);
          if (scene->nt > 4096) k_fg<true><<<(unsigned)pm::div_up((int64_t)c.count * kFG, 128), 128, 0, st>>>(c, sv);
          else k_fg<false><<<(unsigned)pm::div_up((int64_t)c.count * kFG, 128), 128, 0, st>>>(c, sv);
          ctx->launches += 2;
          PM_CUDA(cudaGetLastError());
          PM_CHECK(pm::gather_device(ctx, caustic_map, c.cq_hit, c.cq_brdf, c.count, kK, kCutoff, c.cq_out, false));
          PM_CHECK(pm::gather_device(ctx, global_map, c.gq_hit, c.gq_brdf, (int64_t)c.count * kFG, kK, kCutoff, c.gq_out, true));
          k_resolve<<<nb, 128, 0, st>>>(c);
          ctx->launches++;
          if (global_map && global_map->n >= 1) {
            PM_CUDA(cudaStreamSynchronize(st));
            float ms = 0.f;
            PM_CUDA(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
            ms_gather += ms;
          }
        }
        k_end_sample<<<nb, 128, 0, st>>>(c);
        ctx->launches++;
      }
      k_write<<<nb, 128, 0, st>>>(c, spp, fb_d, rgb_d);
      ctx->launches++;
      PM_CUDA(cudaGetLastError());
    }
  }
  if (!fb_dev) PM_CUDA(cudaMemcpyAsync(fb, fb_d, fb_bytes, cudaMemcpyDeviceToHost, st));
  if (!rgb_dev) PM_CUDA(cudaMemcpyAsync(rgb_linear, rgb_d, rgb_bytes, cudaMemcpyDeviceToHost, st));
  PM_CUDA(cudaStreamSynchronize(st));
  ctx->last_kernel_ms = ms_gather; /* device time of the global-map gather kernels of this call */
  return PM_OK;
}

}  // extern "C"
