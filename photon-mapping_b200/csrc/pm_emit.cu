/* pm_emit.cu — stage 1: photon emission and the multi-bounce walk.
 *
 * Replaces owlRayGenLaunch2D(pointLightRayGen) and its closest-hit/miss programs
 * (photon-mapping/cuda/deviceCode.cu:10-136; launch sizing photon-mapping/src/hostCode.cu:72-90,
 * 102-110). Same walk, same store rules, same RNG streams (PM_RNG_LCG_PARITY); the traversal is the
 * software BVH of pm_scene.cuh.
 *
 * B200 design:
 *   - persistent warps; a lane whose path ended (miss / absorbed / depth) immediately claims the next
 *     path id with a ballot-aggregated atomic, so warps stay full instead of idling on the longest
 *     path of a 32-path group (the reference runs one path per thread to completion);
 *   - stored photons go to a slot buffer in HBM indexed [path][k-th store]; a scan of the per-path
 *     counts and one compaction pass produce the output in the deterministic order
 *     (light, launch index, bounce) — the reference appends through one global atomic that lives in
 *     host-pinned memory (hostCode.cu:92-99), i.e. nondeterministic order and a PCIe round trip per
 *     photon;
 *   - paths are processed in batches so the slot buffer stays bounded.
 */
#include <cub/cub.cuh>
#include "pm_scene.cuh"

namespace {

using namespace pmd;

constexpr int kMaxLights = 64;

struct EmitParams {
  SceneView sc;
  int n_lights;
  int64_t light_first[kMaxLights + 1]; /* prefix of this rank's path counts per light */
  float light_pos[kMaxLights][3];
  float light_rgb[kMaxLights][3];
  float light_shape[kMaxLights][4]; /* normal, side length; side <= 0: point light */
  int max_depth, caustics, rng_mode;
  uint64_t seed;
  int64_t first_photon, photon_stride;
  int64_t batch_begin, batch_count; /* global path numbers [batch_begin, batch_begin + batch_count) */
  int slots_per_path;               /* max_depth - 1 */
  int min_lanes;
};

enum { EV_MISS = 0, EV_ABSORBED = 1, EV_DIFFUSE = 2, EV_SPECULAR = 4, EV_REFRACT = 8 };

__device__ __forceinline__ void store40(pm_photon40* dst, vec3 pos, vec3 dir, vec3 color) {
  float* w = reinterpret_cast<float*>(dst);
  /* 40-byte records are 8-byte aligned: five 8-byte stores */
  reinterpret_cast<float2*>(w)[0] = make_float2(pos.x, pos.y);
  reinterpret_cast<float2*>(w)[1] = make_float2(pos.z, dir.x);
  reinterpret_cast<float2*>(w)[2] = make_float2(dir.y, dir.z);
  reinterpret_cast<float2*>(w)[3] = make_float2(__int_as_float(0), color.x);
  reinterpret_cast<float2*>(w)[4] = make_float2(color.y, color.z);
}

/* fewer traversing lanes than this: leave the traversal loop so that finished lanes can shade and pick up their
 * next segment (or a new path) instead of idling until the slowest ray of the warp is done */
constexpr int kVoteMinTris = 4096;  /* scenes above this (the Morton-built hierarchies) walk with the vote-driven traversal */
constexpr int kMinTraceLanes = 28; /* swept 0..31 on B200: 24-31 within 1 %, 0 (never leave) 8-25 % slower */

/* WARP_VOTE: the vote-driven traversal of pm_scene.cuh (large scenes: twice the lanes per instruction on the
 * 1 M-triangle room, 49 -> 43 ms per 16 M paths) instead of the free-running while-while loop (small scenes: the Cornell
 * box walks 1.7x faster with it — its rays are a handful of steps long and the votes cost more than the idle lanes) */
template <bool WARP_VOTE>
__global__ void __launch_bounds__(128)
k_walk(const EmitParams P, unsigned long long* __restrict__ next_path, pm_photon40* __restrict__ slots,
       int* __restrict__ counts, unsigned long long* __restrict__ segments) {
  const unsigned lane = threadIdx.x & 31u;
  bool active = false; /* this lane owns a live path; its current segment is traced in `st` */
  int64_t local = 0;   /* index inside the batch */
  int depth = 0, stored = 0;
  vec3 org = v3(0, 0, 0), dir = v3(0, 0, 1), color = v3(0, 0, 0);
  Rng rng;
  rng.mode = P.rng_mode; rng.lcg.state = 0; rng.cr.key = 0; rng.cr.ctr = 0;
  unsigned long long my_segments = 0;
  bool exhausted = false;
  TraceState st;
  st.node = kNoChild; st.sp = 0; st.pend = 0;
  while (true) {
    /* 1. lanes whose traversal has finished: optixTrace returned -> triangleMeshClosestHit / miss
     *    (deviceCode.cu:74-136), then shootPhoton / shootCausticsPhoton (deviceCode.cu:25-52) */
    if (active && st.node == kNoChild && st.pend == 0) {
      const Hit h = st.best;
      int ev = EV_MISS;
      vec3 s_org = org, s_dir = dir, s_col = color;
      if (h.prim >= 0) {
        const pm_material m = P.sc.mats[slot_mesh(P.sc, h.slot)];
        const float diffuseProb = m.diffuse;
        const float specularProb = m.specular + diffuseProb;
        const float transmissionProb = m.transmission + specularProb;
        const float r = rng.next();
        if (r < diffuseProb) ev = EV_DIFFUSE;
        else if (r < specularProb) ev = EV_SPECULAR;
        else if (r < transmissionProb) ev = EV_REFRACT;
        else ev = EV_ABSORBED;
        if (ev != EV_ABSORBED) {
          s_org = org + h.t * dir;
          const vec3 n = slot_normal(P.sc, h.slot);
          if (ev == EV_DIFFUSE) s_dir = cosineSampleHemisphere(n, rng);
          else if (ev == EV_SPECULAR) s_dir = reflect(dir, n);
          else s_dir = refract3(dir, n, m.refraction_idx);
          s_col = v3(m.albedo) * color;
        }
      }
      if (ev == EV_DIFFUSE && depth > 0) {
        store40(slots + (size_t)local * P.slots_per_path + stored, s_org, s_dir, color);
        stored++;
      }
      const bool cont = P.caustics ? ((ev & (EV_SPECULAR | EV_REFRACT)) != 0) : (ev == EV_DIFFUSE);
      depth++;
      if (!cont || depth >= P.max_depth) {
        counts[local] = stored;
        active = false;
      } else {
        org = s_org; dir = s_dir; color = s_col;
        my_segments++;
        trace_begin(P.sc, st, dir, 1e30f);
      }
    }
    /* 2. refill: lanes without a path claim consecutive ids */
    const unsigned need = __ballot_sync(0xffffffffu, !active && !exhausted);
    if (need) {
      unsigned long long base = 0;
      const int leader = __ffs(need) - 1;
      if ((int)lane == leader) base = atomicAdd(next_path, (unsigned long long)__popc(need));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (!active && !exhausted) {
        const int64_t mine = (int64_t)base + __popc(need & ((1u << lane) - 1u));
        if (mine >= P.batch_count) {
          exhausted = true;
        } else {
          local = mine;
          const int64_t g = P.batch_begin + mine;
          int l = 0;
          while (l + 1 < P.n_lights && g >= P.light_first[l + 1]) l++;
          const int64_t idx = P.first_photon + (g - P.light_first[l]) * P.photon_stride;
          if (P.rng_mode == PM_RNG_LCG_PARITY) rng.init_lcg((uint32_t)idx, 0u); /* deviceCode.cu:59 */
          else rng.init_counter(P.seed, ((uint64_t)((P.caustics ? P.n_lights : 0) + l) << 32) | (uint64_t)idx);
          color = v3(P.light_rgb[l]);
          org = v3(P.light_pos[l]);
          if (P.light_shape[l][3] > 0.f) sampleSquareLight(org, v3(P.light_shape[l]), P.light_shape[l][3], rng, &org, &dir);
          else dir = randomPointInUnitSphere(rng); /* deviceCode.cu:64 */
          depth = 0; stored = 0;
          active = P.max_depth > 0;
          if (!active) counts[local] = 0;
          else { my_segments++; trace_begin(P.sc, st, dir, 1e30f); }
        }
      }
    }
    if (!__any_sync(0xffffffffu, active)) break;
    /* 3. traversal of the segments in flight (lanes without one have st.node == kNoChild and pass through) */
    if (WARP_VOTE) trace_advance_warp<false>(P.sc, st, org, dir, PMD_EPS, 1e30f, P.min_lanes);
    else if (active) trace_advance<false>(P.sc, st, org, dir, PMD_EPS, 1e30f, P.min_lanes);
  }
  /* one atomic per warp */
  for (int o = 16; o > 0; o >>= 1) my_segments += __shfl_xor_sync(0xffffffffu, my_segments, o);
  if (lane == 0 && my_segments) atomicAdd(segments, my_segments);
}

/* path p copies its counts[p] records to dst[offsets[p] - base ...) while below `limit` records */
/* key of a stored photon: light << 56 | launch index << 8 | ordinal of the store on its path — sorting the
 * shards of several ranks by it restores the single-GPU output order */
struct KeyInfo {
  uint64_t* keys; /* NULL: no keys wanted */
  int n_lights;
  int64_t light_first[kMaxLights + 1];
  int64_t first_photon, photon_stride, batch_begin;
};

__global__ void k_compact(int64_t n_paths, int slots_per_path, const pm_photon40* __restrict__ slots,
                          const int* __restrict__ counts, const int64_t* __restrict__ offsets, int64_t limit,
                          pm_photon40* __restrict__ dst, const KeyInfo ki) {
  /* 5 threads per path-record word pair: thread t handles 8-byte word (t % 5) of record stream */
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t p = t / 5;
  const int w = (int)(t - p * 5);
  if (p >= n_paths) return;
  const int c = counts[p];
  const int64_t o = offsets[p];
  const float2* src = reinterpret_cast<const float2*>(slots + (size_t)p * slots_per_path);
  float2* out = reinterpret_cast<float2*>(dst);
  int l = 0;
  uint64_t idx = 0;
  if (ki.keys && w == 0) {
    const int64_t g = ki.batch_begin + p;
    while (l + 1 < ki.n_lights && g >= ki.light_first[l + 1]) l++;
    idx = (uint64_t)(ki.first_photon + (g - ki.light_first[l]) * ki.photon_stride);
  }
  for (int k = 0; k < c; k++) {
    if (o + k >= limit) break;
    out[(size_t)(o + k) * 5 + w] = src[(size_t)k * 5 + w];
    if (ki.keys && w == 0) ki.keys[o + k] = ((uint64_t)l << 56) | (idx << 8) | (uint64_t)k;
  }
}

/* per-path store counts of a batch as bytes (max_depth <= 64): what pm_exchange_photons gathers */
__global__ void k_counts_u8(int64_t n, const int* __restrict__ counts, uint8_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (uint8_t)counts[i];
}

struct CountToI64 {
  const int* c;
  __device__ int64_t operator()(int64_t i) const { return (int64_t)c[i]; }
};

__global__ void k_conv44(int64_t n, const pm_photon40* __restrict__ in, float power, pm_photon44* __restrict__ out) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t i = t / 11;
  const int w = (int)(t - i * 11);
  if (i >= n) return;
  const uint32_t* s = reinterpret_cast<const uint32_t*>(in) + (size_t)i * 10;
  uint32_t v;
  if (w < 6) v = s[w];                 /* pos, dir */
  else if (w < 9) v = s[w + 1];        /* color sits after the int power */
  else if (w == 9) v = __float_as_uint(power);
  else v = 0u;                         /* quantized_normal, split_dim */
  reinterpret_cast<uint32_t*>(out)[(size_t)i * 11 + w] = v;
}

}  // namespace

extern "C" {

static int emit_impl(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights, const pm_light_shape* shapes, int n_lights, int64_t casted,
                     int max_depth, int caustics_mode, uint64_t seed, int rng_mode, int64_t first_photon,
                     int64_t photon_stride, pm_photon40* out, uint64_t* out_keys, int64_t capacity, int64_t* count,
                     int64_t* n_paths, int64_t* n_segments, uint8_t* out_path_counts = nullptr) {
  PM_REQUIRE(ctx && scene && count, "ctx/scene/count is NULL");
  PM_REQUIRE(!out_path_counts || pm::is_device_ptr(out_path_counts), "path_counts must be device memory");
  PM_REQUIRE(n_lights >= 0 && n_lights <= kMaxLights, "n_lights out of range (max 64)");
  PM_REQUIRE(n_lights == 0 || lights, "lights is NULL");
  PM_REQUIRE(casted >= 0 && casted <= 0x7fffffff, "casted out of range (int in the reference)");
  PM_REQUIRE(max_depth >= 0 && max_depth <= 64, "max_depth out of range");
  PM_REQUIRE(first_photon >= 0 && photon_stride >= 1, "first_photon/photon_stride invalid");
  PM_REQUIRE(capacity >= 0 && (capacity == 0 || out), "out is NULL");
  PM_REQUIRE(rng_mode == PM_RNG_LCG_PARITY || rng_mode == PM_RNG_COUNTER, "unknown rng_mode");
  PM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  *count = 0;
  if (n_paths) *n_paths = 0;
  if (n_segments) *n_segments = 0;

  EmitParams P;
  memset(&P, 0, sizeof P);
  P.sc = scene->view();
  P.n_lights = n_lights;
  double total_power = 0;
  for (int i = 0; i < n_lights; i++) total_power += lights[i].power;
  /* hostCode.cu:102-110: int photonsPerWatt = casted / total (truncation); :86 int(power * perWatt) */
  const int per_watt = n_lights > 0 && total_power > 0 ? (int)((double)(int)casted / total_power) : 0;
  P.light_first[0] = 0;
  for (int l = 0; l < n_lights; l++) {
    const int64_t n = (int64_t)(int)(lights[l].power * per_watt);
    const int64_t mine = n > first_photon ? (n - first_photon + photon_stride - 1) / photon_stride : 0;
    P.light_first[l + 1] = P.light_first[l] + mine;
    for (int d = 0; d < 3; d++) { P.light_pos[l][d] = lights[l].pos[d]; P.light_rgb[l][d] = lights[l].rgb[d]; }
    if (shapes && shapes[l].type == PM_LIGHT_SQUARE) {
      PM_REQUIRE(shapes[l].side_length > 0.f, "square light needs side_length > 0");
      for (int d = 0; d < 3; d++) P.light_shape[l][d] = shapes[l].normal[d];
      P.light_shape[l][3] = shapes[l].side_length;
    } else {
      PM_REQUIRE(!shapes || shapes[l].type == PM_LIGHT_POINT, "unknown light type");
    }
  }
  const int64_t total_paths = P.light_first[n_lights];
  P.max_depth = max_depth; P.caustics = caustics_mode ? 1 : 0; P.rng_mode = rng_mode; P.seed = seed;
  P.first_photon = first_photon; P.photon_stride = photon_stride;
  P.slots_per_path = max_depth > 1 ? max_depth - 1 : 1;
  P.min_lanes = kMinTraceLanes;
  const bool vote = scene->nt > kVoteMinTris;
  if (n_paths) *n_paths = total_paths;
  if (total_paths == 0) return PM_OK;
  PM_REQUIRE(scene->nt == 0 || scene->mats, "scene has triangles but no materials");

  /* batch: slot buffer of at most ~1.5 GB */
  int64_t batch = (int64_t)(1500ll << 20) / ((int64_t)sizeof(pm_photon40) * P.slots_per_path);
  if (batch > total_paths) batch = total_paths;
  if (batch > (int64_t)1 << 30) batch = (int64_t)1 << 30;
  if (batch < 1) batch = 1;

  pm_photon40* slots; int* counts; int64_t* offsets; unsigned long long* ctrs;
  PM_CHECK(ctx->get_scratch(pm::S_BUILD0, (size_t)batch * P.slots_per_path, &slots));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD1, (size_t)batch + 1, &counts));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD2, (size_t)batch + 1, &offsets));
  PM_CHECK(ctx->get_scratch(pm::S_MISC0, (size_t)4, &ctrs));
  size_t scan_bytes = 0;
  {
    CountToI64 op{counts};
    cub::TransformInputIterator<int64_t, CountToI64, cub::CountingInputIterator<int64_t>> it(cub::CountingInputIterator<int64_t>(0), op);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, it, offsets, (int)batch + 1, st);
  }
  void* scan_tmp = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_CUB, scan_bytes, &scan_tmp));
  PM_CUDA(cudaMemsetAsync(ctrs, 0, sizeof(unsigned long long) * 4, st));

  const bool dev_out = pm::is_device_ptr(out);
  PM_REQUIRE(!out_keys || capacity == 0 || pm::is_device_ptr(out_keys) == dev_out, "out and out_keys must both be host or both be device memory");
  KeyInfo ki;
  memset(&ki, 0, sizeof ki);
  ki.n_lights = n_lights;
  for (int l = 0; l <= n_lights; l++) ki.light_first[l] = P.light_first[l];
  ki.first_photon = first_photon; ki.photon_stride = photon_stride;
  int blocks_per_sm = 0;
  if (vote) PM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, k_walk<true>, 128, 0));
  else PM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, k_walk<false>, 128, 0));
  if (blocks_per_sm < 1) blocks_per_sm = 1;
  int64_t written = 0; /* records produced so far (may exceed capacity) */
  float ms_total = 0.f;
  for (int64_t b0 = 0; b0 < total_paths; b0 += batch) {
    const int64_t nb = total_paths - b0 < batch ? total_paths - b0 : batch;
    P.batch_begin = b0; P.batch_count = nb;
    PM_CUDA(cudaMemsetAsync(ctrs, 0, sizeof(unsigned long long), st));
    int64_t grid = (int64_t)ctx->sm_count * blocks_per_sm;
    if (grid > pm::div_up(nb, 128)) grid = pm::div_up(nb, 128);
    PM_CUDA(cudaEventRecord(ctx->ev0, st));
    if (vote) k_walk<true><<<(unsigned)grid, 128, 0, st>>>(P, ctrs, slots, counts, ctrs + 1);
    else k_walk<false><<<(unsigned)grid, 128, 0, st>>>(P, ctrs, slots, counts, ctrs + 1);
    PM_CUDA(cudaEventRecord(ctx->ev1, st));
    PM_CUDA(cudaGetLastError());
    {
      CountToI64 op{counts};
      cub::TransformInputIterator<int64_t, CountToI64, cub::CountingInputIterator<int64_t>> it(cub::CountingInputIterator<int64_t>(0), op);
      /* nb + 1 outputs: offsets[nb] = batch total (the transform reads counts[nb]: scratch is larger) */
      PM_CUDA(cudaMemsetAsync(counts + nb, 0, sizeof(int), st));
      size_t tb = scan_bytes;
      PM_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, tb, it, offsets, (int)nb + 1, st));
    }
    if (out_path_counts) {
      k_counts_u8<<<(unsigned)pm::div_up(nb, 256), 256, 0, st>>>(nb, counts, out_path_counts + b0);
      ctx->launches++;
    }
    int64_t batch_total = 0;
    PM_CUDA(cudaMemcpyAsync(&batch_total, offsets + nb, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    PM_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    PM_CUDA(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ms_total += ms;
    ctx->launches += 3;
    const int64_t room = capacity > written ? capacity - written : 0;
    const int64_t take = batch_total < room ? batch_total : room;
    if (take > 0) {
      pm_photon40* dst = nullptr;
      uint64_t* kdst = nullptr;
      if (dev_out) { dst = out + written; kdst = out_keys ? out_keys + written : nullptr; }
      else {
        PM_CHECK(ctx->get_scratch(pm::S_BUILD3, (size_t)take, &dst));
        if (out_keys) PM_CHECK(ctx->get_scratch(pm::S_BUILD4, (size_t)take, &kdst));
      }
      ki.keys = kdst; ki.batch_begin = b0;
      k_compact<<<(unsigned)pm::div_up(nb * 5, 256), 256, 0, st>>>(nb, P.slots_per_path, slots, counts, offsets, take, dst, ki);
      ctx->launches++;
      PM_CUDA(cudaGetLastError());
      if (!dev_out) {
        PM_CUDA(cudaMemcpyAsync(out + written, dst, sizeof(pm_photon40) * (size_t)take, cudaMemcpyDeviceToHost, st));
        if (out_keys) PM_CUDA(cudaMemcpyAsync(out_keys + written, kdst, sizeof(uint64_t) * (size_t)take, cudaMemcpyDeviceToHost, st));
      }
      PM_CUDA(cudaStreamSynchronize(st));
    }
    written += batch_total;
  }
  unsigned long long segs = 0;
  PM_CUDA(cudaMemcpyAsync(&segs, ctrs + 1, sizeof segs, cudaMemcpyDeviceToHost, st));
  PM_CUDA(cudaStreamSynchronize(st));
  ctx->last_kernel_ms = ms_total;
  if (n_segments) *n_segments = (int64_t)segs;
  *count = written;
  if (written > capacity) {
    pm::set_error("pm_emit_photons: %lld photons stored but capacity is %lld", (long long)written, (long long)capacity);
    return PM_ERR_CAPACITY;
  }
  return PM_OK;
}

int pm_emit_photons(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights, int n_lights, int64_t casted,
                    int max_depth, int caustics_mode, uint64_t seed, int rng_mode, int64_t first_photon,
                    int64_t photon_stride, pm_photon40* out, int64_t capacity, int64_t* count, int64_t* n_paths,
                    int64_t* n_segments) {
  return emit_impl(ctx, scene, lights, nullptr, n_lights, casted, max_depth, caustics_mode, seed, rng_mode, first_photon,
                   photon_stride, out, nullptr, capacity, count, n_paths, n_segments);
}

int pm_emit_photons_keyed(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights, int n_lights, int64_t casted,
                          int max_depth, int caustics_mode, uint64_t seed, int rng_mode, int64_t first_photon,
                          int64_t photon_stride, pm_photon40* out, uint64_t* out_keys, int64_t capacity, int64_t* count,
                          int64_t* n_paths, int64_t* n_segments) {
  PM_REQUIRE(out_keys || capacity == 0, "out_keys is NULL");
  return emit_impl(ctx, scene, lights, nullptr, n_lights, casted, max_depth, caustics_mode, seed, rng_mode, first_photon,
                   photon_stride, out, out_keys, capacity, count, n_paths, n_segments);
}

int pm_emit_photons_counted(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights, const pm_light_shape* shapes,
                            int n_lights, int64_t casted, int max_depth, int caustics_mode, uint64_t seed, int rng_mode,
                            int64_t first_photon, int64_t photon_stride, pm_photon40* out_dev, uint8_t* path_counts_dev,
                            int64_t capacity, int64_t* count, int64_t* n_paths, int64_t* n_segments) {
  PM_REQUIRE(path_counts_dev, "path_counts is NULL");
  PM_REQUIRE(capacity == 0 || pm::is_device_ptr(out_dev), "out must be device memory");
  return emit_impl(ctx, scene, lights, shapes, n_lights, casted, max_depth, caustics_mode, seed, rng_mode, first_photon,
                   photon_stride, out_dev, nullptr, capacity, count, n_paths, n_segments, path_counts_dev);
}

int pm_emit_photons_shaped(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights, const pm_light_shape* shapes, int n_lights,
                           int64_t casted, int max_depth, int caustics_mode, uint64_t seed, int rng_mode, int64_t first_photon,
                           int64_t photon_stride, pm_photon40* out, uint64_t* out_keys, int64_t capacity, int64_t* count,
                           int64_t* n_paths, int64_t* n_segments) {
  return emit_impl(ctx, scene, lights, shapes, n_lights, casted, max_depth, caustics_mode, seed, rng_mode, first_photon,
                   photon_stride, out, out_keys, capacity, count, n_paths, n_segments);
}

int pm_photons_to_maps(pm_ctx* ctx, const pm_photon40* global_in, int64_t n_global, const pm_photon40* caustic_in,
                       int64_t n_caustic, pm_photon44* out_global, pm_photon44* out_caustic) {
  PM_REQUIRE(ctx, "ctx is NULL");
  PM_REQUIRE(n_global >= 0 && n_caustic >= 0, "negative count");
  PM_REQUIRE(n_global == 0 || global_in, "global_in is NULL");
  PM_REQUIRE(n_caustic == 0 || (caustic_in && out_caustic), "caustic_in/out_caustic is NULL");
  PM_REQUIRE(n_global + n_caustic == 0 || out_global, "out_global is NULL");
  PM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const void *g_dev = nullptr, *c_dev = nullptr;
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN0, global_in, sizeof(pm_photon40) * (size_t)n_global, &g_dev));
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN1, caustic_in, sizeof(pm_photon40) * (size_t)n_caustic, &c_dev));
  pm::OutStage og, oc;
  PM_CHECK(pm::stage_out(ctx, pm::S_STAGE_OUT0, out_global, sizeof(pm_photon44) * (size_t)(n_global + n_caustic), &og));
  PM_CHECK(pm::stage_out(ctx, pm::S_STAGE_OUT1, out_caustic, sizeof(pm_photon44) * (size_t)n_caustic, &oc));
  /* hostCode.cu:65-83: PHOTON_POWER 1.0 for the global photons, CAUSTICS_PHOTON_POWER 0.5 for the
   * caustic ones, which are appended to the global map as well */
  if (n_global > 0) {
    k_conv44<<<(unsigned)pm::div_up(n_global * 11, 256), 256, 0, st>>>(n_global, (const pm_photon40*)g_dev, 1.0f, (pm_photon44*)og.dev);
    ctx->launches++;
  }
  if (n_caustic > 0) {
    k_conv44<<<(unsigned)pm::div_up(n_caustic * 11, 256), 256, 0, st>>>(n_caustic, (const pm_photon40*)c_dev, 0.5f, (pm_photon44*)og.dev + n_global);
    k_conv44<<<(unsigned)pm::div_up(n_caustic * 11, 256), 256, 0, st>>>(n_caustic, (const pm_photon40*)c_dev, 0.5f, (pm_photon44*)oc.dev);
    ctx->launches += 2;
  }
  PM_CUDA(cudaGetLastError());
  PM_CHECK(pm::stage_out_finish(ctx, og));
  PM_CHECK(pm::stage_out_finish(ctx, oc));
  PM_CUDA(cudaStreamSynchronize(st));
  return PM_OK;
}

}  // extern "C"
