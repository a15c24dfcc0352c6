/* pm_knn.cu — stage 3: k-nearest-photon search and the fused radiance estimate.
 *
 * Replaces cukd::stackBased::knn<HeapCandidateList<k>> (externals/cudaKDTree/cukd/knn.h:298-364,
 * 536-549; traverse-default-stack-based.h:24-90) as instantiated by
 * ray-tracer/cuda/shading.h:11-18, and gatherPhotons (shading.h:93-121).
 *
 * Result contract: per query, the k smallest (floatBits(dist²), node id) pairs among
 * nodes with dist² < cutoff²; dist² = fma(dz,dz, fma(dx,dx, dy*dy)), d = node - query
 * (the contraction nvcc emits for cukd's dot(d,d) in the reference kernel).
 *
 * B200 design (v1):
 *   - queries are Morton-sorted (24-bit code inside the map's AABB, radix sort of 4-byte ids) so the
 *     32 lanes of a warp walk nearly the same root-to-leaf paths: node fetches coalesce into a few
 *     sectors and the traversal stays convergent;
 *   - the tree is traversed through a 16-byte (x, y, z, split_dim) copy — one LDG.128 per node
 *     instead of a 44-byte-stride AoS access;
 *   - the candidate heap lives in shared memory, laid out [slot][thread] so that every heap access
 *     of a warp hits 32 distinct banks no matter which slot each lane touches (the reference's
 *     per-thread uint64[k] array is dynamically indexed and therefore spills to local memory);
 *     the heap root is mirrored in a register so the common "reject" case costs one compare.
 */
#include <cub/cub.cuh>
#include <cfloat>
#include "pm_knn.cuh"

namespace {

using pmk::encode;

constexpr int kBlock = 128;

/* expands 10 bits into every third bit */
__device__ __forceinline__ uint32_t part1by2(uint32_t v) {
  v &= 0x3FFu;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}

struct Box6 { float lo[3]; float inv[3]; };

__global__ void k_morton(const float* __restrict__ q, const float* __restrict__ skip_if_not_positive, int nq, Box6 box,
                         uint32_t* __restrict__ keys, int* __restrict__ ids) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  ids[i] = i;
  /* queries that will not be evaluated (brdf <= 0) sort to the end so that whole warps retire at once */
  if (skip_if_not_positive && !(skip_if_not_positive[i] > 0.f)) { keys[i] = 0x00FFFFFFu; return; }
  uint32_t code = 0;
#pragma unroll
  for (int d = 0; d < 3; d++) {
    float t = (q[(size_t)i * 3 + d] - box.lo[d]) * box.inv[d];
    t = fminf(fmaxf(t, 0.f), 255.f);
    code |= part1by2((uint32_t)t) << d;
  }
  keys[i] = code;
}

/* Shared-memory max-heap of uint64 entries, one column per thread. */
template <int K>
struct SmemHeap {
  uint64_t* col;  /* &heap[0][threadIdx.x]; slot s is col[s * kBlock] */
  uint64_t top;   /* register mirror of slot 0 */

  __device__ __forceinline__ void init(uint64_t* base, float cutoff2) {
    col = base + threadIdx.x;
    top = encode(cutoff2, -1);
#pragma unroll 8
    for (int s = 0; s < K; s++) col[s * kBlock] = top;
  }
  __device__ __forceinline__ float max_dist2() const { return __uint_as_float((uint32_t)(top >> 32)); }
  /* cukd HeapCandidateList::push (knn.h:327-359): replace the root, sift down */
  __device__ __forceinline__ void push(float d2, int id) {
    const uint64_t e = encode(d2, id);
    if (e >= top) return;
    int pos = 0;
    uint64_t new_top = e;
    while (true) {
      const int c0 = 2 * pos + 1;
      if (c0 >= K) break;
      uint64_t v = col[c0 * kBlock];
      int c = c0;
      if (c0 + 1 < K) {
        const uint64_t v1 = col[(c0 + 1) * kBlock];
        if (v1 > v) { v = v1; c = c0 + 1; }
      }
      if (v < e) break;
      col[pos * kBlock] = v;
      if (pos == 0) new_top = v;
      pos = c;
    }
    col[pos * kBlock] = e;
    top = new_top;
  }
};

/* stack-based traversal, literal in its visiting order (traverse-default-stack-based.h:49-88) */
template <int K>
__device__ __forceinline__ void traverse(const float4* __restrict__ nodes, int n, float qx, float qy, float qz,
                                         SmemHeap<K>& heap) {
  struct Entry { int node; float d2; };
  Entry stack[32];
  int sp = 0;
  int curr = 0;
  float cull = heap.max_dist2();
  while (true) {
    while (curr < n) {
      const float4 nd = __ldg(nodes + curr);
      const int dim = __float_as_int(nd.w);
      const float dx = nd.x - qx, dy = nd.y - qy, dz = nd.z - qz;
      const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
      heap.push(d2, curr);
      cull = heap.max_dist2();
      const float dp = dim == 0 ? -dx : (dim == 1 ? -dy : -dz); /* query - node along the split axis */
      const bool left_close = dp < 0.f;
      const int l = 2 * curr + 1;
      const int close_child = left_close ? l : l + 1;
      const int far_child = left_close ? l + 1 : l;
      const float plane2 = __fmul_rn(dp, dp);
      if (plane2 < cull && far_child < n) { stack[sp].node = far_child; stack[sp].d2 = plane2; sp++; }
      curr = close_child;
    }
    bool found = false;
    while (sp > 0) {
      --sp;
      if (stack[sp].d2 >= cull) continue;
      curr = stack[sp].node;
      found = true;
      break;
    }
    if (!found) return;
  }
}

/* one thread per query; with `list` the queries are list[0 .. *list_count), else order[t] (or t) */
template <int K, bool GATHER>
__global__ void __launch_bounds__(kBlock)
k_kd(const float4* __restrict__ nodes, const float4* __restrict__ color4, int n, pmk::QueryIO io,
     const int* __restrict__ list, const int* __restrict__ list_count) {
  extern __shared__ uint64_t smem_heap[];
  const int total = list ? *list_count : io.nq;
  for (int t = blockIdx.x * kBlock + threadIdx.x; t < total; t += gridDim.x * kBlock) {
    const int qi = list ? list[t] : (io.order ? io.order[t] : t);
    float f = 0.f;
    if (GATHER) {
      f = io.brdf[qi];
      if (!(f > 0.f)) { /* flux = brdf * ... is +0 for brdf == 0; negative marks "no query" (render pipeline) */
        io.rgb[(size_t)qi * 3 + 0] = 0.f; io.rgb[(size_t)qi * 3 + 1] = 0.f; io.rgb[(size_t)qi * 3 + 2] = 0.f;
        continue;
      }
    }
    SmemHeap<K> heap;
    heap.init(smem_heap, io.cutoff2);
    const float qx = io.q[(size_t)qi * 3 + 0], qy = io.q[(size_t)qi * 3 + 1], qz = io.q[(size_t)qi * 3 + 2];
    traverse<K>(nodes, n, qx, qy, qz, heap);
    if (!GATHER) {
      if (io.r2) io.r2[qi] = heap.max_dist2();
      if (io.ids) {
        /* each thread streams its own row; K is even so rows are 8-byte aligned */
        int2* ids2 = reinterpret_cast<int2*>(io.ids + (size_t)qi * K);
        float2* d22 = io.d2 ? reinterpret_cast<float2*>(io.d2 + (size_t)qi * K) : nullptr;
#pragma unroll 4
        for (int s = 0; s < K; s += 2) {
          const uint64_t a = heap.col[s * kBlock], b = heap.col[(s + 1) * kBlock];
          ids2[s >> 1] = make_int2((int)(uint32_t)a, (int)(uint32_t)b);
          if (d22) d22[s >> 1] = make_float2(__uint_as_float((uint32_t)(a >> 32)), __uint_as_float((uint32_t)(b >> 32)));
        }
      }
    } else {
      /* gatherPhotons (shading.h:93-121); arithmetic as written there, one rounding per op */
      const float r2 = heap.max_dist2();
      const float rad = __fsqrt_rn(r2);
      float fr = 0.f, fg = 0.f, fb = 0.f;
      for (int s = 0; s < K; s++) {
        const int id = (int)(uint32_t)heap.col[s * kBlock];
        if (id < 0 || id > n) continue; /* sic, shading.h:105 */
        const float4 p = __ldg(nodes + id);
        const float4 c = __ldg(color4 + id);
        const float dx = p.x - qx, dy = p.y - qy, dz = p.z - qz;
        const float dist = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
        const float w = __fsub_rn(1.f, __fmul_rn(__fdiv_rn(dist, rad), 1.1f));
        const float sc = __fmul_rn(__fmul_rn(f, c.w), w);
        fr = __fadd_rn(fr, __fmul_rn(sc, c.x));
        fg = __fadd_rn(fg, __fmul_rn(sc, c.y));
        fb = __fadd_rn(fb, __fmul_rn(sc, c.z));
      }
      const float k1 = (1 - (2.f / 3.f) * (1.f / 1.1f));
      const float denom = __fmul_rn(__fmul_rn(__fmul_rn(k1, 2.f), 3.14159274101257324f), r2);
      io.rgb[(size_t)qi * 3 + 0] = __fdiv_rn(fr, denom);
      io.rgb[(size_t)qi * 3 + 1] = __fdiv_rn(fg, denom);
      io.rgb[(size_t)qi * 3 + 2] = __fdiv_rn(fb, denom);
    }
    __syncwarp();
  }
}

/* Morton order of the queries inside the map's box → permutation in scratch slot S_PERM */
int sort_queries(pm_ctx* ctx, const pm_map* map, const float* q_dev, const float* brdf_dev, int nq, int** perm_out) {
  uint32_t *keys_in, *keys_out; int *ids_in, *perm;
  PM_CHECK(ctx->get_scratch(pm::S_KEYS0, (size_t)nq, &keys_in));
  PM_CHECK(ctx->get_scratch(pm::S_KEYS1, (size_t)nq, &keys_out));
  PM_CHECK(ctx->get_scratch(pm::S_VALS0, (size_t)nq, &ids_in));
  PM_CHECK(ctx->get_scratch(pm::S_PERM, (size_t)nq, &perm));
  Box6 box;
  for (int d = 0; d < 3; d++) {
    box.lo[d] = map->bounds[d];
    const float w = map->bounds[3 + d] - map->bounds[d];
    box.inv[d] = w > 0.f ? 256.f / w : 0.f;
  }
  k_morton<<<(unsigned)pm::div_up(nq, 256), 256, 0, ctx->stream>>>(q_dev, brdf_dev, nq, box, keys_in, ids_in);
  size_t tb = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tb, keys_in, keys_out, ids_in, perm, nq, 0, 24, ctx->stream);
  void* tmp = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_CUB, tb, &tmp));
  PM_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, keys_in, keys_out, ids_in, perm, nq, 0, 24, ctx->stream));
  ctx->launches += 6; /* k_morton + histogram + scan + 3 onesweep passes */
  *perm_out = perm;
  return PM_OK;
}

template <int K, bool GATHER>
int launch_kd(pm_ctx* ctx, const pm_map* map, const pmk::QueryIO& io, const int* list, const int* list_count) {
  const size_t smem = sizeof(uint64_t) * K * kBlock;
  PM_CUDA(cudaFuncSetAttribute(k_kd<K, GATHER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  /* a list of unknown (device-side) length is served by a fixed, persistent grid */
  int64_t grid = pm::div_up(io.nq, kBlock);
  const int64_t cap = (int64_t)ctx->sm_count * (list ? 2 : 64);
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  k_kd<K, GATHER><<<(unsigned)grid, kBlock, smem, ctx->stream>>>(map->nodes4, map->color4, (int)map->n, io, list, list_count);
  ctx->launches++;
  PM_CUDA(cudaGetLastError());
  return PM_OK;
}

__global__ void k_fill_empty(int64_t total, int32_t* ids, float* d2, float cutoff2) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  if (ids) ids[i] = -1;
  if (d2) d2[i] = cutoff2;
}
__global__ void k_fill_f(int64_t total, float* p, float v) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < total) p[i] = v;
}

/* bounds of a tree-ordered map: ordered-uint atomics, one per block and axis */
__device__ __forceinline__ uint32_t f2ord(float f) { const uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__global__ void k_bounds(const float4* __restrict__ nodes, int n, uint32_t* __restrict__ out /*lo xyz, hi xyz (ordered)*/) {
  uint32_t lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 p = __ldg(nodes + i);
    const uint32_t o[3] = {f2ord(p.x), f2ord(p.y), f2ord(p.z)};
#pragma unroll
    for (int d = 0; d < 3; d++) { lo[d] = min(lo[d], o[d]); hi[d] = max(hi[d], o[d]); }
  }
#pragma unroll
  for (int d = 0; d < 3; d++) {
    lo[d] = __reduce_min_sync(0xffffffffu, lo[d]);
    hi[d] = __reduce_max_sync(0xffffffffu, hi[d]);
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int d = 0; d < 3; d++) { atomicMin(out + d, lo[d]); atomicMax(out + 3 + d, hi[d]); }
  }
}

template <bool GATHER>
int dispatch_kd(pm_ctx* ctx, const pm_map* map, const pmk::QueryIO& io, const int* list, const int* list_count) {
  switch (io.k) {
    case 8: return launch_kd<8, GATHER>(ctx, map, io, list, list_count);
    case 16: return launch_kd<16, GATHER>(ctx, map, io, list, list_count);
    case 32: return launch_kd<32, GATHER>(ctx, map, io, list, list_count);
    case 50: return launch_kd<50, GATHER>(ctx, map, io, list, list_count);
    case 64: return launch_kd<64, GATHER>(ctx, map, io, list, list_count);
    default: pm::set_error("k must be one of 8, 16, 32, 50, 64 (got %d)", io.k); return PM_ERR_UNSUPPORTED;
  }
}

/* answers one batch on device pointers: Morton order, grid kernel, kd-tree kernel for the rest */
int run_queries(pm_ctx* ctx, const pm_map* map, pmk::QueryIO io, bool gather, bool time_it) {
  int* perm = nullptr;
  PM_CHECK(sort_queries(ctx, map, io.q, gather ? io.brdf : nullptr, io.nq, &perm));
  io.order = perm;
  ctx->fb_count_dev = nullptr;
  ctx->ho_count_dev = nullptr;
  if (time_it) PM_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  static const bool kd_only = getenv("PM_KD_ONLY") != nullptr; /* measurement knob: every query through the kd-tree kernel */
  if (map->grid_pts && !kd_only) {
    int *fb_list = nullptr, *fb_count = nullptr;
    PM_CHECK(ctx->get_scratch(pm::S_MISC1, (size_t)io.nq, &fb_list));
    PM_CHECK(ctx->get_scratch(pm::S_MISC2, (size_t)4, &fb_count));
    PM_CUDA(cudaMemsetAsync(fb_count, 0, sizeof(int), ctx->stream));
    ctx->fb_count_dev = fb_count;
    PM_CHECK(pm::grid_query(ctx, map, io, gather, fb_list, fb_count));
    PM_CHECK(pm::kd_query(ctx, map, io, gather, fb_list, fb_count));
  } else {
    PM_CHECK(pm::kd_query(ctx, map, io, gather, nullptr, nullptr));
  }
  if (time_it) PM_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
  return PM_OK;
}

/* end of a synchronous query call: wait, then read the kernel time and the fallback count */
int finish_stats(pm_ctx* ctx, const pm_map* map) {
  int fb = 0, ho = 0;
  const bool have = map && map->n >= 1;
  if (have && ctx->fb_count_dev) PM_CUDA(cudaMemcpyAsync(&fb, ctx->fb_count_dev, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  if (have && ctx->ho_count_dev) PM_CUDA(cudaMemcpyAsync(&ho, ctx->ho_count_dev, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  if (have) PM_CUDA(cudaEventElapsedTime(&ctx->last_kernel_ms, ctx->ev0, ctx->ev1));
  ctx->last_fallbacks = have && ctx->fb_count_dev ? fb : (have ? -1 : 0);
  ctx->last_handovers = have && ctx->ho_count_dev ? ho : 0;
  return PM_OK;
}

/* pm_knn with HOST output buffers: the batch is cut into chunks so that the device-to-host copy of
 * chunk c (PCIe, the slow part: k * 8 bytes per query) overlaps the kernels of chunk c + 1. Outputs are
 * staged in two alternating device buffers; a chunk's kernels wait for the copy that last used its buffer. */
constexpr int64_t kPipeChunk = 131072;

int knn_pipelined(pm_ctx* ctx, const pm_map* map, const float* q_dev, int64_t nq, int k, float cutoff2, int32_t* out_ids,
                  float* out_d2, float* out_r2) {
  if (!ctx->copy_stream) {
    PM_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) {
      PM_CUDA(cudaEventCreateWithFlags(&ctx->ev_kernel[i], cudaEventDisableTiming));
      PM_CUDA(cudaEventCreateWithFlags(&ctx->ev_copy[i], cudaEventDisableTiming));
    }
  }
  int32_t* s_ids = nullptr; float *s_d2 = nullptr, *s_r2 = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_STAGE_OUT0, (size_t)(2 * kPipeChunk * k), &s_ids));
  if (out_d2) PM_CHECK(ctx->get_scratch(pm::S_STAGE_OUT1, (size_t)(2 * kPipeChunk * k), &s_d2));
  if (out_r2) PM_CHECK(ctx->get_scratch(pm::S_STAGE_OUT2, (size_t)(2 * kPipeChunk), &s_r2));
  PM_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  int c = 0;
  for (int64_t b = 0; b < nq; b += kPipeChunk, c++) {
    const int64_t cnt = nq - b < kPipeChunk ? nq - b : kPipeChunk;
    const int buf = c & 1;
    if (c >= 2) PM_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_copy[buf], 0));
    pmk::QueryIO io;
    io.q = q_dev + 3 * b; io.nq = (int)cnt; io.k = k; io.cutoff2 = cutoff2;
    io.ids = s_ids + (size_t)buf * kPipeChunk * k;
    io.d2 = s_d2 ? s_d2 + (size_t)buf * kPipeChunk * k : nullptr;
    io.r2 = s_r2 ? s_r2 + (size_t)buf * kPipeChunk : nullptr;
    PM_CHECK(run_queries(ctx, map, io, false, false));
    PM_CUDA(cudaEventRecord(ctx->ev_kernel[buf], ctx->stream));
    PM_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_kernel[buf], 0));
    PM_CUDA(cudaMemcpyAsync(out_ids + (size_t)b * k, io.ids, sizeof(int32_t) * (size_t)cnt * k, cudaMemcpyDeviceToHost, ctx->copy_stream));
    if (out_d2) PM_CUDA(cudaMemcpyAsync(out_d2 + (size_t)b * k, io.d2, sizeof(float) * (size_t)cnt * k, cudaMemcpyDeviceToHost, ctx->copy_stream));
    if (out_r2) PM_CUDA(cudaMemcpyAsync(out_r2 + b, io.r2, sizeof(float) * (size_t)cnt, cudaMemcpyDeviceToHost, ctx->copy_stream));
    PM_CUDA(cudaEventRecord(ctx->ev_copy[buf], ctx->copy_stream));
  }
  PM_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
  PM_CUDA(cudaStreamSynchronize(ctx->copy_stream));
  return PM_OK;
}

}  // namespace

namespace pm {

int kd_query(pm_ctx* ctx, const pm_map* map, const pmk::QueryIO& io, bool gather, const int* list, const int* list_count) {
  /* the warp-cooperative kernel answers everything; the per-thread restatement of cukd's traversal stays
   * selectable (PM_KD_KERNEL=thread) as the literal reference order */
  static const bool per_thread = getenv("PM_KD_KERNEL") && !strcmp(getenv("PM_KD_KERNEL"), "thread");
  if (!per_thread) return kdw_query(ctx, map, io, gather, list, list_count);
  return gather ? dispatch_kd<true>(ctx, map, io, list, list_count) : dispatch_kd<false>(ctx, map, io, list, list_count);
}

int map_bounds(pm_ctx* ctx, pm_map* m) {
  uint32_t* d = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_MISC3, (size_t)8, &d));
  const uint32_t init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
  PM_CUDA(cudaMemcpyAsync(d, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream));
  int64_t grid = div_up(m->n, 256);
  if (grid > ctx->sm_count * 8) grid = ctx->sm_count * 8;
  k_bounds<<<(unsigned)grid, 256, 0, ctx->stream>>>(m->nodes4, (int)m->n, d);
  ctx->launches++;
  uint32_t h[6];
  PM_CUDA(cudaMemcpyAsync(h, d, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  for (int i = 0; i < 6; i++) {
    const uint32_t u = (h[i] & 0x80000000u) ? (h[i] & 0x7fffffffu) : ~h[i];
    memcpy(&m->bounds[i], &u, 4);
  }
  return PM_OK;
}

/* device-pointer entry used by pm_gather and by the render pipeline */
int gather_device(pm_ctx* ctx, const pm_map* map, const float* hit_dev, const float* brdf_dev, int64_t nq, int k,
                  float cutoff, float* rgb_dev, bool time_it) {
  if (nq <= 0) return PM_OK;
  if (!map || map->n < 1) {
    k_fill_f<<<(unsigned)div_up(nq * 3, 256), 256, 0, ctx->stream>>>(nq * 3, rgb_dev, 0.f);
    ctx->launches++;
    return PM_OK;
  }
  if (!(k == 8 || k == 16 || k == 32 || k == 50 || k == 64)) { set_error("k must be one of 8, 16, 32, 50, 64 (got %d)", k); return PM_ERR_UNSUPPORTED; }
  pmk::QueryIO io;
  io.q = hit_dev; io.brdf = brdf_dev; io.nq = (int)nq; io.k = k; io.cutoff2 = cutoff * cutoff; io.rgb = rgb_dev;
  return run_queries(ctx, map, io, true, time_it);
}

}  // namespace pm

extern "C" {

int pm_knn(pm_ctx* ctx, const pm_map* map, const float* queries, int64_t nq, int k, float cutoff_radius,
           int32_t* out_ids, float* out_dist2, float* out_r2) {
  PM_REQUIRE(ctx, "ctx is NULL");
  PM_REQUIRE(!map || !map->cells_only, "this map was built by pm_map_build_cells (no kd-tree): fixed-radius gathers only");
  PM_REQUIRE(nq >= 0 && nq < ((int64_t)1 << 31), "nq out of range");
  PM_REQUIRE(k == 8 || k == 16 || k == 32 || k == 50 || k == 64, "k must be one of 8, 16, 32, 50, 64");
  if (nq == 0) return PM_OK;
  PM_REQUIRE(queries && out_ids, "queries/out_ids is NULL");
  PM_CUDA(cudaSetDevice(ctx->device));
  const float cutoff2 = cutoff_radius * cutoff_radius;
  const void* q_dev = nullptr;
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN0, queries, sizeof(float) * 3 * (size_t)nq, &q_dev));
  if (map && map->n >= 1 && nq >= 2 * kPipeChunk && !pm::is_device_ptr(out_ids) && !pm::is_device_ptr(out_dist2) &&
      !pm::is_device_ptr(out_r2)) {
    PM_CHECK(knn_pipelined(ctx, map, (const float*)q_dev, nq, k, cutoff2, out_ids, out_dist2, out_r2));
    PM_CHECK(finish_stats(ctx, map));
    return PM_OK;
  }
  pm::OutStage o_ids, o_d2, o_r2;
  PM_CHECK(pm::stage_out(ctx, pm::S_STAGE_OUT0, out_ids, sizeof(int32_t) * (size_t)k * nq, &o_ids));
  PM_CHECK(pm::stage_out(ctx, pm::S_STAGE_OUT1, out_dist2, sizeof(float) * (size_t)k * nq, &o_d2));
  PM_CHECK(pm::stage_out(ctx, pm::S_STAGE_OUT2, out_r2, sizeof(float) * (size_t)nq, &o_r2));
  if (!map || map->n < 1) {
    k_fill_empty<<<(unsigned)pm::div_up(nq * k, 256), 256, 0, ctx->stream>>>(nq * k, (int32_t*)o_ids.dev, (float*)o_d2.dev, cutoff2);
    if (o_r2.dev) k_fill_f<<<(unsigned)pm::div_up(nq, 256), 256, 0, ctx->stream>>>(nq, (float*)o_r2.dev, cutoff2);
    ctx->launches += 2;
  } else {
    pmk::QueryIO io;
    io.q = (const float*)q_dev; io.nq = (int)nq; io.k = k; io.cutoff2 = cutoff2;
    io.ids = (int32_t*)o_ids.dev; io.d2 = (float*)o_d2.dev; io.r2 = (float*)o_r2.dev;
    PM_CHECK(run_queries(ctx, map, io, false, true));
  }
  PM_CHECK(pm::stage_out_finish(ctx, o_ids));
  PM_CHECK(pm::stage_out_finish(ctx, o_d2));
  PM_CHECK(pm::stage_out_finish(ctx, o_r2));
  PM_CHECK(finish_stats(ctx, map));
  return PM_OK;
}

int pm_gather(pm_ctx* ctx, const pm_map* map, const float* hit, const float* brdf, int64_t nq, int k,
              float cutoff_radius, float* out_rgb) {
  PM_REQUIRE(ctx, "ctx is NULL");
  PM_REQUIRE(!map || !map->cells_only, "this map was built by pm_map_build_cells (no kd-tree): fixed-radius gathers only");
  PM_REQUIRE(nq >= 0 && nq < ((int64_t)1 << 31), "nq out of range");
  PM_REQUIRE(k == 8 || k == 16 || k == 32 || k == 50 || k == 64, "k must be one of 8, 16, 32, 50, 64");
  if (nq == 0) return PM_OK;
  PM_REQUIRE(hit && brdf && out_rgb, "hit/brdf/out_rgb is NULL");
  PM_CUDA(cudaSetDevice(ctx->device));
  const void *h_dev = nullptr, *b_dev = nullptr;
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN0, hit, sizeof(float) * 3 * (size_t)nq, &h_dev));
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN1, brdf, sizeof(float) * (size_t)nq, &b_dev));
  pm::OutStage o;
  PM_CHECK(pm::stage_out(ctx, pm::S_STAGE_OUT0, out_rgb, sizeof(float) * 3 * (size_t)nq, &o));
  PM_CHECK(pm::gather_device(ctx, map, (const float*)h_dev, (const float*)b_dev, nq, k, cutoff_radius, (float*)o.dev, true));
  PM_CHECK(pm::stage_out_finish(ctx, o));
  PM_CHECK(finish_stats(ctx, map));
  return PM_OK;
}

/* fixed-radius gather: one progressive-photon-mapping pass over a set of hit points */
int pm_gather_radius(pm_ctx* ctx, const pm_map* map, const float* hit, const float* brdf, const float* radius2, int64_t nq,
                     float* out_flux, int32_t* out_count) {
  PM_REQUIRE(ctx, "ctx is NULL");
  PM_REQUIRE(nq >= 0 && nq < ((int64_t)1 << 31), "nq out of range");
  if (nq == 0) return PM_OK;
  PM_REQUIRE(hit && brdf && radius2 && out_flux && out_count, "hit/brdf/radius2/out_flux/out_count is NULL");
  PM_CUDA(cudaSetDevice(ctx->device));
  const void *h_dev = nullptr, *b_dev = nullptr, *r_dev = nullptr;
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN0, hit, sizeof(float) * 3 * (size_t)nq, &h_dev));
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN1, brdf, sizeof(float) * (size_t)nq, &b_dev));
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN2, radius2, sizeof(float) * (size_t)nq, &r_dev));
  pm::OutStage of, oc;
  PM_CHECK(pm::stage_out(ctx, pm::S_STAGE_OUT0, out_flux, sizeof(float) * 3 * (size_t)nq, &of));
  PM_CHECK(pm::stage_out(ctx, pm::S_STAGE_OUT1, out_count, sizeof(int32_t) * (size_t)nq, &oc));
  ctx->fb_count_dev = nullptr;
  if (!map || map->n < 1) {
    PM_CUDA(cudaMemsetAsync(of.dev, 0, sizeof(float) * 3 * (size_t)nq, ctx->stream));
    PM_CUDA(cudaMemsetAsync(oc.dev, 0, sizeof(int32_t) * (size_t)nq, ctx->stream));
  } else {
    int* perm = nullptr;
    PM_CHECK(sort_queries(ctx, map, (const float*)h_dev, (const float*)b_dev, (int)nq, &perm));
    int *fb_list = nullptr, *fb_count = nullptr;
    PM_CHECK(ctx->get_scratch(pm::S_MISC1, (size_t)nq, &fb_list));
    PM_CHECK(ctx->get_scratch(pm::S_MISC2, (size_t)4, &fb_count));
    PM_CUDA(cudaMemsetAsync(fb_count, 0, sizeof(int), ctx->stream));
    if (map->grid_pts) ctx->fb_count_dev = fb_count;
    PM_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
    PM_CHECK(pm::radius_query(ctx, map, (const float*)h_dev, (const float*)b_dev, (const float*)r_dev, perm, (int)nq,
                              (float*)of.dev, (int32_t*)oc.dev, fb_list, fb_count));
    PM_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
  }
  PM_CHECK(pm::stage_out_finish(ctx, of));
  PM_CHECK(pm::stage_out_finish(ctx, oc));
  PM_CHECK(finish_stats(ctx, map));
  return PM_OK;
}

}  // extern "C"
