/* pm_knn.cu — stage 3 entry points: k-nearest-photon search and the fused radiance estimate.
 *
 * Replaces cukd::stackBased::knn<HeapCandidateList<k>> (externals/cudaKDTree/cukd/knn.h:298-364,
 * 536-549; traverse-default-stack-based.h:24-90) as instantiated by
 * ray-tracer/cuda/shading.h:11-18, and gatherPhotons (shading.h:93-121).
 *
 * Result contract: per query, the k smallest (floatBits(dist²), node id) pairs among
 * nodes with dist² < cutoff²; dist² = fma(dz,dz, fma(dx,dx, dy*dy)), d = node - query
 * (the contraction nvcc emits for cukd's dot(d,d) in the reference kernel).
 *
 * This file holds the query ordering (a counting sort of the queries into 2^21 Morton bins of the map's box, so
 * that the warps of a CTA work on neighbouring queries and share their candidate rows in L1/L2), the pipelined
 * host-buffer variant of pm_knn and the ABI entry points. The kernels that answer the queries live in
 * pm_grid.cu (uniform grid, one warp per query) and pm_kdwarp.cu (warp-cooperative kd-tree walk).
 */
#include <cfloat>
#include <cub/cub.cuh>
#include "pm_knn.cuh"

namespace {

using pmk::encode;

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

/* ---- query ordering: counting sort into Morton bins ------------------------------------------------
 * 4 to 7 bits per axis inside the map's box (queries outside are clamped onto its faces) + one last bin for queries that
 * will not be evaluated (brdf <= 0), so whole warps retire at once. Three kernels: histogram (one global atomic per
 * query, the last bin warp-aggregated), exclusive scan of the counters (single pass, decoupled look-back),
 * scatter (the atomic's return value is the slot inside the bin). The order inside a bin is whatever the atomics
 * hand out; results do not depend on it. */
constexpr int kQBitsMax = 7;
constexpr int kScanTile = 2048;                       /* counters per CTA of the scan: 256 threads x 8 */

__device__ __forceinline__ uint32_t bin_of(const float* __restrict__ q, const float* __restrict__ skip_if_not_positive, int i,
                                           const Box6& box, int bits) {
  if (skip_if_not_positive && !(skip_if_not_positive[i] > 0.f)) return 1u << (3 * bits);
  uint32_t code = 0;
#pragma unroll
  for (int d = 0; d < 3; d++) {
    float t = (q[(size_t)i * 3 + d] - box.lo[d]) * box.inv[d];
    t = fminf(fmaxf(t, 0.f), (float)((1 << bits) - 1)); /* drops NaN */
    code |= part1by2((uint32_t)t) << d;
  }
  return code;
}

/* kQItems queries per thread (striped: coalesced): the atomics of a thread are independent, so several are in flight */
constexpr int kQItems = 4;

__global__ void __launch_bounds__(256)
k_qbin_count(const float* __restrict__ q, const float* __restrict__ skip_if_not_positive, int nq, Box6 box, int bits,
             uint32_t* __restrict__ codes, uint32_t* __restrict__ hist) {
#pragma unroll
  for (int it = 0; it < kQItems; it++) {
    const int i = (blockIdx.x * kQItems + it) * blockDim.x + threadIdx.x;
    const bool in = i < nq;
    const uint32_t code = in ? bin_of(q, skip_if_not_positive, i, box, bits) : 0u;
    if (in) codes[i] = code;
    const bool last = in && code == 1u << (3 * bits);
    const unsigned lb = __ballot_sync(0xffffffffu, last);
    if (last) { if ((threadIdx.x & 31) == __ffs(lb) - 1) atomicAdd(hist + code, (uint32_t)__popc(lb)); }
    else if (in) atomicAdd(hist + code, 1u);
  }
}

/* in place: counts -> exclusive offsets. status[t] = flag << 32 | value with flag 1 = the tile's own sum, 2 = the
 * inclusive sum up to and including the tile; tiles take their number from a counter so that a tile never waits for
 * one that has not started */
__global__ void __launch_bounds__(256)
k_qbin_scan(uint32_t* __restrict__ hist, int nbins, unsigned long long* __restrict__ status, int* __restrict__ tile_counter) {
  __shared__ int s_tile;
  __shared__ uint32_t s_warp[8];
  __shared__ uint32_t s_prefix;
  if (threadIdx.x == 0) s_tile = atomicAdd(tile_counter, 1);
  __syncthreads();
  const int tile = s_tile;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int base = tile * kScanTile + threadIdx.x * 8;
  uint32_t v[8];
#pragma unroll
  for (int e = 0; e < 8; e++) v[e] = base + e < nbins ? hist[base + e] : 0u;
  uint32_t sum = 0;
#pragma unroll
  for (int e = 0; e < 8; e++) sum += v[e];
  uint32_t incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += u;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t wbase = 0, total = 0;
#pragma unroll
  for (int w = 0; w < 8; w++) {
    if (w < warp) wbase += s_warp[w];
    total += s_warp[w];
  }
  if (warp == 0) {
    uint32_t prefix = 0;
    if (tile == 0) {
      if (lane == 0) atomicExch(status, (2ull << 32) | total);
    } else {
      if (lane == 0) atomicExch(status + tile, (1ull << 32) | total);
      /* look back 32 tiles at a time */
      int first = tile - 1;
      while (true) {
        const int tt = first - lane;
        unsigned long long st = 0;
        bool ready;
        do {
          st = tt >= 0 ? *((volatile unsigned long long*)(status + tt)) : (2ull << 32);
          ready = (st >> 32) != 0;
        } while (!__all_sync(0xffffffffu, ready));
        const unsigned full = __ballot_sync(0xffffffffu, (st >> 32) == 2);
        /* lanes up to and including the nearest tile with an inclusive sum contribute */
        const int stop = full ? __ffs(full) - 1 : 32;
        uint32_t c = lane <= stop ? (uint32_t)st : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        prefix += c;
        if (full) break;
        first -= 32;
      }
      if (lane == 0) atomicExch(status + tile, (2ull << 32) | (prefix + total));
    }
    if (lane == 0) s_prefix = prefix;
  }
  __syncthreads();
  uint32_t run = s_prefix + wbase + incl - sum;
#pragma unroll
  for (int e = 0; e < 8; e++) {
    if (base + e < nbins) hist[base + e] = run;
    run += v[e];
  }
}

__global__ void __launch_bounds__(256)
k_qbin_scatter(const float* __restrict__ q, const uint32_t* __restrict__ codes, int nq, int bits,
               uint32_t* __restrict__ cursor, int* __restrict__ perm, float4* __restrict__ sq,
               const float* __restrict__ brdf, float* __restrict__ sbrdf) {
  uint32_t pos[kQItems];
  float qx[kQItems], qy[kQItems], qz[kQItems];
#pragma unroll
  for (int it = 0; it < kQItems; it++) {
    const int i = (blockIdx.x * kQItems + it) * blockDim.x + threadIdx.x;
    const bool in = i < nq;
    const uint32_t code = in ? codes[i] : 0u;
    const bool last = in && code == 1u << (3 * bits);
    const unsigned lb = __ballot_sync(0xffffffffu, last);
    pos[it] = 0u;
    if (last) {
      const int leader = __ffs(lb) - 1;
      if ((threadIdx.x & 31) == leader) pos[it] = atomicAdd(cursor + code, (uint32_t)__popc(lb));
      pos[it] = __shfl_sync(lb, pos[it], leader) + __popc(lb & ((1u << (threadIdx.x & 31)) - 1u));
    } else if (in) {
      pos[it] = atomicAdd(cursor + code, 1u);
    }
    if (in && sq) { qx[it] = q[(size_t)i * 3 + 0]; qy[it] = q[(size_t)i * 3 + 1]; qz[it] = q[(size_t)i * 3 + 2]; }
  }
#pragma unroll
  for (int it = 0; it < kQItems; it++) {
    const int i = (blockIdx.x * kQItems + it) * blockDim.x + threadIdx.x;
    if (i < nq) {
      if (perm) perm[pos[it]] = i;
      if (sq) sq[pos[it]] = make_float4(qx[it], qy[it], qz[it], __int_as_float(i));
      if (sbrdf) sbrdf[pos[it]] = brdf[i];
    }
  }
}

/* Morton-bin order of the queries inside the map's box -> permutation in scratch slot S_PERM and, with `sq_out`, the
 * queries themselves in that order as float4 (x, y, z, bits(index)) in S_VALS1 */
int sort_queries(pm_ctx* ctx, const pm_map* map, const float* q_dev, const float* brdf_dev, int nq, int** perm_out,
                 float4** sq_out, float** sbrdf_out = nullptr) {
  uint32_t *codes, *hist; int* perm; float4* sq = nullptr; float* sbrdf = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_KEYS0, (size_t)nq, &codes));
  PM_CHECK(ctx->get_scratch(pm::S_PERM, (size_t)nq, &perm));
  if (sq_out) PM_CHECK(ctx->get_scratch(pm::S_VALS1, (size_t)nq, &sq));
  if (sq_out && sbrdf_out && brdf_dev) PM_CHECK(ctx->get_scratch(pm::S_VALS0, (size_t)nq, &sbrdf));
  /* bins per axis: about four queries per bin, 2^4 .. 2^7 (a finer grid only costs the scan) */
  int bits = 4;
  while (bits < kQBitsMax && ((int64_t)1 << (3 * bits)) * 4 < nq) bits++;
  const int bins = (1 << (3 * bits)) + 1, tiles = (bins + kScanTile - 1) / kScanTile;
  /* counters, then the scan's tile states and its tile counter: one memset */
  const size_t hist_words = ((size_t)bins + 3) & ~(size_t)3;
  const size_t bytes = hist_words * 4 + (size_t)tiles * 8 + 16;
  PM_CHECK(ctx->get_scratch(pm::S_KEYS1, bytes, &hist));
  unsigned long long* status = reinterpret_cast<unsigned long long*>(hist + hist_words);
  int* tile_counter = reinterpret_cast<int*>(status + tiles);
  PM_CUDA(cudaMemsetAsync(hist, 0, bytes, ctx->stream));
  Box6 box;
  for (int d = 0; d < 3; d++) {
    box.lo[d] = map->bounds[d];
    const float w = map->bounds[3 + d] - map->bounds[d];
    box.inv[d] = w > 0.f ? (float)(1 << bits) / w : 0.f;
  }
  const unsigned grid = (unsigned)pm::div_up(nq, 256 * kQItems);
  k_qbin_count<<<grid, 256, 0, ctx->stream>>>(q_dev, brdf_dev, nq, box, bits, codes, hist);
  k_qbin_scan<<<tiles, 256, 0, ctx->stream>>>(hist, bins, status, tile_counter);
  /* the fast path reads the ordered queries themselves (sq) and nothing else reads the permutation then */
  k_qbin_scatter<<<grid, 256, 0, ctx->stream>>>(q_dev, codes, nq, bits, hist, sq ? nullptr : perm, sq, brdf_dev, sbrdf);
  ctx->launches += 3;
  PM_CUDA(cudaGetLastError());
  *perm_out = perm;
  if (sq_out) *sq_out = sq;
  if (sbrdf_out) *sbrdf_out = sbrdf;
  return PM_OK;
}

/* The same Morton-bin order, but STABLE (input order inside a bin): a radix sort of (bin, index) pairs over the bits
 * in use. pm_gather_radius needs it: its group kernel sums a hit point's flux in an order that depends on which
 * neighbours share its warp, so the order of the list has to be a function of the input alone for the result to be
 * reproducible bit for bit — and input order inside a bin is raster order, which is what makes neighbours neighbours. */
__global__ void k_qbin_codes(const float* __restrict__ q, const float* __restrict__ skip_if_not_positive, int nq, Box6 box, int bits,
                             uint32_t* __restrict__ codes, int* __restrict__ iota) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  codes[i] = bin_of(q, skip_if_not_positive, i, box, bits);
  iota[i] = i;
}

int sort_queries_stable(pm_ctx* ctx, const float* bounds /* lo[3], hi[3] */, const float* q_dev, const float* brdf_dev, int nq,
                        int** perm_out) {
  uint32_t *codes, *codes_sorted; int *iota, *perm;
  PM_CHECK(ctx->get_scratch(pm::S_KEYS0, (size_t)nq, &codes));
  PM_CHECK(ctx->get_scratch(pm::S_KEYS1, (size_t)nq, &codes_sorted));
  PM_CHECK(ctx->get_scratch(pm::S_VALS0, (size_t)nq, &iota));
  PM_CHECK(ctx->get_scratch(pm::S_PERM, (size_t)nq, &perm));
  int bits = 4;
  while (bits < kQBitsMax && ((int64_t)1 << (3 * bits)) * 4 < nq) bits++;
  Box6 box;
  for (int d = 0; d < 3; d++) {
    box.lo[d] = bounds[d];
    const float w = bounds[3 + d] - bounds[d];
    box.inv[d] = w > 0.f ? (float)(1 << bits) / w : 0.f;
  }
  k_qbin_codes<<<(unsigned)pm::div_up(nq, 256), 256, 0, ctx->stream>>>(q_dev, brdf_dev, nq, box, bits, codes, iota);
  size_t tb = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tb, codes, codes_sorted, iota, perm, nq, 0, 3 * bits + 1, ctx->stream);
  void* tmp = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_CUB, tb, &tmp));
  PM_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, codes, codes_sorted, iota, perm, nq, 0, 3 * bits + 1, ctx->stream));
  ctx->launches += 1 + (3 * bits + 1 + 7) / 8 + 2;
  PM_CUDA(cudaGetLastError());
  *perm_out = perm;
  return PM_OK;
}

/* bounding box of the hit points that will be evaluated (brdf > 0): ordered-integer min / max */
__device__ __forceinline__ uint32_t ordered_bits(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__global__ void k_hit_bounds(const float* __restrict__ q, const float* __restrict__ brdf, int nq, uint32_t* __restrict__ out /* lo[3], hi[3] */) {
  uint32_t lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nq; i += gridDim.x * blockDim.x) {
    if (!(brdf[i] > 0.f)) continue;
#pragma unroll
    for (int d = 0; d < 3; d++) {
      const float v = q[(size_t)i * 3 + d];
      if (v != v) continue;
      const uint32_t k = ordered_bits(v);
      lo[d] = min(lo[d], k); hi[d] = max(hi[d], k);
    }
  }
#pragma unroll
  for (int d = 0; d < 3; d++) {
    const uint32_t l = __reduce_min_sync(0xffffffffu, lo[d]), h = __reduce_max_sync(0xffffffffu, hi[d]);
    if ((threadIdx.x & 31) == 0) { atomicMin(out + d, l); atomicMax(out + 3 + d, h); }
  }
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

/* pm_selftest_math: the call-free division / square root of pm_knn.cuh against the IEEE intrinsics on hashed
 * operands — random bit patterns (every exponent, both signs), and values drawn like the gather's (distances and
 * radii in [1e-6, 1e3], flux sums of any sign). Counts operands the predicates accept whose results differ. */
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
__global__ void k_selftest_math(int64_t n, uint32_t seed, unsigned long long* __restrict__ out /* mismatches, accepted, rejected */) {
  unsigned long long bad = 0, acc = 0, rej = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t h0 = mix32((uint32_t)i * 2u + seed), h1 = mix32((uint32_t)i * 2u + 1u + seed * 31u);
    float a, b;
    if ((i & 3) == 0) { a = __uint_as_float(h0); b = __uint_as_float(h1); }                 /* any bit pattern */
    else {
      /* magnitudes 10^(-6 .. 3), a signed in a quarter of the cases, exact zero now and then */
      a = exp2f(-20.f + 30.f * (float)(h0 >> 8) * (1.f / 16777216.f));
      b = exp2f(-20.f + 30.f * (float)(h1 >> 8) * (1.f / 16777216.f));
      if ((h0 & 3u) == 0u) a = -a;
      if ((h0 & 0xfcu) == 0u) a = 0.f;
    }
    if (pmk::div_fast_ok(a, b)) {
      acc++;
      if (__float_as_uint(pmk::div_rn_fast(a, b)) != __float_as_uint(__fdiv_rn(a, b))) bad++;
    } else rej++;
    const float x = fabsf(a);
    if (pmk::sqrt_fast_ok(x)) {
      acc++;
      if (__float_as_uint(pmk::sqrt_rn_fast(x)) != __float_as_uint(__fsqrt_rn(x))) bad++;
    } else rej++;
  }
  atomicAdd(out, bad); atomicAdd(out + 1, acc); atomicAdd(out + 2, rej);
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

/* answers one batch on device pointers: Morton-bin order, grid kernels, kd-tree kernel for the rest. The event pair
 * (time_it) brackets the whole batch, ordering pass included. */
int run_queries(pm_ctx* ctx, const pm_map* map, pmk::QueryIO io, bool gather, bool time_it) {
  if (time_it) PM_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  int* perm = nullptr;
  float4* sq = nullptr;
  float* sbrdf = nullptr;
  PM_CHECK(sort_queries(ctx, map, io.q, gather ? io.brdf : nullptr, io.nq, &perm, map->fast_path ? &sq : nullptr, &sbrdf));
  io.order = sq ? nullptr : perm; /* with sq the permutation is not written (sort_queries) */
  io.sq = sq;
  io.sbrdf = sbrdf;
  ctx->fb_count_dev = nullptr;
  ctx->ho_count_dev = nullptr;
  if (map->grid_pts) {
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

/* end of a synchronous query call: the two counters travel to pinned host memory on the stream, one wait, then the
 * kernel time */
int finish_stats(pm_ctx* ctx, const pm_map* map) {
  const bool have = map && map->n >= 1;
  ctx->host_counts[0] = ctx->host_counts[1] = 0;
  if (have && ctx->fb_count_dev) PM_CUDA(cudaMemcpyAsync(ctx->host_counts, ctx->fb_count_dev, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  if (have && ctx->ho_count_dev) PM_CUDA(cudaMemcpyAsync(ctx->host_counts + 1, ctx->ho_count_dev, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  if (have) PM_CUDA(cudaEventElapsedTime(&ctx->last_kernel_ms, ctx->ev0, ctx->ev1));
  ctx->last_fallbacks = have && ctx->fb_count_dev ? ctx->host_counts[0] : (have ? -1 : 0);
  ctx->last_handovers = have && ctx->ho_count_dev ? ctx->host_counts[1] : 0;
  return PM_OK;
}

/* pm_knn with HOST output buffers: the batch is cut into chunks so that the device-to-host copy of
 * chunk c (PCIe, the slow part: k * 8 bytes per query) overlaps the kernels of chunk c + 1. Outputs are
 * staged in two alternating device buffers; a chunk's kernels wait for the copy that last used its buffer.
 * The first chunks are small (16 Ki queries, doubling up to kPipeChunk): the result copies — which bound the call —
 * start after ~0.05 ms instead of after a full chunk. Host QUERIES travel the same way: only the first chunk's
 * queries are copied ahead of its kernels, the rest follow on a third stream (the other PCIe direction) while chunk 0
 * is being answered. */
constexpr int64_t kPipeChunk = 131072, kPipeFirstChunk = 16384;

int knn_pipelined(pm_ctx* ctx, const pm_map* map, const float* q_dev, const float* q_host, int64_t nq, int k, float cutoff2,
                  int32_t* out_ids, float* out_d2, float* out_r2) {
  if (!ctx->copy_stream) {
    PM_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) {
      PM_CUDA(cudaEventCreateWithFlags(&ctx->ev_kernel[i], cudaEventDisableTiming));
      PM_CUDA(cudaEventCreateWithFlags(&ctx->ev_copy[i], cudaEventDisableTiming));
    }
  }
  if (q_host && !ctx->in_stream) {
    PM_CUDA(cudaStreamCreateWithFlags(&ctx->in_stream, cudaStreamNonBlocking));
    PM_CUDA(cudaEventCreateWithFlags(&ctx->ev_in, cudaEventDisableTiming));
  }
  int32_t* s_ids = nullptr; float *s_d2 = nullptr, *s_r2 = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_STAGE_OUT0, (size_t)(2 * kPipeChunk * k), &s_ids));
  if (out_d2) PM_CHECK(ctx->get_scratch(pm::S_STAGE_OUT1, (size_t)(2 * kPipeChunk * k), &s_d2));
  if (out_r2) PM_CHECK(ctx->get_scratch(pm::S_STAGE_OUT2, (size_t)(2 * kPipeChunk), &s_r2));
  PM_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  int c = 0;
  int64_t size = kPipeFirstChunk;
  for (int64_t b = 0; b < nq; b += size, size = std::min<int64_t>(2 * size, kPipeChunk), c++) {
    const int64_t cnt = std::min<int64_t>(nq - b, size);
    const int buf = c & 1;
    if (q_host) {
      float* qd = const_cast<float*>(q_dev);
      if (c == 0) {
        PM_CUDA(cudaMemcpyAsync(qd, q_host, sizeof(float) * 3 * (size_t)cnt, cudaMemcpyHostToDevice, ctx->stream));
        if (cnt < nq) {
          PM_CUDA(cudaMemcpyAsync(qd + 3 * cnt, q_host + 3 * cnt, sizeof(float) * 3 * (size_t)(nq - cnt), cudaMemcpyHostToDevice, ctx->in_stream));
          PM_CUDA(cudaEventRecord(ctx->ev_in, ctx->in_stream));
        }
      } else if (c == 1) {
        PM_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_in, 0));
      }
    }
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
  return kdw_query(ctx, map, io, gather, list, list_count);
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

/* pm_gather with HOST buffers on both sides (16 B in, 12 B out per query): the three legs — queries in, kernels,
 * colours out — run on three streams chunk by chunk, so the call costs about what its
 * slowest leg costs instead of their sum. The staging buffers hold the whole batch: no buffer is reused inside a call. */
/* few, large chunks: a chunk costs about ten launches on the host, and here the kernels are the slowest leg */
constexpr int64_t kGatherFirstChunk = 65536, kGatherChunk = 262144;
static int gather_pipelined(pm_ctx* ctx, const pm_map* map, const float* hit, const float* brdf, int64_t nq, int k, float cutoff,
                            float* out_rgb) {
  if (!ctx->copy_stream) {
    PM_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) {
      PM_CUDA(cudaEventCreateWithFlags(&ctx->ev_kernel[i], cudaEventDisableTiming));
      PM_CUDA(cudaEventCreateWithFlags(&ctx->ev_copy[i], cudaEventDisableTiming));
    }
  }
  if (!ctx->in_stream) {
    PM_CUDA(cudaStreamCreateWithFlags(&ctx->in_stream, cudaStreamNonBlocking));
    PM_CUDA(cudaEventCreateWithFlags(&ctx->ev_in, cudaEventDisableTiming));
  }
  float *h_dev = nullptr, *b_dev = nullptr, *rgb_dev = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_STAGE_IN0, (size_t)nq * 3, &h_dev));
  PM_CHECK(ctx->get_scratch(pm::S_STAGE_IN1, (size_t)nq, &b_dev));
  PM_CHECK(ctx->get_scratch(pm::S_STAGE_OUT0, (size_t)nq * 3, &rgb_dev));
  size_t chunks = 0;
  for (int64_t b = 0, size = kGatherFirstChunk; b < nq; b += size, size = kGatherChunk) chunks++;
  while (ctx->pipe_events.size() < 2 * chunks) {
    cudaEvent_t e;
    PM_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    ctx->pipe_events.push_back(e);
  }
  /* leg 1: every chunk's inputs, in order, on the input stream */
  size_t c = 0;
  for (int64_t b = 0, size = kGatherFirstChunk; b < nq; b += size, size = kGatherChunk, c++) {
    const int64_t cnt = std::min<int64_t>(nq - b, size);
    PM_CUDA(cudaMemcpyAsync(h_dev + 3 * b, hit + 3 * b, sizeof(float) * 3 * (size_t)cnt, cudaMemcpyHostToDevice, ctx->in_stream));
    PM_CUDA(cudaMemcpyAsync(b_dev + b, brdf + b, sizeof(float) * (size_t)cnt, cudaMemcpyHostToDevice, ctx->in_stream));
    PM_CUDA(cudaEventRecord(ctx->pipe_events[2 * c], ctx->in_stream));
  }
  /* legs 2 and 3 */
  PM_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  c = 0;
  for (int64_t b = 0, size = kGatherFirstChunk; b < nq; b += size, size = kGatherChunk, c++) {
    const int64_t cnt = std::min<int64_t>(nq - b, size);
    PM_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->pipe_events[2 * c], 0));
    PM_CHECK(pm::gather_device(ctx, map, h_dev + 3 * b, b_dev + b, cnt, k, cutoff, rgb_dev + 3 * b, false));
    PM_CUDA(cudaEventRecord(ctx->pipe_events[2 * c + 1], ctx->stream));
    PM_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->pipe_events[2 * c + 1], 0));
    PM_CUDA(cudaMemcpyAsync(out_rgb + 3 * b, rgb_dev + 3 * b, sizeof(float) * 3 * (size_t)cnt, cudaMemcpyDeviceToHost, ctx->copy_stream));
  }
  PM_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
  PM_CUDA(cudaStreamSynchronize(ctx->copy_stream));
  return PM_OK;
}

extern "C" {

int pm_selftest_math(pm_ctx* ctx, int64_t n, uint32_t seed, int64_t* mismatches, int64_t* accepted, int64_t* rejected) {
  PM_REQUIRE(ctx && mismatches && n >= 0, "bad arguments");
  PM_CUDA(cudaSetDevice(ctx->device));
  unsigned long long* d = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_MISC3, (size_t)4, &d));
  PM_CUDA(cudaMemsetAsync(d, 0, sizeof(unsigned long long) * 4, ctx->stream));
  k_selftest_math<<<(unsigned)(ctx->sm_count * 8), 256, 0, ctx->stream>>>(n, seed, d);
  ctx->launches++;
  unsigned long long h[3] = {0, 0, 0};
  PM_CUDA(cudaMemcpyAsync(h, d, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  *mismatches = (int64_t)h[0];
  if (accepted) *accepted = (int64_t)h[1];
  if (rejected) *rejected = (int64_t)h[2];
  return PM_OK;
}

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
  if (map && map->n >= 1 && nq >= 2 * kPipeChunk && !pm::is_device_ptr(out_ids) && !pm::is_device_ptr(out_dist2) &&
      !pm::is_device_ptr(out_r2)) {
    const float* q_host = nullptr;
    if (pm::is_device_ptr(queries)) q_dev = queries;
    else { /* host queries are copied chunk-wise by the pipeline itself */
      float* d = nullptr;
      PM_CHECK(ctx->get_scratch(pm::S_STAGE_IN0, (size_t)nq * 3, &d));
      q_dev = d; q_host = queries;
    }
    PM_CHECK(knn_pipelined(ctx, map, (const float*)q_dev, q_host, nq, k, cutoff2, out_ids, out_dist2, out_r2));
    PM_CHECK(finish_stats(ctx, map));
    return PM_OK;
  }
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN0, queries, sizeof(float) * 3 * (size_t)nq, &q_dev));
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
  if (map && map->n >= 1 && nq >= 2 * kPipeChunk && !pm::is_device_ptr(hit) && !pm::is_device_ptr(brdf) && !pm::is_device_ptr(out_rgb)) {
    PM_CHECK(gather_pipelined(ctx, map, hit, brdf, nq, k, cutoff_radius, out_rgb));
    PM_CHECK(finish_stats(ctx, map));
    return PM_OK;
  }
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
static int gather_radius_impl(pm_ctx* ctx, const pm_map* map, const pm_hit_order* order, const float* hit, const float* brdf,
                              const float* radius2, int64_t nq, float* out_flux, int32_t* out_count) {
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
  ctx->ho_count_dev = nullptr;
  if (!map || map->n < 1) {
    PM_CUDA(cudaMemsetAsync(of.dev, 0, sizeof(float) * 3 * (size_t)nq, ctx->stream));
    PM_CUDA(cudaMemsetAsync(oc.dev, 0, sizeof(int32_t) * (size_t)nq, ctx->stream));
  } else {
    int* perm = nullptr;
    if (order) perm = order->perm;
    else PM_CHECK(sort_queries_stable(ctx, map->bounds, (const float*)h_dev, (const float*)b_dev, (int)nq, &perm));
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

int pm_gather_radius(pm_ctx* ctx, const pm_map* map, const float* hit, const float* brdf, const float* radius2, int64_t nq,
                     float* out_flux, int32_t* out_count) {
  return gather_radius_impl(ctx, map, nullptr, hit, brdf, radius2, nq, out_flux, out_count);
}

int pm_hit_order_create(pm_ctx* ctx, const float* hit, const float* brdf, int64_t nq, pm_hit_order** out) {
  PM_REQUIRE(ctx && out, "ctx/out is NULL");
  *out = nullptr;
  PM_REQUIRE(nq >= 0 && nq < ((int64_t)1 << 31), "nq out of range");
  if (nq == 0) return PM_OK;
  PM_REQUIRE(hit && brdf, "hit/brdf is NULL");
  PM_CUDA(cudaSetDevice(ctx->device));
  const void *h_dev = nullptr, *b_dev = nullptr;
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN0, hit, sizeof(float) * 3 * (size_t)nq, &h_dev));
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN1, brdf, sizeof(float) * (size_t)nq, &b_dev));
  uint32_t* bb = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_MISC3, (size_t)8, &bb));
  const uint32_t init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
  PM_CUDA(cudaMemcpyAsync(bb, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream));
  k_hit_bounds<<<(unsigned)std::min<int64_t>(pm::div_up(nq, 256), (int64_t)ctx->sm_count * 8), 256, 0, ctx->stream>>>(
      (const float*)h_dev, (const float*)b_dev, (int)nq, bb);
  uint32_t got[6];
  PM_CUDA(cudaMemcpyAsync(got, bb, sizeof got, cudaMemcpyDeviceToHost, ctx->stream));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  float bounds[6];
  for (int d = 0; d < 6; d++) {
    const uint32_t k = got[d], u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    memcpy(&bounds[d], &u, 4);
  }
  if (got[0] > got[3]) for (int d = 0; d < 6; d++) bounds[d] = 0.f; /* no hit point is lit: any order will do */
  int* perm = nullptr;
  PM_CHECK(sort_queries_stable(ctx, bounds, (const float*)h_dev, (const float*)b_dev, (int)nq, &perm));
  pm_hit_order* o = new pm_hit_order();
  o->device = ctx->device; o->nq = nq;
  if (pm::map_alloc(&o->perm, sizeof(int) * (size_t)nq, ctx->stream) != cudaSuccess) {
    cudaGetLastError(); delete o; pm::set_error("pm_hit_order_create: out of device memory"); return PM_ERR_NOMEM;
  }
  PM_CUDA(cudaMemcpyAsync(o->perm, perm, sizeof(int) * (size_t)nq, cudaMemcpyDeviceToDevice, ctx->stream));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->launches++;
  *out = o;
  return PM_OK;
}

void pm_hit_order_destroy(pm_hit_order* order) {
  if (!order) return;
  cudaSetDevice(order->device);
  cudaDeviceSynchronize();
  pm::map_free(order->perm);
  delete order;
}

int pm_gather_radius_ordered(pm_ctx* ctx, const pm_map* map, const pm_hit_order* order, const float* hit, const float* brdf,
                             const float* radius2, int64_t nq, float* out_flux, int32_t* out_count) {
  PM_REQUIRE(order || nq == 0, "order is NULL");
  PM_REQUIRE(!order || (order->nq == nq && ctx && order->device == ctx->device), "order was made for another set of hit points (size or device)");
  return gather_radius_impl(ctx, map, order, hit, brdf, radius2, nq, out_flux, out_count);
}

}  // extern "C"

/* ---- pm_knn_static: pm_knn without any device-memory allocation -----------------------------------------------
 * No counterpart in the reference; added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). The scratch the query
 * kernels need — Morton codes, permutation, ordered queries, bin counters, the hand-over lists of k_knn_fast and
 * k_grid_query — lives in ONE file-scope device array instead of the context's growable cudaMalloc'ed slots.
 * Every GPU that loads the library holds the array (12.0 MiB) until the process exits. All calls share it: two
 * calls that overlap on the GPU (two contexts, two host threads) race. Nothing in it has to survive a call; what
 * must start at zero (bin counters, scan states, list counters) is cleared by the same memsets pm_knn issues. */
namespace {

constexpr int64_t kStaticMaxQueries = PM_KNN_STATIC_MAX_QUERIES;

/* sort_queries' counter-block size for a batch of nq queries (the same arithmetic, restated for the layout) */
constexpr size_t static_hist_bytes(int64_t nq) {
  int bits = 4;
  while (bits < kQBitsMax && ((int64_t)1 << (3 * bits)) * 4 < nq) bits++;
  const size_t bins = ((size_t)1 << (3 * bits)) + 1, tiles = (bins + kScanTile - 1) / kScanTile;
  return ((bins + 3) & ~(size_t)3) * 4 + tiles * 8 + 16;
}
constexpr size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }

/* slot -> (offset, bytes) inside the array: exactly the slots run_queries / sort_queries / grid_query ask for */
struct StaticSlot { int slot; size_t bytes; };
constexpr StaticSlot kStaticSlots[] = {
    {pm::S_KEYS0, align256((size_t)kStaticMaxQueries * sizeof(uint32_t))},   /* Morton bin per query */
    {pm::S_PERM, align256((size_t)kStaticMaxQueries * sizeof(int))},         /* processing order */
    {pm::S_VALS1, align256((size_t)kStaticMaxQueries * sizeof(float4))},     /* queries in that order */
    /* bin counters + scan states; x 4: sort_queries asks for that many uint32_t, not bytes */
    {pm::S_KEYS1, align256(static_hist_bytes(kStaticMaxQueries) * sizeof(uint32_t))},
    {pm::S_MISC1, align256((size_t)kStaticMaxQueries * sizeof(int))},        /* kd-tree hand-over list */
    {pm::S_MISC2, 256},                                                      /* ... its counter */
    {pm::S_FAST1, align256((size_t)kStaticMaxQueries * sizeof(int))},        /* fast-path hand-over list */
    {pm::S_FAST2, 256},                                                      /* ... its counter */
};
constexpr size_t static_total() {
  size_t t = 0;
  for (const StaticSlot& s : kStaticSlots) t += s.bytes;
  return t;
}
constexpr size_t kStaticBytes = static_total();
static_assert(kStaticBytes <= ((size_t)13 << 20), "pm_knn_static: the resident scratch is meant to stay modest");

/* In HOST code this name is only the host-side stand-in of the symbol, not a device address: the address the kernels
 * get comes from cudaGetSymbolAddress (static_scratch_base). */
static __device__ __align__(256) unsigned char g_knn_static_scratch[kStaticBytes];

constexpr int kStaticMaxDevices = 64;
unsigned char* g_static_base[kStaticMaxDevices] = {};

/* device address of the array on the CURRENT device (the caller has made ctx->device current), looked up once per
 * device */
int static_scratch_base(int device, unsigned char** out) {
  if (device < 0 || device >= kStaticMaxDevices) { pm::set_error("pm_knn_static: device %d out of range", device); return PM_ERR_INVALID; }
  unsigned char* p = __atomic_load_n(&g_static_base[device], __ATOMIC_ACQUIRE);
  if (!p) {
    void* sym = nullptr;
    PM_CUDA(cudaGetSymbolAddress(&sym, g_knn_static_scratch));
    if (!sym) { pm::set_error("pm_knn_static: cudaGetSymbolAddress returned NULL"); return PM_ERR_CUDA; }
    p = static_cast<unsigned char*>(sym);
    __atomic_store_n(&g_static_base[device], p, __ATOMIC_RELEASE); /* every thread computes the same value */
  }
  *out = p;
  return PM_OK;
}

/* Swaps the context's scratch table for one whose slots point into the static array, for the duration of one call.
 * A slot the array has no room for is marked (non-NULL pointer, capacity 0) so that pm_ctx::get_scratch fails on it
 * (it would have to cudaFree an address cudaMalloc never returned) instead of allocating behind the caller's back. */
struct StaticScratchScope {
  pm_ctx* ctx;
  pm::Scratch saved[pm::S_COUNT];
  StaticScratchScope(pm_ctx* c, unsigned char* base) : ctx(c) {
    for (int i = 0; i < pm::S_COUNT; i++) {
      saved[i] = c->scratch[i];
      c->scratch[i].ptr = base;
      c->scratch[i].cap = 0;
    }
    size_t off = 0;
    for (const StaticSlot& s : kStaticSlots) {
      c->scratch[s.slot].ptr = base + off;
      c->scratch[s.slot].cap = s.bytes;
      off += s.bytes;
    }
  }
  ~StaticScratchScope() {
    for (int i = 0; i < pm::S_COUNT; i++) ctx->scratch[i] = saved[i];
  }
};

}  // namespace

extern "C" int pm_knn_static(pm_ctx* ctx, const pm_map* map, const float* queries_dev, int64_t nq, int k, float cutoff_radius,
                             int32_t* out_ids_dev, float* out_dist2_dev, float* out_r2_dev) {
  PM_REQUIRE(ctx, "ctx is NULL");
  PM_REQUIRE(!map || !map->cells_only, "this map was built by pm_map_build_cells (no kd-tree): fixed-radius gathers only");
  PM_REQUIRE(k == 8 || k == 16 || k == 32 || k == 50 || k == 64, "k must be one of 8, 16, 32, 50, 64");
  PM_REQUIRE(nq >= 0, "nq is negative");
  if (nq > kStaticMaxQueries) {
    pm::set_error("pm_knn_static: nq = %lld exceeds PM_KNN_STATIC_MAX_QUERIES = %lld (the resident scratch array is sized for "
                  "that; split the batch or call pm_knn)", (long long)nq, (long long)kStaticMaxQueries);
    return PM_ERR_CAPACITY;
  }
  if (nq == 0) return PM_OK;
  PM_REQUIRE(queries_dev && out_ids_dev, "queries/out_ids is NULL");
  PM_CUDA(cudaSetDevice(ctx->device));
  /* host buffers would need a staging allocation: refused */
  PM_REQUIRE(pm::is_device_ptr(queries_dev) && pm::is_device_ptr(out_ids_dev) && (!out_dist2_dev || pm::is_device_ptr(out_dist2_dev)) &&
                 (!out_r2_dev || pm::is_device_ptr(out_r2_dev)),
             "queries and outputs must be device memory (this entry point allocates nothing, so it cannot stage host buffers)");
  PM_REQUIRE(!map || map->device == ctx->device, "the map lives on another device");
  const float cutoff2 = cutoff_radius * cutoff_radius;
  if (!map || map->n < 1) {
    k_fill_empty<<<(unsigned)pm::div_up(nq * k, 256), 256, 0, ctx->stream>>>(nq * k, out_ids_dev, out_dist2_dev, cutoff2);
    if (out_r2_dev) k_fill_f<<<(unsigned)pm::div_up(nq, 256), 256, 0, ctx->stream>>>(nq, out_r2_dev, cutoff2);
    ctx->launches += 2;
    PM_CUDA(cudaGetLastError());
    ctx->fb_count_dev = nullptr;
    ctx->ho_count_dev = nullptr;
    return finish_stats(ctx, map);
  }
  unsigned char* base = nullptr;
  PM_CHECK(static_scratch_base(ctx->device, &base));
  int r;
  {
    StaticScratchScope scope(ctx, base);
    pmk::QueryIO io;
    io.q = queries_dev; io.nq = (int)nq; io.k = k; io.cutoff2 = cutoff2;
    io.ids = out_ids_dev; io.d2 = out_dist2_dev; io.r2 = out_r2_dev;
    r = run_queries(ctx, map, io, false, true);
  }
  if (r != PM_OK) { cudaStreamSynchronize(ctx->stream); return r; }
  return finish_stats(ctx, map);
}
