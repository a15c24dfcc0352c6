/* pm_kdwarp.cu — stage 3, the density-independent path: warp-cooperative k-nearest-photon search in the
 * left-balanced kd-tree itself.
 *
 * Same result contract as the other stage-3 kernels (pm_knn.cu, pm_grid.cu; reference:
 * cukd::stackBased::knn<HeapCandidateList<k>>, externals/cudaKDTree/cukd/knn.h:298-364,536-549, driven by
 * ray-tracer/cuda/shading.h:11-18,93-121): per query the k smallest (floatBits(dist²), node id) pairs among nodes
 * with dist² < cutoff².
 *
 * Why it exists: the uniform grid answers a query only if its neighbourhood fits a block of cells. Photon maps of
 * real scenes are not like that — the caustic map is a few dense clusters, and most pixels query it from far away;
 * the reference's per-thread traversal (k_kd) answers those at ~30 M queries/s on B200 (dependent 16-byte loads, a
 * heap sift per accepted node, divergent lanes). Here ONE WARP walks the tree for one query:
 *   - the tree is stored in level order, so the descendants of node b at relative depth d are the contiguous
 *     range [(b << d) + 2^d - 1, ...): a whole subtree of <= 255 nodes ("bucket": the last 8 levels below a node
 *     of level numLevels - 8) is read with 8 coalesced LDG.128 steps instead of being traversed;
 *   - above the bucket level the walk is the reference's (near child first, far child pushed with its plane
 *     distance²), executed uniformly by the warp, stack in shared memory;
 *   - candidates closer than the current bound are ballot-compacted into a 256-entry shared buffer; when it fills
 *     up (and at the end) the k smallest are selected by histogram narrowing over dist² + exact ranking of the
 *     last <= 32 by 64-bit key, and the bound becomes the k-th key. Pruning keeps subtrees whose plane distance²
 *     EQUALS the bound's, so ties resolve by node id exactly as a brute-force search would (SURVEY §8c tie rule).
 */
#include <cfloat>
#include "pm_knn.cuh"

namespace {

using pmk::encode;

constexpr int kWarpsPerCta = 8;
constexpr int kCap = 256;            /* candidate buffer per warp */
constexpr int kSlotsW = kCap / 32;   /* keys per lane during a selection */
constexpr int kBucketLevels = 7;     /* a bucket = the last 8 levels below its root: <= 255 nodes */
constexpr int kBucketNodes = (1 << kBucketLevels) - 1;
constexpr int kStack = 40;

struct Box6f { float lo[3], hi[3]; };

struct __align__(16) KdwScratch {
  uint64_t keys[kCap];
  uint64_t win[kCap];      /* window members during a selection */
  int hist[32];
  int stack_node[kStack];
  float4 stack_box[kStack]; /* per-axis distance from the query to the subtree's domain box, and its dist² bound */
};

__device__ __forceinline__ uint64_t warp_max_u64(uint64_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const uint64_t w = __shfl_xor_sync(0xffffffffu, v, o);
    v = w > v ? w : v;
  }
  return v;
}

/* keys[0 .. cnt) -> the k smallest in keys[0 .. k) (any order); returns the largest of them. cnt > k, cnt <= kCap. */
__device__ __noinline__ uint64_t select_k(KdwScratch& ws, int cnt, int k, int lane) {
  const unsigned lt = (1u << lane) - 1u;
  uint64_t key[kSlotsW];
  int st[kSlotsW]; /* 0: in the window, 1: accepted, 2: rejected / empty */
#pragma unroll
  for (int s = 0; s < kSlotsW; s++) {
    const int idx = lane + 32 * s;
    st[s] = 2; key[s] = 0;
    if (idx < cnt) { key[s] = ws.keys[idx]; st[s] = 0; }
  }
  __syncwarp();
  int need = k, wcount = cnt;
  while (true) {
    if (wcount == need) { /* everything left is wanted */
#pragma unroll
      for (int s = 0; s < kSlotsW; s++) if (st[s] == 0) st[s] = 1;
      break;
    }
    /* dist² range of the window (non-negative floats order like their bit patterns) */
    uint32_t lo = 0xffffffffu, hi = 0u;
#pragma unroll
    for (int s = 0; s < kSlotsW; s++)
      if (st[s] == 0) { const uint32_t b = (uint32_t)(key[s] >> 32); lo = min(lo, b); hi = max(hi, b); }
    lo = __reduce_min_sync(0xffffffffu, lo);
    hi = __reduce_max_sync(0xffffffffu, hi);
    if (wcount <= 32 || lo == hi) {
      /* exact ranks inside the window by full key */
      int base = 0;
#pragma unroll
      for (int s = 0; s < kSlotsW; s++) {
        const bool in = st[s] == 0;
        const unsigned b = __ballot_sync(0xffffffffu, in);
        if (in) ws.win[base + __popc(b & lt)] = key[s];
        base += __popc(b);
      }
      __syncwarp();
      if (wcount <= 32) {
        uint64_t mine = ~0ull;
        if (lane < wcount) mine = ws.win[lane];
        int rank = 0;
        for (int j = 0; j < wcount; j++) rank += ws.win[j] < mine;
        /* write the verdict back to whoever holds the key: accepted iff rank < need */
        __syncwarp();
        if (lane < wcount) ws.win[lane] = rank < need ? 1ull : 0ull;
        __syncwarp();
        base = 0;
#pragma unroll
        for (int s = 0; s < kSlotsW; s++) {
          const bool in = st[s] == 0;
          const unsigned b = __ballot_sync(0xffffffffu, in);
          if (in) st[s] = ws.win[base + __popc(b & lt)] ? 1 : 2;
          base += __popc(b);
        }
      } else {
        /* more than 32 candidates at one distance: rank each against all window members */
#pragma unroll
        for (int s = 0; s < kSlotsW; s++) {
          if (st[s] == 0) {
            int rank = 0;
            for (int j = 0; j < wcount; j++) rank += ws.win[j] < key[s];
            st[s] = rank < need ? 1 : 2;
          }
        }
      }
      __syncwarp();
      break;
    }
    /* 32 linear bins over [lo, hi] */
    const float flo = __uint_as_float(lo), fhi = __uint_as_float(hi);
    const float scale = __fdividef(32.f, fhi - flo);
    ws.hist[lane] = 0;
    __syncwarp();
    int bin[kSlotsW];
#pragma unroll
    for (int s = 0; s < kSlotsW; s++) {
      bin[s] = 99;
      if (st[s] == 0) {
        bin[s] = min(max(__float2int_rd((__uint_as_float((uint32_t)(key[s] >> 32)) - flo) * scale), 0), 31);
        atomicAdd(&ws.hist[bin[s]], 1);
      }
    }
    __syncwarp();
    const int c = ws.hist[lane];
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    const int kbin = __popc(__ballot_sync(0xffffffffu, incl < need)); /* wcount > need: kbin <= 31 */
    const int below = __shfl_sync(0xffffffffu, incl - c, kbin);
    const int inbin = __shfl_sync(0xffffffffu, c, kbin);
#pragma unroll
    for (int s = 0; s < kSlotsW; s++)
      if (st[s] == 0) st[s] = bin[s] < kbin ? 1 : (bin[s] == kbin ? 0 : 2);
    need -= below;
    /* the bins are linear in a float difference: lo == hi was excluded, so at least the extremes separate and the
     * window shrinks unless everything fell into one bin by rounding — then fall back to exact ranking next round */
    if (inbin == wcount) {
      /* no progress (values closer than the bin arithmetic resolves): narrow by bit patterns instead */
      const uint32_t mid = lo + ((hi - lo) >> 1);
      int cl = 0;
#pragma unroll
      for (int s = 0; s < kSlotsW; s++) cl += st[s] == 0 && (uint32_t)(key[s] >> 32) <= mid;
      cl = __reduce_add_sync(0xffffffffu, cl);
      if (cl >= need) {
#pragma unroll
        for (int s = 0; s < kSlotsW; s++) if (st[s] == 0 && (uint32_t)(key[s] >> 32) > mid) st[s] = 2;
        wcount = cl;
      } else {
#pragma unroll
        for (int s = 0; s < kSlotsW; s++) if (st[s] == 0 && (uint32_t)(key[s] >> 32) <= mid) st[s] = 1;
        need -= cl;
        wcount -= cl;
      }
    } else {
      wcount = inbin;
    }
    if (need == 0) {
#pragma unroll
      for (int s = 0; s < kSlotsW; s++) if (st[s] == 0) st[s] = 2;
      break;
    }
    __syncwarp();
  }
  /* compact the accepted keys to the front; the bound is their maximum */
  int base = 0;
  uint64_t mx = 0;
#pragma unroll
  for (int s = 0; s < kSlotsW; s++) {
    const bool a = st[s] == 1;
    const unsigned b = __ballot_sync(0xffffffffu, a);
    if (a) { ws.keys[base + __popc(b & lt)] = key[s]; mx = key[s] > mx ? key[s] : mx; }
    base += __popc(b);
  }
  __syncwarp();
  return warp_max_u64(mx);
}

/* one warp per query of list[0 .. *list_count) (or order[t] / t when list == NULL) */
template <bool GATHER>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 3)
k_kdw(const float4* __restrict__ nodes, const float4* __restrict__ color4, int n, int first_bucket_node, const Box6f box,
      pmk::QueryIO io, const int* __restrict__ list, const int* __restrict__ list_count) {
  __shared__ KdwScratch scratch[kWarpsPerCta];
  const int wid = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); /* warp-uniform to the compiler */
  KdwScratch& ws = scratch[wid];
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  const int total = list ? *list_count : io.nq;
  const int k = io.k;
  for (int t = blockIdx.x * kWarpsPerCta + wid; t < total; t += gridDim.x * kWarpsPerCta) {
    /* loaded values pass through a shuffle so that the compiler sees them as warp-uniform */
    const int qi = __shfl_sync(0xffffffffu, list ? list[t] : (io.order ? io.order[t] : t), 0);
    float brdf = 0.f;
    if (GATHER) {
      brdf = __shfl_sync(0xffffffffu, io.brdf[qi], 0);
      if (!(brdf > 0.f)) {
        if (lane < 3) io.rgb[(size_t)qi * 3 + lane] = 0.f;
        continue;
      }
    }
    const float qx = io.q[(size_t)qi * 3 + 0], qy = io.q[(size_t)qi * 3 + 1], qz = io.q[(size_t)qi * 3 + 2];
    uint64_t bound = encode(io.cutoff2, 0); /* candidates must have key < bound */
    int cnt = 0, sp = 0, curr = 0;
    /* per-axis distance from the query to the domain box of `curr` (root: the map's bounds). Culling by the box
     * distance instead of the reference's single plane distance is what keeps far-away queries (a caustic map seen
     * from the other side of the room) from visiting most of the tree; the factor below absorbs the different
     * rounding of this sum of squares and pmk::dist2, so the bound stays conservative */
    float ox = fmaxf(fmaxf(box.lo[0] - qx, qx - box.hi[0]), 0.f), oy = fmaxf(fmaxf(box.lo[1] - qy, qy - box.hi[1]), 0.f),
          oz = fmaxf(fmaxf(box.lo[2] - qz, qz - box.hi[2]), 0.f);
    while (true) {
      while (curr < n) {
        if (curr >= first_bucket_node) {
          /* bucket: node j of the subtree (level order inside it) is array element (curr << d) + j, d = floor(log2(j + 1)) */
          for (int j0 = 0; j0 < kBucketNodes; j0 += 32) {
            const int j = j0 + lane;
            const int d = 31 - __clz(j + 1);
            const long long idx = ((long long)curr << d) + j;
            if (((long long)curr << (31 - __clz(j0 + 1))) + j0 >= n) break; /* this and all deeper levels are absent */
            bool pass = false;
            uint64_t key = 0;
            if (j < kBucketNodes && idx < n) {
              const float4 nd = __ldg(nodes + idx);
              key = encode(pmk::dist2(nd.x, nd.y, nd.z, qx, qy, qz), (int)idx);
              pass = key < bound;
            }
            const unsigned b = __ballot_sync(0xffffffffu, pass);
            if (pass) ws.keys[cnt + __popc(b & lt)] = key;
            cnt += __popc(b);
            if (cnt > kCap - 32) {
              __syncwarp();
              bound = select_k(ws, cnt, k, lane);
              cnt = k;
            }
          }
          break;
        }
        const float4 nd = __ldg(nodes + curr);
        const uint64_t key = encode(pmk::dist2(nd.x, nd.y, nd.z, qx, qy, qz), curr);
        if (key < bound) {
          if (lane == 0) ws.keys[cnt] = key;
          cnt++;
          if (cnt > kCap - 32) {
            __syncwarp();
            bound = select_k(ws, cnt, k, lane);
            cnt = k;
          }
        }
        const int dim = __float_as_int(nd.w);
        const float dp = dim == 0 ? qx - nd.x : (dim == 1 ? qy - nd.y : qz - nd.z);
        const int l = 2 * curr + 1;
        const int far_child = dp < 0.f ? l + 1 : l;
        const float ad = fabsf(dp);
        const float fx = dim == 0 ? fmaxf(ox, ad) : ox, fy = dim == 1 ? fmaxf(oy, ad) : oy, fz = dim == 2 ? fmaxf(oz, ad) : oz;
        const float fd2 = (fx * fx + fy * fy + fz * fz) * 0.999999f;
        /* the far subtree can only hold keys >= encode(fd2, 0) */
        if (far_child < n && encode(fd2, 0) < bound && sp < kStack) {
          if (lane == 0) { ws.stack_node[sp] = far_child; ws.stack_box[sp] = make_float4(fx, fy, fz, fd2); }
          sp++;
        }
        curr = dp < 0.f ? l : l + 1;
      }
      __syncwarp();
      bool found = false;
      while (sp > 0) {
        --sp;
        const float4 b4 = ws.stack_box[sp];
        if (encode(b4.w, 0) < bound) { curr = ws.stack_node[sp]; ox = b4.x; oy = b4.y; oz = b4.z; found = true; break; }
      }
      __syncwarp();
      if (!found) break;
    }
    if (cnt > k) { bound = select_k(ws, cnt, k, lane); cnt = k; }
    const int kk = cnt; /* <= k entries in keys[0 .. kk) */
    float r2 = io.cutoff2;
    if (kk == k) {
      uint64_t mx = 0;
      for (int p = lane; p < kk; p += 32) { const uint64_t e = ws.keys[p]; mx = e > mx ? e : mx; }
      r2 = __uint_as_float((uint32_t)(warp_max_u64(mx) >> 32));
    }
    if (!GATHER) {
      if (io.ids) {
        int2* ids2 = reinterpret_cast<int2*>(io.ids + (size_t)qi * k);
        float2* d22 = io.d2 ? reinterpret_cast<float2*>(io.d2 + (size_t)qi * k) : nullptr;
        for (int p = 2 * lane; p < k; p += 64) {
          int2 iv = make_int2(-1, -1);
          float2 dv = make_float2(io.cutoff2, io.cutoff2); /* unfilled slots keep (cutoff², -1) (knn.h:298-306) */
          if (p < kk) { const uint64_t e = ws.keys[p]; iv.x = (int)(uint32_t)e; dv.x = __uint_as_float((uint32_t)(e >> 32)); }
          if (p + 1 < kk) { const uint64_t e = ws.keys[p + 1]; iv.y = (int)(uint32_t)e; dv.y = __uint_as_float((uint32_t)(e >> 32)); }
          ids2[p >> 1] = iv;
          if (d22) d22[p >> 1] = dv;
        }
      }
      if (io.r2 && lane == 0) io.r2[qi] = r2;
    } else {
      /* gatherPhotons (ray-tracer/cuda/shading.h:93-121), summed across the warp */
      const float rad = __fsqrt_rn(r2);
      float fr = 0.f, fg = 0.f, fb = 0.f;
      for (int p = lane; p < kk; p += 32) {
        const uint64_t e = ws.keys[p];
        const float4 c = __ldg(color4 + (uint32_t)e);
        const float dist = __fsqrt_rn(__uint_as_float((uint32_t)(e >> 32)));
        const float w = __fsub_rn(1.f, __fmul_rn(__fdiv_rn(dist, rad), 1.1f));
        const float sc = __fmul_rn(__fmul_rn(brdf, c.w), w);
        fr = __fadd_rn(fr, __fmul_rn(sc, c.x));
        fg = __fadd_rn(fg, __fmul_rn(sc, c.y));
        fb = __fadd_rn(fb, __fmul_rn(sc, c.z));
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        fr = __fadd_rn(fr, __shfl_xor_sync(0xffffffffu, fr, o));
        fg = __fadd_rn(fg, __shfl_xor_sync(0xffffffffu, fg, o));
        fb = __fadd_rn(fb, __shfl_xor_sync(0xffffffffu, fb, o));
      }
      const float k1 = (1 - (2.f / 3.f) * (1.f / 1.1f));
      const float denom = __fmul_rn(__fmul_rn(__fmul_rn(k1, 2.f), 3.14159274101257324f), r2);
      if (lane == 0) {
        io.rgb[(size_t)qi * 3 + 0] = __fdiv_rn(fr, denom);
        io.rgb[(size_t)qi * 3 + 1] = __fdiv_rn(fg, denom);
        io.rgb[(size_t)qi * 3 + 2] = __fdiv_rn(fb, denom);
      }
    }
    __syncwarp();
  }
}

}  // namespace

namespace pm {

int kdw_query(pm_ctx* ctx, const pm_map* map, const pmk::QueryIO& io, bool gather, const int* list, const int* list_count) {
  const int n = (int)map->n;
  int levels = 0;
  while (((int64_t)1 << levels) <= n) levels++; /* floor(log2 n) + 1 */
  const int lb = levels > kBucketLevels ? levels - kBucketLevels : 0;
  const int first_bucket_node = (1 << lb) - 1;
  int64_t grid = div_up(io.nq, kWarpsPerCta);
  const int64_t cap = (int64_t)ctx->sm_count * (list ? 16 : 64);
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  Box6f box;
  for (int d = 0; d < 3; d++) { box.lo[d] = map->bounds[d]; box.hi[d] = map->bounds[3 + d]; }
  if (gather) k_kdw<true><<<(unsigned)grid, kWarpsPerCta * 32, 0, ctx->stream>>>(map->nodes4, map->color4, n, first_bucket_node, box, io, list, list_count);
  else k_kdw<false><<<(unsigned)grid, kWarpsPerCta * 32, 0, ctx->stream>>>(map->nodes4, map->color4, n, first_bucket_node, box, io, list, list_count);
  ctx->launches++;
  PM_CUDA(cudaGetLastError());
  return PM_OK;
}

}  // namespace pm
