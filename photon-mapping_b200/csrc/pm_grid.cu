/* pm_grid.cu — stage 3 fast path: warp-cooperative k-nearest-photon search over a uniform grid.
 *
 * Same result contract as the kd-tree kernels of pm_knn.cu (which restate
 * cukd::stackBased::knn<HeapCandidateList<k>>, externals/cudaKDTree/cukd/knn.h:298-364,536-549):
 * per query the k smallest (floatBits(dist²), node id) pairs among nodes with dist² < cutoff²,
 * ids being positions in the left-balanced kd-tree array that pm_map_build produced. Only the way
 * the candidates are found differs.
 *
 * Why not the tree: a per-thread traversal is a chain of dependent, divergent 16-byte loads plus a
 * heap sift per accepted node; on B200 it runs at ~1 % of the HBM roofline. Here
 *   - the map keeps a second copy of the node positions, sorted by uniform-grid cell (x fastest), so
 *     the 5x5x5-cell neighbourhood of a query is 25 contiguous runs of (x, y, z, id) records;
 *   - one warp answers one query: the 25 runs are flattened into one index space (a bitmap of run
 *     starts in shared memory turns "which run is element j in" into one popc), every lane pulls
 *     one record per step with a coalesced LDG.128, computes dist² and compacts the candidates
 *     inside the search radius into shared memory with a ballot;
 *   - the k smallest are found by COUNTING, not by a heap: two thresholds placed by interpolating
 *     the count/dist² relation bracket the k-th candidate to <= 32 candidates in (almost always) one
 *     sweep over <= 5 registers per lane; the bracket is ranked exactly on full 64-bit keys, one
 *     candidate per lane;
 *   - exactness: every node closer than G (the distance from the query to the nearest face of the
 *     5x5x5 block that is not also a face of the whole grid, less a rounding margin) lies inside the
 *     block, so once k candidates with dist² < G² exist the k nearest are among them. Queries for
 *     which that fails (sparse neighbourhood), or whose neighbourhood is too dense for the shared
 *     buffers, are appended to a list that the kd-tree kernel answers.
 * The cell size comes from a sample of exact k-th-neighbour distances taken when the map is built.
 */
#include <cub/cub.cuh>
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include "pm_knn.cuh"

namespace {

using pmk::encode;

constexpr int kWarps = 8;          /* warps (= queries in flight) per CTA */
constexpr int kBuf = 128;          /* compacted candidates per query */
constexpr int kSlots = kBuf / 32;  /* candidates per lane in the selection phase */
constexpr int kMaxScan = 1024;     /* nodes in the 5x5x5 block beyond which the kd-tree kernel takes over */
constexpr int kWords = kMaxScan / 32;
constexpr int kMaxH = 8;          /* widest block tried: (2 * 8 + 1)^3 cells */
constexpr int kMaxAttempts = 12;
constexpr int kTuneK = 64;         /* the cell size is tuned for the largest supported k */
constexpr int kTuneSamples = 2048;
constexpr int64_t kMinGridNodes = 4096;
/* cells of the offset table (4 bytes each): at least 2^24, up to 32 per photon, never more than 2^30 (4 GB).
 * Photons on surfaces occupy a vanishing share of a 3-D grid, so a dense table at the wanted resolution can be out
 * of reach (C3: 92 M photons on the Cornell walls want 4e10 cells); the cap then coarsens the cells and the
 * ball-trimmed rows keep the candidate count in check. */
constexpr int64_t kMinCellCap = (int64_t)1 << 24, kMaxCellCap = (int64_t)1 << 30;

struct GridView {
  const float4* pts;
  const uint32_t* cell_start;
  const uint32_t* cell_start_t; /* fast-path maps: the same offsets with x slowest, [(x * nz + z) * ny + y], x in [0, nx] */
  float lo[3];
  float cell, inv_cell, margin;
  int nx, ny, nz;
};

struct __align__(16) WarpScratch {
  uint64_t keys[kBuf];
  uint2 words[kWords]; /* x: bitmap of run starts among elements [32 i, 32 i + 32); y: runs started before 32 i */
  int run_off[32];     /* per non-empty run: (first record of the run) - (its first flat index) */
};

__device__ __forceinline__ int cell_of(float v, float lo, float inv_cell, int n) {
  const float u = (v - lo) * inv_cell;
  /* fminf/fmaxf drop NaN; the clamp keeps out-of-grid queries on the boundary cells */
  return (int)fminf(fmaxf(floorf(u), 0.f), (float)(n - 1));
}

/* x / d for x < 1024 and 1 <= d <= 17 by multiplication with 2^16 / d + 1 (checked exhaustively):
 * which row of a block a lane owns */
__constant__ int c_rdiv[18] = {0, 65537, 32769, 21846, 16385, 13108, 10923, 9363, 8193, 7282, 6554, 5958, 5462,
                               5042, 4682, 4370, 4097, 3856};
__device__ __forceinline__ int div_small(int x, int d) { return (x * c_rdiv[d]) >> 16; }

/* cell index of coordinate v: the same floor((v - lo) * inv_cell) the build uses, clamped */
__device__ __forceinline__ int cell_idx(float v, float lo, float inv_cell, int n) {
  return min(max(__float2int_rd((v - lo) * inv_cell), 0), n - 1);
}

/* x^(2/3) for the radius guesses: two MUFU ops, accuracy is irrelevant there */
__device__ __forceinline__ float pow23(float x) { return __powf(x, 2.f / 3.f); }

template <bool GATHER>
__global__ void __launch_bounds__(kWarps * 32, 5)
k_grid_query(const GridView g, const float4* __restrict__ color4, const pmk::QueryIO io, const int* __restrict__ nq_dev,
             int* __restrict__ fb_list, int* __restrict__ fb_count) {
  __shared__ WarpScratch scratch[kWarps];
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u, le_not0 = (lt | (1u << lane)) & ~1u;
  /* the warp index through a shuffle: the compiler then knows it (and every branch on the query index) to be
   * warp-uniform and drops the divergence guards around the warp-level primitives below */
  const int wid = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  WarpScratch& ws = scratch[wid];
  const int k = io.k;
  const float cell3 = g.cell * g.cell * g.cell;

  const int total = nq_dev ? *nq_dev : io.nq; /* a hand-over list has a device-side length */
  for (int t = blockIdx.x * kWarps + wid; t < total; t += gridDim.x * kWarps) {
    /* loaded values pass through a shuffle so that the compiler sees them as warp-uniform */
    const int qi = __shfl_sync(0xffffffffu, io.order ? io.order[t] : t, 0);
    float brdf = 0.f;
    if (GATHER) {
      brdf = __shfl_sync(0xffffffffu, io.brdf[qi], 0);
      if (!(brdf > 0.f)) {
        if (lane < 3) io.rgb[(size_t)qi * 3 + lane] = 0.f;
        continue;
      }
    }
    const float qx = io.q[(size_t)qi * 3 + 0], qy = io.q[(size_t)qi * 3 + 1], qz = io.q[(size_t)qi * 3 + 2];
    const float fx = (qx - g.lo[0]) * g.inv_cell, fy = (qy - g.lo[1]) * g.inv_cell, fz = (qz - g.lo[2]) * g.inv_cell;
    const int cx = min(max(__float2int_rd(fx), 0), g.nx - 1), cy = min(max(__float2int_rd(fy), 0), g.ny - 1),
              cz = min(max(__float2int_rd(fz), 0), g.nz - 1);

    /* ---- candidate search: block of (2h+1)^3 cells, ball of radius sqrt(R2) inside it ------ */
    int h = 2, n = 0;
    float R2 = 0.f, R2lo = 0.f, R2hi = INFINITY;
    bool give_up = false, take_max = false;
    for (int attempt = 0;; attempt++) {
      const int x0 = max(cx - h, 0), x1 = min(cx + h, g.nx - 1);
      const int y0 = max(cy - h, 0), y1 = min(cy + h, g.ny - 1);
      const int z0 = max(cz - h, 0), z1 = min(cz + h, g.nz - 1);
      /* G: distance to the nearest block face that is not a face of the grid itself (measured in
       * cells first; the margin covers the rounding of both this and the build's cell assignment) */
      float Gc = INFINITY;
      if (x0 > 0) Gc = fminf(Gc, fx - (float)x0);
      if (x1 < g.nx - 1) Gc = fminf(Gc, (float)(x1 + 1) - fx);
      if (y0 > 0) Gc = fminf(Gc, fy - (float)y0);
      if (y1 < g.ny - 1) Gc = fminf(Gc, (float)(y1 + 1) - fy);
      if (z0 > 0) Gc = fminf(Gc, fz - (float)z0);
      if (z1 < g.nz - 1) Gc = fminf(Gc, (float)(z1 + 1) - fz);
      const float G = Gc * g.cell - g.margin;
      const float Rmax2 = fminf(G > 0.f ? G * G : 0.f, io.cutoff2);
      const int ny_r = y1 - y0 + 1, nrows = ny_r * (z1 - z0 + 1);
      if (attempt == 0) {
        /* first radius from the block's mean density: expect ~1.3 k + 12 nodes inside it */
        int c = 0;
        if (lane < nrows) {
          const int rzo = div_small(lane, ny_r);
          const size_t base = ((size_t)(z0 + rzo) * g.ny + (y0 + lane - rzo * ny_r)) * g.nx;
          c = (int)(__ldg(g.cell_start + base + x1 + 1) - __ldg(g.cell_start + base + x0));
        }
        const int m0 = __reduce_add_sync(0xffffffffu, c);
        R2 = Rmax2;
        if (m0 > 0) {
          const float vol = (float)((x1 - x0 + 1) * nrows) * cell3;
          R2 = fminf(Rmax2, pow23(__fdividef((1.3f * (float)k + 12.f) * vol, 4.18879f * (float)m0)));
        }
      } else if (take_max) {
        R2 = Rmax2;
        take_max = false;
      }

      /* rows of the block, 32 at a time: trim each to the ball, flatten, scan */
      n = 0;
      bool dense = false;
      for (int r0 = 0; r0 < nrows; r0 += 32) {
        const int row = r0 + lane;
        int len = 0, st = 0;
        if (row < nrows) {
          const int rzo = div_small(row, ny_r);
          const int ry = y0 + row - rzo * ny_r, rz = z0 + rzo;
          const float ylo = g.lo[1] + (float)ry * g.cell, zlo = g.lo[2] + (float)rz * g.cell;
          const float dy = fmaxf(fmaxf(ylo - qy, qy - (ylo + g.cell)) - g.margin, 0.f);
          const float dz = fmaxf(fmaxf(zlo - qz, qz - (zlo + g.cell)) - g.margin, 0.f);
          const float dyz2 = dy * dy + dz * dz;
          if (dyz2 < R2) {
            const float t2 = R2 - dyz2; /* > 0; half-width of the ball in this row, rounded up generously */
            const float w = (t2 < 1e30f ? t2 * rsqrtf(t2) * 1.0001f : INFINITY) + g.margin;
            const int xa = max(cell_idx(qx - w, g.lo[0], g.inv_cell, g.nx), x0);
            const int xb = min(cell_idx(qx + w, g.lo[0], g.inv_cell, g.nx), x1);
            if (xa <= xb) {
              const size_t base = ((size_t)rz * g.ny + ry) * g.nx;
              st = (int)__ldg(g.cell_start + base + xa);
              len = (int)__ldg(g.cell_start + base + xb + 1) - st;
            }
          }
        }
        int incl = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int v = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += v;
        }
        const int mc = __shfl_sync(0xffffffffu, incl, 31);
        if (mc == 0) continue;
        if (mc > kMaxScan) { dense = true; break; }
        /* flat index space of this chunk: words[i].x = bitmap of run starts among elements
         * [32 i, 32 i + 32), words[i].y = the run holding element 32 i, run_off = address offsets */
        const int P = incl - len;
        __syncwarp();
        ws.words[lane] = make_uint2(0u, 0u);
        __syncwarp();
        const unsigned nonempty = __ballot_sync(0xffffffffu, len > 0);
        if (len > 0) {
          const int d = __popc(nonempty & lt);
          atomicOr(&ws.words[P >> 5].x, 1u << (P & 31));
          ws.run_off[d] = st - P;
          for (int i = (P + 31) >> 5; i * 32 < P + len; i++) ws.words[i].y = (unsigned)d;
        }
        __syncwarp();
        for (int i = 0; i * 32 < mc; i++) {
          const int j = i * 32 + lane;
          const int jj = min(j, mc - 1); /* lanes past the end re-read the last record and drop it */
          const uint2 w = ws.words[i];
          const int run = (int)w.y + __popc(w.x & le_not0);
          const float4 p = __ldg(g.pts + (ws.run_off[run] + jj));
          float d2 = pmk::dist2(p.x, p.y, p.z, qx, qy, qz);
          if (j >= mc) d2 = INFINITY;
          const bool pass = d2 < R2;
          const unsigned b = __ballot_sync(0xffffffffu, pass);
          /* beyond the buffer everything lands in the last slot; n > kBuf is detected below */
          if (pass) ws.keys[min(n + __popc(b & lt), kBuf - 1)] = encode(d2, __float_as_int(p.w));
          n += __popc(b);
        }
      }
      if (dense) { give_up = true; break; }
      if (n >= k && n <= kBuf) break;
      if (n < k && !(R2 < io.cutoff2)) break; /* the whole cutoff ball has been searched */
      if (attempt >= kMaxAttempts) { give_up = true; break; }
      /* adjust the radius inside the bracket (R2lo: too few, R2hi: more than the buffer holds) */
      if (n > kBuf) {
        R2hi = R2;
        float c = R2 * pow23(__fdividef(1.2f * (float)k + 8.f, (float)n));
        if (!(c > R2lo)) c = sqrtf(R2lo * R2hi);
        R2 = c;
      } else {
        R2lo = R2;
        const float c = R2 * pow23(__fdividef(1.4f * (float)k + 16.f, (float)max(n, k >> 2)));
        if (c < R2hi) {
          if (c < Rmax2) R2 = c;
          else if (R2 < Rmax2) take_max = true;           /* everything this block guarantees */
          else if (h < kMaxH) { h++; take_max = true; }   /* a wider block */
          else { give_up = true; break; }
        } else {
          R2 = sqrtf(R2lo * R2hi);
        }
      }
    }
    if (give_up) {
      if (lane == 0) fb_list[atomicAdd(fb_count, 1)] = qi;
      __syncwarp();
      continue;
    }
    __syncwarp();

    /* ---- selection: the kk smallest keys among the n candidates --------------------------
     * One 32-bin histogram over dist² (linear bins between the distances at which k -/+ 3 sigma
     * candidates are expected, everything nearer in bin 0, everything farther in bin 31) finds the
     * bin holding the k-th candidate; that bin's few candidates are ranked exactly by 64-bit key. */
    const int kk = min(n, k);
    const int ns = (n + 31) >> 5; /* occupied register slots per lane */
    int hb[kSlots], bin[kSlots];
    uint32_t idv[kSlots];
#pragma unroll
    for (int s = 0; s < kSlots; s++) {
      const int idx = lane + 32 * s;
      hb[s] = 0x7fffffff; idv[s] = 0xffffffffu; bin[s] = 64; /* never accepted */
      if (s < ns && idx < n) { const uint64_t e = ws.keys[idx]; hb[s] = (int)(e >> 32); idv[s] = (uint32_t)e; bin[s] = 0; }
    }
    __syncwarp();
    int kbin = 1, clo = n; /* n <= k: bin 0 = everything is accepted */
    int nb = 0, rank = 0;
    uint64_t mine = ~0ull;
    if (n > k) {
      int mx = 0;
#pragma unroll
      for (int s = 0; s < kSlots; s++) if (s < ns && bin[s] == 0) mx = max(mx, hb[s]);
      const float fhi = __int_as_float(__reduce_max_sync(0xffffffffu, mx));
      /* count(dist² < T) ~ n (T / fhi)^1.5; its spread at the k-th candidate is binomial */
      const float fk = (float)k, fn = (float)n;
      const float delta = 3.f * sqrtf(fk * (1.f - __fdividef(fk, fn))) + 2.f;
      const float Ta = fhi * pow23(__fdividef(fmaxf(fk - delta, 0.5f), fn));
      const float Tb = fhi * pow23(__fdividef(fminf(fk + delta, fn - 0.5f), fn));
      const float scale = Tb > Ta ? __fdividef(30.f, Tb - Ta) : 0.f;
      ws.run_off[lane] = 0; /* the histogram */
      __syncwarp();
#pragma unroll
      for (int s = 0; s < kSlots; s++) {
        if (s < ns && bin[s] == 0) {
          /* monotone in dist²: subtraction, multiplication by a positive constant, floor, clamp */
          bin[s] = min(max(__float2int_rd((__int_as_float(hb[s]) - Ta) * scale) + 1, 0), 31);
          atomicAdd(&ws.run_off[bin[s]], 1);
        }
      }
      __syncwarp();
      const int cnt = ws.run_off[lane];
      int incl = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      kbin = __popc(__ballot_sync(0xffffffffu, incl < k)); /* first bin whose running count reaches k */
      clo = __shfl_sync(0xffffffffu, incl - cnt, kbin);
      nb = __shfl_sync(0xffffffffu, cnt, kbin);
      if (nb > 32) {
        /* rare (the k-th candidate fell outside the window, or many equal distances): narrow the
         * bracket (tlo, thi] of dist² bit patterns with pairs of counting thresholds */
        int a = -1, b2 = 0;
#pragma unroll
        for (int s = 0; s < kSlots; s++) {
          if (s < ns && bin[s] < kbin) a = max(a, hb[s]);
          if (s < ns && bin[s] <= kbin) b2 = max(b2, hb[s]);
        }
        int tlo = __reduce_max_sync(0xffffffffu, a), thi = __reduce_max_sync(0xffffffffu, b2), chi = clo + nb;
        float flo = tlo < 0 ? 0.f : __int_as_float(tlo), fh = __int_as_float(thi);
        int it = 0;
        while (chi - clo > 32) {
          if (thi - tlo <= 2) { give_up = true; break; } /* > 32 candidates at one distance: degenerate */
          int ta, tb;
          if (it < 6) {
            const float bw = (float)(chi - clo);
            const float dl = fmaxf(0.85f * bw * rsqrtf(bw), 2.f);
            const float fa = __fdividef(fmaxf(fk + 0.5f - dl, (float)clo + 0.5f) - (float)clo, bw);
            const float fb = __fdividef(fminf(fk + 0.5f + dl, (float)chi - 0.5f) - (float)clo, bw);
            ta = min(max(__float_as_int(flo + (fh - flo) * fa), tlo + 1), thi - 2);
            tb = min(max(__float_as_int(flo + (fh - flo) * fb), ta + 1), thi - 1);
          } else {
            const int w3 = (thi - tlo) / 3;
            ta = tlo + w3; tb = tlo + 2 * w3;
          }
          int ca = 0, cb = 0;
#pragma unroll
          for (int s = 0; s < kSlots; s++) if (s < ns) { ca += hb[s] <= ta; cb += hb[s] <= tb; }
          ca = __reduce_add_sync(0xffffffffu, ca);
          cb = __reduce_add_sync(0xffffffffu, cb);
          if (k <= ca) { thi = ta; chi = ca; fh = __int_as_float(ta); }
          else if (k <= cb) { tlo = ta; clo = ca; flo = __int_as_float(ta); thi = tb; chi = cb; fh = __int_as_float(tb); }
          else { tlo = tb; clo = cb; flo = __int_as_float(tb); }
          it++;
        }
        if (give_up) {
          if (lane == 0) fb_list[atomicAdd(fb_count, 1)] = qi;
          __syncwarp();
          continue;
        }
        nb = chi - clo;
        kbin = 1;
#pragma unroll
        for (int s = 0; s < kSlots; s++)
          if (s < ns && bin[s] < 64) bin[s] = hb[s] <= tlo ? 0 : (hb[s] <= thi ? 1 : 2);
      }
      /* the bracket: nb <= 32 candidates, one per lane, exact rank by key */
      int filled = 0;
#pragma unroll
      for (int s = 0; s < kSlots; s++) {
        if (s < ns) {
          const bool inb = bin[s] == kbin;
          const unsigned b = __ballot_sync(0xffffffffu, inb);
          if (inb) ws.keys[filled + __popc(b & lt)] = ((uint64_t)(uint32_t)hb[s] << 32) | idv[s];
          filled += __popc(b);
        }
      }
      __syncwarp();
      if (lane < nb) mine = ws.keys[lane];
      for (int j = 0; j < nb; j++) rank += ws.keys[j] < mine;
      __syncwarp();
    }
    const int need = kk - clo; /* accepted from the bracket (0 when n <= k) */
    const bool acc_b = lane < nb && rank < need;
    /* returned radius: the k-th smallest dist² (the last accepted bracket entry), cutoff² if not full */
    float r2 = io.cutoff2;
    if (kk == k) {
      int r2b = 0;
      if (n > k) {
        const unsigned last = __ballot_sync(0xffffffffu, acc_b && rank == need - 1);
        r2b = __shfl_sync(0xffffffffu, (int)(mine >> 32), __ffs(last) - 1);
      } else {
#pragma unroll
        for (int s = 0; s < kSlots; s++) if (s < ns && bin[s] == 0) r2b = max(r2b, hb[s]);
        r2b = __reduce_max_sync(0xffffffffu, r2b);
      }
      r2 = __int_as_float(r2b);
    }

    /* ---- output ------------------------------------------------------------------- */
    if (!GATHER) {
      /* accepted entries -> shared memory in slot order, then 8-byte coalesced row stores */
      int base = 0;
#pragma unroll
      for (int s = 0; s < kSlots; s++) {
        if (s < ns) {
          const bool a = bin[s] < kbin;
          const unsigned b = __ballot_sync(0xffffffffu, a);
          if (a) ws.keys[base + __popc(b & lt)] = ((uint64_t)(uint32_t)hb[s] << 32) | idv[s];
          base += __popc(b);
        }
      }
      if (acc_b) ws.keys[clo + rank] = mine;
      __syncwarp();
      int2* ids2 = reinterpret_cast<int2*>(io.ids + (size_t)qi * k); /* k is even: rows are 8-byte aligned */
      float2* d22 = io.d2 ? reinterpret_cast<float2*>(io.d2 + (size_t)qi * k) : nullptr;
      for (int p = 2 * lane; p < k; p += 64) {
        /* unfilled slots of the candidate list keep (cutoff², -1) (knn.h:298-306) */
        int2 iv = make_int2(-1, -1);
        float2 dv = make_float2(io.cutoff2, io.cutoff2);
        if (p < kk) { const uint64_t e = ws.keys[p]; iv.x = (int)(uint32_t)e; dv.x = __uint_as_float((uint32_t)(e >> 32)); }
        if (p + 1 < kk) { const uint64_t e = ws.keys[p + 1]; iv.y = (int)(uint32_t)e; dv.y = __uint_as_float((uint32_t)(e >> 32)); }
        ids2[p >> 1] = iv;
        if (d22) d22[p >> 1] = dv;
      }
      if (io.r2 && lane == 0) io.r2[qi] = r2;
    } else {
      /* gatherPhotons (ray-tracer/cuda/shading.h:93-121), summed across the warp. Division and square root are the
       * call-free fast paths of pm_knn.cuh; a query with an operand outside their range goes to the kd-tree kernel */
      const float rad = pmk::sqrt_rn_fast(r2);
      /* rad is checked once; an accepted distance is 0 or lies in [2^-50.5, rad] as soon as its square passes the
       * square root's own test (>= 2^-101), which is then all div_fast_ok(dist, rad) would ask */
      bool ok = pmk::sqrt_fast_ok(r2) && pmk::div_fast_ok(rad, rad);
      float fr = 0.f, fg = 0.f, fb = 0.f;
#pragma unroll
      for (int s = 0; s <= kSlots; s++) {
        bool a; int hv; uint32_t id;
        if (s < kSlots) { a = s < ns && bin[s] < kbin; hv = hb[s]; id = idv[s]; }
        else { a = acc_b; hv = (int)(mine >> 32); id = (uint32_t)mine; }
        if (a) {
          const float4 c = __ldg(color4 + id);
          const float d2v = __int_as_float(hv);
          const float dist = pmk::sqrt_rn_fast(d2v);
          ok = ok && pmk::sqrt_fast_ok(d2v);
          const float w = __fsub_rn(1.f, __fmul_rn(pmk::div_rn_fast(dist, rad), 1.1f));
          const float sc = __fmul_rn(__fmul_rn(brdf, c.w), w);
          fr = __fadd_rn(fr, __fmul_rn(sc, c.x));
          fg = __fadd_rn(fg, __fmul_rn(sc, c.y));
          fb = __fadd_rn(fb, __fmul_rn(sc, c.z));
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        fr = __fadd_rn(fr, __shfl_xor_sync(0xffffffffu, fr, o));
        fg = __fadd_rn(fg, __shfl_xor_sync(0xffffffffu, fg, o));
        fb = __fadd_rn(fb, __shfl_xor_sync(0xffffffffu, fb, o));
      }
      const float k1 = (1 - (2.f / 3.f) * (1.f / 1.1f));
      const float denom = __fmul_rn(__fmul_rn(__fmul_rn(k1, 2.f), 3.14159274101257324f), r2);
      ok = ok && pmk::div_fast_ok(fr, denom) && pmk::div_fast_ok(fg, denom) && pmk::div_fast_ok(fb, denom);
      if (!__all_sync(0xffffffffu, ok)) { /* exotic operands: the kd-tree kernel divides with the IEEE intrinsics */
        if (lane == 0) fb_list[atomicAdd(fb_count, 1)] = qi;
      } else if (lane == 0) {
        io.rgb[(size_t)qi * 3 + 0] = pmk::div_rn_fast(fr, denom);
        io.rgb[(size_t)qi * 3 + 1] = pmk::div_rn_fast(fg, denom);
        io.rgb[(size_t)qi * 3 + 2] = pmk::div_rn_fast(fb, denom);
      }
    }
    __syncwarp();
  }
}

/* ---- fast path for uniform-like clouds --------------------------------------------------
 * k_grid_query spends a third of its instructions on a density pass (50 offset-table reads per query) and on a
 * selection that keeps every candidate's bin and position. On a uniform-like cloud neither is needed: the radius
 * that holds about 1.25 k + 10 photons is a property of the MAP (sampled when the grid is built), and the cells are
 * sized so that a 5x5x5 block guarantees it. k_knn_fast therefore starts every query at that radius:
 *   - the lanes own the rows of the block (row = r0 + lane), trim them to the ball in cell units, and a warp scan
 *     flattens them (same bitmap trick as k_grid_query) — no descriptor, no second kernel;
 *   - candidates inside the radius are compacted into shared memory as (dist² bits, id);
 *   - ONE histogram over dist² (window from the count actually found) locates the bin of the k-th candidate, its
 *     <= 32 members are ranked, and the k-th dist² becomes a threshold tau: every candidate with dist² <= tau is
 *     written straight to its output row at a ballot-compacted position. A tie on tau that would need the id order
 *     (the exact (dist², id) contract) is left to k_grid_query;
 *   - a query whose count falls outside [k, 128] retries with the radius rescaled by the count it saw, on a wider
 *     block if the exactness bound needs one (balls cut by the cloud's faces take this path); what still fails
 *     after kFastAttempts goes to the hand-over list that k_grid_query answers.
 * Queries arrive pre-ordered as float4 (x, y, z, bits(index)) from the binning pass of pm_knn.cu. */
constexpr int kFastWords = 24;                /* flat index space of one round of rows: 768 candidates (the rows of a */
constexpr int kFastMaxScan = kFastWords * 32; /* retry on a 7- or 9-cell-wide block are up to 9 cells long)            */
constexpr int kFastPairs = 6;                 /* steps taken two at a time by the unrolled part of the scan (a 5x5x5 block needs ~8) */
constexpr int kFastMaxH = 4;                  /* widest block of the fast path: 9x9 rows (3 rounds) */
constexpr int kFastAttempts = 6;
constexpr int kFastProbes = 10;               /* interpolated counting probes per query ... */
constexpr int kFastBisect = 28;               /* ... then bisection steps, before it is handed over */

struct __align__(16) FastScratch {
  uint2 keys[kBuf];         /* (dist² bits, node id) of the candidates inside the search radius */
  uint2 words[kFastWords];  /* x: bitmap of row starts among elements [32 i, 32 i + 32); y: row holding element 32 i */
  int off[32];              /* per non-empty row: first record - first flat index */
};

/* first counting threshold per candidate count n: T = R2 * win[n], the dist² below which k of n candidates are
 * expected (count ~ r^3) */
constexpr int kWinSize = kBuf + 1;
struct FastParams {
  const float4* sq;   /* queries in processing order */
  float R2;           /* first search radius², already capped by the block guarantee and the cutoff */
  float Rc2;          /* (its radius in cells + the margins)² */
  float margin_c;     /* GridView::margin in cells */
  float win[kWinSize];
};

__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

template <bool GATHER>
__global__ void __launch_bounds__(kWarps * 32, GATHER ? 6 : 8) /* no spills: a reloaded value is no longer warp-uniform to the compiler */
k_knn_fast(const GridView g, const float4* __restrict__ color4, const pmk::QueryIO io, const __grid_constant__ FastParams fp,
           int* __restrict__ ho_list, int* __restrict__ ho_count) {
  __shared__ FastScratch scratch[kWarps];
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u, le_not0 = (lt | (1u << lane)) & ~1u;
  /* the warp index through a shuffle: the compiler then knows it (and every branch on the query index) to be
   * warp-uniform and drops the divergence guards around the warp-level primitives below */
  const int wid = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  FastScratch& ws = scratch[wid];
  const int k = io.k;
  const float fk = (float)k;
  /* rows of the 5x5 block of the first attempt: lane -> (oy, oz) */
  const int oz5 = div_small(lane, 5) - 2, oy5 = lane - (oz5 + 2) * 5 - 2;
  /* GATHER: the next query (position and factor, both in processing order) is loaded while this one is answered — the
   * two dependent loads at the head of the loop were where a quarter of the kernel's stall samples sat */
  float4 qnext = make_float4(0.f, 0.f, 0.f, 0.f);
  float bnext = 0.f;
  if (GATHER) {
    const int t0 = blockIdx.x * kWarps + wid;
    if (t0 < io.nq) { qnext = __ldg(fp.sq + t0); bnext = __ldg(io.sbrdf + t0); }
  }
  for (int t = blockIdx.x * kWarps + wid; t < io.nq; t += gridDim.x * kWarps) {
    float4 qv;
    float brdf = 0.f;
    if (GATHER) {
      qv = qnext; brdf = bnext;
      const int tn = t + gridDim.x * kWarps;
      if (tn < io.nq) { qnext = __ldg(fp.sq + tn); bnext = __ldg(io.sbrdf + tn); }
    } else {
      qv = __ldg(fp.sq + t);
    }
    const int qi = __float_as_int(qv.w);
    if (GATHER) {
      brdf = __shfl_sync(0xffffffffu, brdf, 0); /* warp-uniform to the compiler */
      if (!(brdf > 0.f)) {
        if (lane < 3) io.rgb[(size_t)qi * 3 + lane] = 0.f;
        continue;
      }
    }
    const float qx = qv.x, qy = qv.y, qz = qv.z;
    const float fx = (qx - g.lo[0]) * g.inv_cell, fy = (qy - g.lo[1]) * g.inv_cell, fz = (qz - g.lo[2]) * g.inv_cell;
    const int cx = min(max(__float2int_rd(fx), 0), g.nx - 1), cy = min(max(__float2int_rd(fy), 0), g.ny - 1),
              cz = min(max(__float2int_rd(fz), 0), g.nz - 1);
    const float uy = fy - (float)cy, uz = fz - (float)cz; /* position inside the own cell, in cells */

    /* ---- candidates: rows of the (2h+1)^3 block trimmed to the ball, flattened, scanned. Everything is in cell
     * units; the margins cover the rounding of this and of the build's cell assignment. ---- */
    int h = 2, n = 0, attempt = 0, r0 = 0, nrows = 25;
    int oy = oy5, oz = oz5;
    float R2 = fp.R2, Rc2 = fp.Rc2;
    bool give_up = false;
    while (true) {
      {
        const int ry = cy + oy, rz = cz + oz;
        const float dyc = fmaxf(fmaxf((float)oy - uy, uy - (float)(oy + 1)), 0.f);
        const float dzc = fmaxf(fmaxf((float)oz - uz, uz - (float)(oz + 1)), 0.f);
        const float t2 = Rc2 - dyc * dyc - dzc * dzc;
        int st = 0, len = 0;
        if (r0 + lane < nrows && (unsigned)ry < (unsigned)g.ny && (unsigned)rz < (unsigned)g.nz && t2 > 0.f) {
          const float w = sqrt_approx(t2) * 1.0001f; /* half-width of the ball in this row, rounded up */
          const int xa = max(__float2int_rd(fx - w), max(cx - h, 0)), xb = min(__float2int_rd(fx + w), min(cx + h, g.nx - 1));
          if (xa <= xb) {
            /* the x-slowest copy of the offset table: the five rows of a z-slab that share xa sit in one 20-byte run,
             * so the 25 lanes touch about a dozen 128-byte lines per load instead of 22 (< 2^30 cells: 32-bit index) */
            const int zy = rz * g.ny + ry, plane = g.nz * g.ny;
            st = (int)__ldg(g.cell_start_t + (uint32_t)(xa * plane + zy));
            len = (int)__ldg(g.cell_start_t + (uint32_t)((xb + 1) * plane + zy)) - st;
          }
        }
        int incl = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int v = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += v;
        }
        const int mc = __shfl_sync(0xffffffffu, incl, 31);
        if (mc > kFastMaxScan) { give_up = true; break; }
        if (mc > 0) {
          const int P = incl - len;
          __syncwarp();
          if (lane < kFastWords) ws.words[lane] = make_uint2(0u, 0u);
          __syncwarp();
          const unsigned nonempty = __ballot_sync(0xffffffffu, len > 0);
          if (len > 0) {
            const int d = __popc(nonempty & lt);
            atomicOr(&ws.words[P >> 5].x, 1u << (P & 31));
            ws.off[d] = st - P;
#pragma unroll 1
            for (int i = (P + 31) >> 5; i * 32 < P + len; i++) ws.words[i].y = (unsigned)d;
          }
          __syncwarp();
          /* the per-row offsets move from shared memory into registers (lane d holds row d's): a step fetches its
           * offset with one shuffle */
          const int offreg = ws.off[lane];
          /* full steps two at a time: both addresses are known before the first record arrives, so two loads per lane
           * are in flight (one step's chain LDS -> SHFL -> LDG -> compare is ~400 cycles; with one load in flight the
           * warps of an SM cannot cover it). Unrolled over the 12 words of the flat space: constant smem addresses. */
          int i = 0;
#pragma unroll
          for (int pr = 0; pr < kFastPairs; pr++) {
            if ((2 * pr + 2) * 32 > mc) break;
            float4 p[2];
#pragma unroll
            for (int u = 0; u < 2; u++) {
              const uint2 w = ws.words[2 * pr + u];
              const int run = (int)w.y + __popc(w.x & le_not0);
              p[u] = __ldg(g.pts + (__shfl_sync(0xffffffffu, offreg, run) + (2 * pr + u) * 32 + lane));
            }
#pragma unroll
            for (int u = 0; u < 2; u++) {
              const float d2 = pmk::dist2(p[u].x, p[u].y, p[u].z, qx, qy, qz);
              const bool pass = d2 < R2;
              const unsigned b = __ballot_sync(0xffffffffu, pass);
              /* beyond the buffer everything lands in the last slot; n > kBuf is detected below */
              if (pass) ws.keys[min(n + __popc(b & lt), kBuf - 1)] = make_uint2(__float_as_uint(d2), __float_as_uint(p[u].w));
              n += __popc(b);
            }
            i = 2 * pr + 2;
          }
          for (; i * 32 < mc; i++) {
            const int j = i * 32 + lane;
            const uint2 w = ws.words[i];
            const int run = (int)w.y + __popc(w.x & le_not0);
            /* lanes past the end read a record behind the last row (the array is padded) and drop it */
            const float4 p = __ldg(g.pts + (__shfl_sync(0xffffffffu, offreg, run) + j));
            const float d2 = pmk::dist2(p.x, p.y, p.z, qx, qy, qz);
            const bool pass = j < mc && d2 < R2;
            const unsigned b = __ballot_sync(0xffffffffu, pass);
            if (pass) ws.keys[min(n + __popc(b & lt), kBuf - 1)] = make_uint2(__float_as_uint(d2), __float_as_uint(p.w));
            n += __popc(b);
          }
        }
      }
      r0 += 32;
      if (r0 >= nrows) {
        if (n >= k && n <= kBuf) break;
        /* rescale the radius by the count seen (count ~ r^3), widen the block if the exactness bound asks for it */
        if (++attempt >= kFastAttempts || (n < k && !(R2 < io.cutoff2))) { give_up = true; break; }
        R2 = fminf(R2 * pow23(__fdividef(1.3f * fk + 12.f, (float)max(n, k >> 2))), io.cutoff2);
        const float Rc = sqrtf(R2) * g.inv_cell + 2.f * fp.margin_c;
        h = max(__float2int_ru(Rc + 2.f * fp.margin_c), 2);
        if (h > kFastMaxH) { give_up = true; break; }
        Rc2 = Rc * Rc;
        nrows = (2 * h + 1) * (2 * h + 1);
        r0 = 0; n = 0;
      }
      {
        const int row = r0 + lane, side = 2 * h + 1;
        const int rzo = div_small(row, side);
        oy = row - rzo * side - h; oz = rzo - h;
      }
    }
    if (give_up) {
      if (lane == 0) ho_list[atomicAdd(ho_count, 1)] = qi;
      __syncwarp();
      continue;
    }
    __syncwarp();

    /* ---- the k-th smallest dist², by counting: every lane holds its 3 - 4 candidates in registers, a probe counts
     * those with dist² <= T (per-lane compares + one warp reduction) and the bracket (Tlo, Thi] is narrowed by linear
     * interpolation on the counts until exactly k lie below T. No histogram: shared-memory atomics cost about one
     * L1 wavefront per lane and were half of this kernel's shared-memory traffic. The first T is the quantile the
     * count ~ r^3 law predicts for the n candidates found. ---- */
    uint2 kv[kSlots];
    const bool four = n > 3 * 32; /* the fourth register slot is needed by about one query in five */
    kv[kSlots - 1] = make_uint2(0xffffffffu, 0u);
#pragma unroll
    for (int s = 0; s < kSlots; s++) {
      if (s == kSlots - 1 && !four) break;
      kv[s] = ws.keys[lane + 32 * s]; /* branch-free: slots past n read stale entries and become +inf */
      if (lane + 32 * s >= n) kv[s].x = 0xffffffffu;
    }
    uint32_t tau = 0u;
    bool exact = false;
    {
      float Tlo = 0.f, Thi = R2, T = n == k ? R2 : R2 * fp.win[n]; /* every candidate has dist² < R2 */
      int clo = 0, chi = n;
#pragma unroll 1
      for (int it = 0; it < kFastProbes; it++) {
        const uint32_t tb = __float_as_uint(T);
        int c = (kv[0].x <= tb) + (kv[1].x <= tb) + (kv[2].x <= tb) + (kv[3].x <= tb); /* an unused slot is +inf */
        c = __reduce_add_sync(0xffffffffu, c);
        if (c == k) { exact = true; break; }
        if (c < k) { Tlo = T; clo = c; } else { Thi = T; chi = c; }
        T = Tlo + (Thi - Tlo) * ((float)(k - clo) * rcp_approx((float)(chi - clo))); /* clo < k < chi */
        if (!(T > Tlo && T < Thi)) break; /* degenerate step */
      }
      /* the interpolation stalls now and then (regula falsi: one end of the bracket stops moving) or steps onto an end
       * of the bracket; what is left of the bracket is then bisected. Handing such a query to k_grid_query instead
       * cost a kernel of its own: 1 590 of 1 M queries and 29 us on C2. Only a real tie at the k-th place (no float
       * between the bracket's ends) still leaves here without a threshold. */
      if (!exact) {
#pragma unroll 1
        for (int it = 0; it < kFastBisect; it++) {
          T = 0.5f * (Tlo + Thi);
          if (!(T > Tlo && T < Thi)) break;
          const uint32_t tb = __float_as_uint(T);
          int c = (kv[0].x <= tb) + (kv[1].x <= tb) + (kv[2].x <= tb) + (kv[3].x <= tb);
          c = __reduce_add_sync(0xffffffffu, c);
          if (c == k) { exact = true; break; }
          if (c < k) Tlo = T; else Thi = T;
        }
      }
      if (exact) { /* the k-th smallest itself */
        const uint32_t tb = __float_as_uint(T);
        uint32_t m = 0u;
#pragma unroll
        for (int s = 0; s < kSlots; s++) m = max(m, kv[s].x <= tb ? kv[s].x : 0u);
        tau = __reduce_max_sync(0xffffffffu, m);
      }
    }
    if (!exact) { /* equal distances at the k-th place (the id order decides), or an odd distribution */
      if (lane == 0) ho_list[atomicAdd(ho_count, 1)] = qi;
      __syncwarp();
      continue;
    }

    /* ---- output: every candidate with dist² <= tau ---- */
    if (!GATHER) {
      /* ballot-compacted into keys[0 .. k) (every lane holds its candidates in registers by now), then whole rows
       * leave as coalesced 8-byte stores */
      __syncwarp();
      int base = 0;
#pragma unroll
      for (int s = 0; s < kSlots; s++) {
        if (s == kSlots - 1 && !four) break;
        const bool a = kv[s].x <= tau;
        const unsigned b = __ballot_sync(0xffffffffu, a);
        if (a) ws.keys[base + __popc(b & lt)] = kv[s];
        base += __popc(b);
      }
      __syncwarp();
      int2* ids2 = reinterpret_cast<int2*>(io.ids + (size_t)qi * k); /* k is even: rows are 8-byte aligned */
      float2* d22 = io.d2 ? reinterpret_cast<float2*>(io.d2 + (size_t)qi * k) : nullptr;
      if (2 * lane < k) { /* k <= 64: one pair of entries (2 lane, 2 lane + 1) per lane */
        const uint4 e = reinterpret_cast<const uint4*>(ws.keys)[lane];
        ids2[lane] = make_int2((int)e.y, (int)e.w);
        if (d22) d22[lane] = make_float2(__uint_as_float(e.x), __uint_as_float(e.z));
      }
      if (io.r2 && lane == 0) io.r2[qi] = __uint_as_float(tau);
    } else {
      /* gatherPhotons (ray-tracer/cuda/shading.h:93-121), summed across the warp; call-free division / square root
       * (pm_knn.cuh), exotic operands go to the general kernel */
      const float r2 = __uint_as_float(tau);
      const float rad = pmk::sqrt_rn_fast(r2);
      bool ok = pmk::sqrt_fast_ok(r2) && pmk::div_fast_ok(rad, rad); /* see k_grid_query */
      float fr = 0.f, fg = 0.f, fb = 0.f;
      /* the k accepted entries are compacted first (as for the id output above): the sum then runs over two full-width
       * steps instead of four register slots of which 70 % hold an accepted entry (191 -> ~130 instructions per query) */
      __syncwarp();
      int base = 0;
#pragma unroll
      for (int s = 0; s < kSlots; s++) {
        if (s == kSlots - 1 && !four) break;
        const bool a = kv[s].x <= tau;
        const unsigned b = __ballot_sync(0xffffffffu, a);
        if (a) ws.keys[base + __popc(b & lt)] = kv[s];
        base += __popc(b);
      }
      __syncwarp();
#pragma unroll
      for (int u = 0; u < 2; u++) { /* k <= 64 */
        const int p = lane + 32 * u;
        if (p < k) {
          const uint2 e = ws.keys[p];
          const float4 c = __ldg(color4 + e.y);
          const float d2v = __uint_as_float(e.x);
          const float dist = pmk::sqrt_rn_fast(d2v);
          ok = ok && pmk::sqrt_fast_ok(d2v);
          const float w = __fsub_rn(1.f, __fmul_rn(pmk::div_rn_fast(dist, rad), 1.1f));
          const float sc = __fmul_rn(__fmul_rn(brdf, c.w), w);
          fr = __fadd_rn(fr, __fmul_rn(sc, c.x));
          fg = __fadd_rn(fg, __fmul_rn(sc, c.y));
          fb = __fadd_rn(fb, __fmul_rn(sc, c.z));
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        fr = __fadd_rn(fr, __shfl_xor_sync(0xffffffffu, fr, o));
        fg = __fadd_rn(fg, __shfl_xor_sync(0xffffffffu, fg, o));
        fb = __fadd_rn(fb, __shfl_xor_sync(0xffffffffu, fb, o));
      }
      const float k1 = (1 - (2.f / 3.f) * (1.f / 1.1f));
      const float denom = __fmul_rn(__fmul_rn(__fmul_rn(k1, 2.f), 3.14159274101257324f), r2);
      ok = ok && pmk::div_fast_ok(fr, denom) && pmk::div_fast_ok(fg, denom) && pmk::div_fast_ok(fb, denom);
      if (!__all_sync(0xffffffffu, ok)) {
        if (lane == 0) ho_list[atomicAdd(ho_count, 1)] = qi;
      } else if (lane == 0) {
        io.rgb[(size_t)qi * 3 + 0] = pmk::div_rn_fast(fr, denom);
        io.rgb[(size_t)qi * 3 + 1] = pmk::div_rn_fast(fg, denom);
        io.rgb[(size_t)qi * 3 + 2] = pmk::div_rn_fast(fb, denom);
      }
    }
    __syncwarp();
  }
}

/* flux += power * colour of one accepted photon; PREMUL: the product was formed when the cells-only map was built */
template <bool PREMUL>
__device__ __forceinline__ void add_flux(const float4 c, float& sr, float& sg, float& sb) {
  sr = __fadd_rn(sr, PREMUL ? c.x : __fmul_rn(c.w, c.x));
  sg = __fadd_rn(sg, PREMUL ? c.y : __fmul_rn(c.w, c.y));
  sb = __fadd_rn(sb, PREMUL ? c.z : __fmul_rn(c.w, c.z));
}

constexpr int kMaxHRadius = 48;
constexpr int kThreadMaxH = 6;
constexpr int kRadiusSamples = 2048;
constexpr int kRadiusDenseMean = 200; /* mean candidates per sampled hit point from which the group kernel is faster */
struct RadiusPick { int candidates, lit, heavy; };
__device__ __forceinline__ bool radius_dense(const RadiusPick* pick) {
  return pick->candidates > kRadiusDenseMean * max(pick->lit, 1);
}

/* ---- fixed-radius gather (progressive photon mapping pass; SURVEY appendix B — no reference code) ------
 * Per hit point: flux = brdf * sum(power * colour) over the nodes with dist² < radius², count = their number.
 * Same candidate enumeration as k_grid_query (ball-trimmed rows of the smallest block that covers the ball),
 * no selection. Rows are taken 32 at a time (one per lane); a chunk of rows with at most kMaxScan candidates is
 * flattened into one index space so that short rows still fill the warp, a denser chunk is streamed row by row
 * (its rows are long: every step is a coalesced 512-byte read). Early passes of a progressive render hold
 * thousands of photons per ball on blocks of 15..20 cells a side: both used to fall to the per-thread kd-tree
 * range kernel below, which now only sees balls that no block of (2 kMaxHRadius + 1)^3 cells covers.
 * Since the per-group and per-thread kernels below exist, this one takes what they leave behind (heavy_list:
 * groups that are not compact, blocks too wide for a thread), or the whole list when called without a pick. */
template <bool PREMUL>
__global__ void __launch_bounds__(kWarps * 32)
k_grid_radius(const GridView g, const float4* __restrict__ color4, const float* __restrict__ q, const float* __restrict__ brdf,
              const float* __restrict__ radius2, const int* __restrict__ order, int nq_host, const int* __restrict__ heavy_list,
              const RadiusPick* __restrict__ pick, int max_h, float* __restrict__ out_flux,
              int32_t* __restrict__ out_count, int* __restrict__ fb_list, int* __restrict__ fb_count) {
  __shared__ WarpScratch scratch[kWarps];
  /* after the per-thread / per-group kernels: only what they left behind */
  int nq = nq_host;
  if (pick) { order = heavy_list; nq = pick->heavy; }
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u, le_not0 = (lt | (1u << lane)) & ~1u;
  const int wid = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); /* warp-uniform to the compiler */
  WarpScratch& ws = scratch[wid];
  for (int t = blockIdx.x * kWarps + wid; t < nq; t += gridDim.x * kWarps) {
    const int qi = __shfl_sync(0xffffffffu, order ? order[t] : t, 0); /* warp-uniform to the compiler */
    const float f = __shfl_sync(0xffffffffu, brdf[qi], 0), R2 = __shfl_sync(0xffffffffu, radius2[qi], 0);
    if (!(f > 0.f) || !(R2 > 0.f)) { /* no hit point here */
      if (lane < 3) out_flux[(size_t)qi * 3 + lane] = 0.f;
      if (lane == 0) out_count[qi] = 0;
      continue;
    }
    const float qx = q[(size_t)qi * 3 + 0], qy = q[(size_t)qi * 3 + 1], qz = q[(size_t)qi * 3 + 2];
    const float fx = (qx - g.lo[0]) * g.inv_cell, fy = (qy - g.lo[1]) * g.inv_cell, fz = (qz - g.lo[2]) * g.inv_cell;
    const int cx = min(max(__float2int_rd(fx), 0), g.nx - 1), cy = min(max(__float2int_rd(fy), 0), g.ny - 1),
              cz = min(max(__float2int_rd(fz), 0), g.nz - 1);
    /* smallest block whose guaranteed radius covers the ball */
    int h = min(max(__float2int_ru(sqrtf(R2) * g.inv_cell), 1), max_h);
    int x0, x1, y0, y1, z0, z1;
    bool covered = false;
    for (; h <= max_h; h++) {
      x0 = max(cx - h, 0); x1 = min(cx + h, g.nx - 1);
      y0 = max(cy - h, 0); y1 = min(cy + h, g.ny - 1);
      z0 = max(cz - h, 0); z1 = min(cz + h, g.nz - 1);
      float Gc = INFINITY;
      if (x0 > 0) Gc = fminf(Gc, fx - (float)x0);
      if (x1 < g.nx - 1) Gc = fminf(Gc, (float)(x1 + 1) - fx);
      if (y0 > 0) Gc = fminf(Gc, fy - (float)y0);
      if (y1 < g.ny - 1) Gc = fminf(Gc, (float)(y1 + 1) - fy);
      if (z0 > 0) Gc = fminf(Gc, fz - (float)z0);
      if (z1 < g.nz - 1) Gc = fminf(Gc, (float)(z1 + 1) - fz);
      const float G = Gc * g.cell - g.margin;
      if (G > 0.f && !(G * G < R2)) { covered = true; break; }
    }
    bool give_up = !covered;
    float sr = 0.f, sg = 0.f, sb = 0.f;
    int cnt = 0;
    if (!give_up) {
      const int ny_r = y1 - y0 + 1, nrows = ny_r * (z1 - z0 + 1);
      for (int r0 = 0; r0 < nrows; r0 += 32) {
        const int row = r0 + lane;
        int len = 0, st = 0;
        if (row < nrows) {
          const int rzo = row / ny_r;
          const int ry = y0 + row - rzo * ny_r, rz = z0 + rzo;
          const float ylo = g.lo[1] + (float)ry * g.cell, zlo = g.lo[2] + (float)rz * g.cell;
          const float dy = fmaxf(fmaxf(ylo - qy, qy - (ylo + g.cell)) - g.margin, 0.f);
          const float dz = fmaxf(fmaxf(zlo - qz, qz - (zlo + g.cell)) - g.margin, 0.f);
          const float dyz2 = dy * dy + dz * dz;
          if (dyz2 < R2) {
            const float t2 = R2 - dyz2;
            const float w = (t2 < 1e30f ? t2 * rsqrtf(t2) * 1.0001f : INFINITY) + g.margin;
            const int xa = max(cell_idx(qx - w, g.lo[0], g.inv_cell, g.nx), x0);
            const int xb = min(cell_idx(qx + w, g.lo[0], g.inv_cell, g.nx), x1);
            if (xa <= xb) {
              const size_t base = ((size_t)rz * g.ny + ry) * g.nx;
              st = (int)__ldg(g.cell_start + base + xa);
              len = (int)__ldg(g.cell_start + base + xb + 1) - st;
            }
          }
        }
        int incl = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int v = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += v;
        }
        const int mc = __shfl_sync(0xffffffffu, incl, 31);
        if (mc == 0) continue;
        if (mc > kMaxScan) { /* dense chunk: row by row */
          unsigned rows_left = __ballot_sync(0xffffffffu, len > 0);
          while (rows_left) {
            const int r = __ffs(rows_left) - 1;
            rows_left &= rows_left - 1;
            const int rs = __shfl_sync(0xffffffffu, st, r), rl = __shfl_sync(0xffffffffu, len, r);
            for (int j = lane; j < rl; j += 32) {
              const float4 p = __ldg(g.pts + (rs + j));
              const float d2 = pmk::dist2(p.x, p.y, p.z, qx, qy, qz);
              if (d2 < R2) {
                const float4 c = __ldg(color4 + __float_as_int(p.w));
                add_flux<PREMUL>(c, sr, sg, sb);
                cnt++;
              }
            }
          }
          continue;
        }
        const int P = incl - len;
        __syncwarp();
        ws.words[lane] = make_uint2(0u, 0u);
        __syncwarp();
        const unsigned nonempty = __ballot_sync(0xffffffffu, len > 0);
        if (len > 0) {
          const int d = __popc(nonempty & lt);
          atomicOr(&ws.words[P >> 5].x, 1u << (P & 31));
          ws.run_off[d] = st - P;
          for (int i = (P + 31) >> 5; i * 32 < P + len; i++) ws.words[i].y = (unsigned)d;
        }
        __syncwarp();
        for (int i = 0; i * 32 < mc; i++) {
          const int j = i * 32 + lane;
          const int jj = min(j, mc - 1);
          const uint2 w = ws.words[i];
          const int run = (int)w.y + __popc(w.x & le_not0);
          const float4 p = __ldg(g.pts + (ws.run_off[run] + jj));
          const float d2 = pmk::dist2(p.x, p.y, p.z, qx, qy, qz);
          if (j < mc && d2 < R2) {
            const float4 c = __ldg(color4 + __float_as_int(p.w));
            add_flux<PREMUL>(c, sr, sg, sb);
            cnt++;
          }
        }
      }
    }
    if (give_up) {
      if (lane == 0) fb_list[atomicAdd(fb_count, 1)] = qi;
      __syncwarp();
      continue;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sr = __fadd_rn(sr, __shfl_xor_sync(0xffffffffu, sr, o));
      sg = __fadd_rn(sg, __shfl_xor_sync(0xffffffffu, sg, o));
      sb = __fadd_rn(sb, __shfl_xor_sync(0xffffffffu, sb, o));
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0) {
      out_flux[(size_t)qi * 3 + 0] = __fmul_rn(f, sr);
      out_flux[(size_t)qi * 3 + 1] = __fmul_rn(f, sg);
      out_flux[(size_t)qi * 3 + 2] = __fmul_rn(f, sb);
      out_count[qi] = cnt;
    }
    __syncwarp();
  }
}

/* ---- the same gather, one warp per GROUP of kGroup neighbouring hit points ------------------------------
 * Dense balls (thousands of photons each): neighbouring hit points of the sorted list — neighbouring pixels — see
 * almost the same photons, yet one warp per hit point pulls every candidate through L1 once per hit point (ncu on
 * the C5 frame, 2800 photons per ball: 12.5 TB/s between L2 and L1, long-scoreboard 16.8 stalls per issue, and 37
 * instructions per 32 (hit point, candidate) pairs, most of them row bookkeeping). Here a warp walks the trimmed
 * rows of the UNION of kGroup blocks once and tests every candidate against the kGroup balls held in registers:
 * one load and one round of bookkeeping per kGroup tests. Same candidates and more (a superset of every member's
 * rows), same distance arithmetic and comparison: counts are unchanged. A group that is not compact (its union
 * block holds more than twice the cells of its largest member's: a scanline or bin boundary in the order), or
 * that has a member no block covers, is appended to heavy_list for k_grid_radius (1 .. 4 % of the C5 hit points).
 * C5 frame, one GPU, 29 M photons, 2800 per ball: 13.0 ms against 27.8 ms for one warp per hit point. */
constexpr int kGroup = 4;
struct __align__(16) GroupScratch {
  uint2 words[kWords];
  int run_off[32];
};

template <bool PREMUL>
__global__ void __launch_bounds__(kWarps * 32, 4)
k_grid_radius_group(const GridView g, const float4* __restrict__ color4, const float* __restrict__ q, const float* __restrict__ brdf,
                    const float* __restrict__ radius2, const int* __restrict__ order, int nq, int max_h, float* __restrict__ out_flux,
                    int32_t* __restrict__ out_count, int* __restrict__ heavy_list, RadiusPick* __restrict__ pick) {
  __shared__ GroupScratch scratch[kWarps];
  if (!radius_dense(pick)) return; /* the per-thread kernel has taken the list */
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u, le_not0 = (lt | (1u << lane)) & ~1u;
  const int wid = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); /* warp-uniform to the compiler */
  GroupScratch& ws = scratch[wid];
  const int ngroups = (nq + kGroup - 1) / kGroup;
  for (int t = blockIdx.x * kWarps + wid; t < ngroups; t += gridDim.x * kWarps) {
    /* lanes 0 .. kGroup-1 own one member each: block and cover test as in k_grid_radius */
    const int my = t * kGroup + lane;
    const bool have = lane < kGroup && my < nq;
    const int qi_l = have ? (order ? order[my] : my) : -1;
    const float f_l = have ? brdf[qi_l] : 0.f, R2_l = have ? radius2[qi_l] : 0.f;
    const bool lit_l = f_l > 0.f && R2_l > 0.f;
    float qx_l = 0.f, qy_l = 0.f, qz_l = 0.f;
    int x0 = INT_MAX, x1 = INT_MIN, y0 = INT_MAX, y1 = INT_MIN, z0 = INT_MAX, z1 = INT_MIN, vol_l = 0;
    bool covered = true;
    if (lit_l) {
      qx_l = q[(size_t)qi_l * 3 + 0]; qy_l = q[(size_t)qi_l * 3 + 1]; qz_l = q[(size_t)qi_l * 3 + 2];
      const float fx = (qx_l - g.lo[0]) * g.inv_cell, fy = (qy_l - g.lo[1]) * g.inv_cell, fz = (qz_l - g.lo[2]) * g.inv_cell;
      const int cx = min(max(__float2int_rd(fx), 0), g.nx - 1), cy = min(max(__float2int_rd(fy), 0), g.ny - 1),
                cz = min(max(__float2int_rd(fz), 0), g.nz - 1);
      covered = false;
      for (int h = min(max(__float2int_ru(sqrtf(R2_l) * g.inv_cell), 1), max_h); h <= max_h; h++) {
        x0 = max(cx - h, 0); x1 = min(cx + h, g.nx - 1);
        y0 = max(cy - h, 0); y1 = min(cy + h, g.ny - 1);
        z0 = max(cz - h, 0); z1 = min(cz + h, g.nz - 1);
        float Gc = INFINITY;
        if (x0 > 0) Gc = fminf(Gc, fx - (float)x0);
        if (x1 < g.nx - 1) Gc = fminf(Gc, (float)(x1 + 1) - fx);
        if (y0 > 0) Gc = fminf(Gc, fy - (float)y0);
        if (y1 < g.ny - 1) Gc = fminf(Gc, (float)(y1 + 1) - fy);
        if (z0 > 0) Gc = fminf(Gc, fz - (float)z0);
        if (z1 < g.nz - 1) Gc = fminf(Gc, (float)(z1 + 1) - fz);
        const float G = Gc * g.cell - g.margin;
        if (G > 0.f && !(G * G < R2_l)) { covered = true; break; }
      }
      /* cells of the member's block, saturated (a cells-only map's block may be the whole grid) */
      vol_l = (int)fminf((float)(x1 - x0 + 1) * (float)(y1 - y0 + 1) * (float)(z1 - z0 + 1), 1e9f);
      if (!covered) { x0 = y0 = z0 = INT_MAX; x1 = y1 = z1 = INT_MIN; }
    }
    const unsigned litmask = __ballot_sync(0xffffffffu, lit_l);
    if (litmask == 0u) { /* nothing but misses */
      if (have) {
        out_flux[(size_t)qi_l * 3 + 0] = 0.f; out_flux[(size_t)qi_l * 3 + 1] = 0.f; out_flux[(size_t)qi_l * 3 + 2] = 0.f;
        out_count[qi_l] = 0;
      }
      continue;
    }
    const int X0 = __reduce_min_sync(0xffffffffu, x0), X1 = __reduce_max_sync(0xffffffffu, x1);
    const int Y0 = __reduce_min_sync(0xffffffffu, y0), Y1 = __reduce_max_sync(0xffffffffu, y1);
    const int Z0 = __reduce_min_sync(0xffffffffu, z0), Z1 = __reduce_max_sync(0xffffffffu, z1);
    const int vmax = __reduce_max_sync(0xffffffffu, vol_l);
    const float volU = (float)(X1 - X0 + 1) * (float)(Y1 - Y0 + 1) * (float)(Z1 - Z0 + 1);
    if (__any_sync(0xffffffffu, lit_l && !covered) || volU > 2.f * (float)vmax) {
      if (have) heavy_list[atomicAdd(&pick->heavy, 1)] = qi_l;
      continue;
    }
    float qx[kGroup], qy[kGroup], qz[kGroup], R2[kGroup], sr[kGroup], sg[kGroup], sb[kGroup];
    int cnt[kGroup];
#pragma unroll
    for (int m = 0; m < kGroup; m++) {
      qx[m] = __shfl_sync(0xffffffffu, qx_l, m); qy[m] = __shfl_sync(0xffffffffu, qy_l, m); qz[m] = __shfl_sync(0xffffffffu, qz_l, m);
      const float r2m = __shfl_sync(0xffffffffu, R2_l, m);
      R2[m] = (litmask >> m) & 1u ? r2m : -1.f; /* a miss accepts nothing */
      sr[m] = sg[m] = sb[m] = 0.f; cnt[m] = 0;
    }
    /* candidate at position `at` of the cell-ordered arrays. In a cells-only map (PREMUL) the colour sits at the
     * same position, so both loads are issued together; in a full map the colour is looked up by node id. */
    auto test = [&](const int at, const bool valid) {
      const float4 p = __ldg(g.pts + at);
      float4 c;
      if (PREMUL) c = __ldg(color4 + at);
      bool ps[kGroup], any = false;
#pragma unroll
      for (int m = 0; m < kGroup; m++) {
        ps[m] = valid && pmk::dist2(p.x, p.y, p.z, qx[m], qy[m], qz[m]) < R2[m];
        any = any || ps[m];
      }
      if (PREMUL || any) {
        if (!PREMUL) c = __ldg(color4 + __float_as_int(p.w));
#pragma unroll
        for (int m = 0; m < kGroup; m++)
          if (ps[m]) { add_flux<PREMUL>(c, sr[m], sg[m], sb[m]); cnt[m]++; }
      }
    };
    const int ny_r = Y1 - Y0 + 1, nrows = ny_r * (Z1 - Z0 + 1);
    for (int r0 = 0; r0 < nrows; r0 += 32) {
      const int row = r0 + lane;
      int len = 0, st = 0;
      if (row < nrows) {
        const int rzo = row / ny_r;
        const int ry = Y0 + row - rzo * ny_r, rz = Z0 + rzo;
        const float ylo = g.lo[1] + (float)ry * g.cell, zlo = g.lo[2] + (float)rz * g.cell;
        int xa = INT_MAX, xb = INT_MIN;
#pragma unroll
        for (int m = 0; m < kGroup; m++) { /* the row's x range: union of the members' trimmed ranges */
          const float dy = fmaxf(fmaxf(ylo - qy[m], qy[m] - (ylo + g.cell)) - g.margin, 0.f);
          const float dz = fmaxf(fmaxf(zlo - qz[m], qz[m] - (zlo + g.cell)) - g.margin, 0.f);
          const float dyz2 = dy * dy + dz * dz;
          if (dyz2 < R2[m]) {
            const float t2 = R2[m] - dyz2;
            const float w = (t2 < 1e30f ? t2 * rsqrtf(t2) * 1.0001f : INFINITY) + g.margin;
            xa = min(xa, max(cell_idx(qx[m] - w, g.lo[0], g.inv_cell, g.nx), X0));
            xb = max(xb, min(cell_idx(qx[m] + w, g.lo[0], g.inv_cell, g.nx), X1));
          }
        }
        if (xa <= xb) {
          const size_t base = ((size_t)rz * g.ny + ry) * g.nx;
          st = (int)__ldg(g.cell_start + base + xa);
          len = (int)__ldg(g.cell_start + base + xb + 1) - st;
        }
      }
      int incl = len;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      const int mc = __shfl_sync(0xffffffffu, incl, 31);
      if (mc == 0) continue;
      if (mc > kMaxScan) { /* dense chunk: row by row (long rows: every step is a coalesced 512-byte read) */
        unsigned rows_left = __ballot_sync(0xffffffffu, len > 0);
        while (rows_left) {
          const int r = __ffs(rows_left) - 1;
          rows_left &= rows_left - 1;
          const int rs = __shfl_sync(0xffffffffu, st, r), rl = __shfl_sync(0xffffffffu, len, r);
#pragma unroll 4
          for (int j = lane; j < rl; j += 32) test(rs + j, true);
        }
        continue;
      }
      const int P = incl - len;
      __syncwarp();
      ws.words[lane] = make_uint2(0u, 0u);
      __syncwarp();
      const unsigned nonempty = __ballot_sync(0xffffffffu, len > 0);
      if (len > 0) {
        const int d = __popc(nonempty & lt);
        atomicOr(&ws.words[P >> 5].x, 1u << (P & 31));
        ws.run_off[d] = st - P;
        for (int i = (P + 31) >> 5; i * 32 < P + len; i++) ws.words[i].y = (unsigned)d;
      }
      __syncwarp();
#pragma unroll 2
      for (int i = 0; i * 32 < mc; i++) {
        const int j = i * 32 + lane;
        const uint2 w = ws.words[i];
        const int run = (int)w.y + __popc(w.x & le_not0);
        test(ws.run_off[run] + min(j, mc - 1), j < mc);
      }
    }
#pragma unroll
    for (int m = 0; m < kGroup; m++) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        sr[m] = __fadd_rn(sr[m], __shfl_xor_sync(0xffffffffu, sr[m], o));
        sg[m] = __fadd_rn(sg[m], __shfl_xor_sync(0xffffffffu, sg[m], o));
        sb[m] = __fadd_rn(sb[m], __shfl_xor_sync(0xffffffffu, sb[m], o));
      }
      cnt[m] = __reduce_add_sync(0xffffffffu, cnt[m]);
    }
    /* lane m writes member m */
    float wr = 0.f, wg = 0.f, wb = 0.f;
    int wc = 0;
#pragma unroll
    for (int m = 0; m < kGroup; m++)
      if (lane == m) { wr = sr[m]; wg = sg[m]; wb = sb[m]; wc = cnt[m]; }
    if (have) {
      out_flux[(size_t)qi_l * 3 + 0] = lit_l ? __fmul_rn(f_l, wr) : 0.f;
      out_flux[(size_t)qi_l * 3 + 1] = lit_l ? __fmul_rn(f_l, wg) : 0.f;
      out_flux[(size_t)qi_l * 3 + 2] = lit_l ? __fmul_rn(f_l, wb) : 0.f;
      out_count[qi_l] = lit_l ? wc : 0;
    }
    __syncwarp();
  }
}

/* ---- the same gather, one THREAD per hit point -----------------------------------------------------
 * With up to a couple of hundred photons per ball (C5 on eight ranks after the first passes: 90 per lit hit point)
 * even a warp per group spends most of its instructions on set-up, scans and reductions. Here every thread walks
 * the ball-trimmed rows of its own hit point (same row bounds as k_grid_radius, so the same candidates);
 * neighbouring hit points of the sorted list share their rows through L1. Measured on the C5 hit points (B200,
 * candidates = photons in the trimmed rows, per lit hit point; group kernel second): 1.45 vs 1.6 ms at 170
 * candidates, then behind: 2.29 vs 2.09 at 265, 2.49 vs 2.13 at 300, 3.90 vs 3.04 at 500, 5.4 vs 3.2 at 540 of a
 * 14 M-photon shard. So a sample of the hit points decides per call (SAMPLE pass: mean number of candidates, 2048
 * warps, kRadiusDenseMean), both kernels are launched and the one that is not wanted returns at once. A hit point
 * whose block is wider than 2 * kThreadMaxH + 1 cells is appended to heavy_list and left to k_grid_radius. */
template <bool SAMPLE, bool PREMUL>
__global__ void __launch_bounds__(256)
k_grid_radius_thread(const GridView g, const float4* __restrict__ color4, const float* __restrict__ q, const float* __restrict__ brdf,
                     const float* __restrict__ radius2, const int* __restrict__ order, int nq, int stride,
                     float* __restrict__ out_flux, int32_t* __restrict__ out_count, int* __restrict__ heavy_list,
                     RadiusPick* __restrict__ pick) {
  /* SAMPLE: one warp per sampled hit point (its rows are dealt to the lanes), every stride-th of the list */
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int row_first = SAMPLE ? (int)(threadIdx.x & 31) : 0, row_step = SAMPLE ? 32 : 1;
  if (SAMPLE) t = (t >> 5) * stride;
  if (t >= nq) return;
  if (!SAMPLE && radius_dense(pick)) return; /* the warp kernel takes the whole list */
  const int qi = order ? order[t] : t;
  const float f = brdf[qi], R2 = radius2[qi];
  float sr = 0.f, sg = 0.f, sb = 0.f;
  int cnt = 0, scanned = 0;
  bool heavy = false;
  const bool lit = f > 0.f && R2 > 0.f;
  if (lit) {
    const float qx = q[(size_t)qi * 3 + 0], qy = q[(size_t)qi * 3 + 1], qz = q[(size_t)qi * 3 + 2];
    const float fx = (qx - g.lo[0]) * g.inv_cell, fy = (qy - g.lo[1]) * g.inv_cell, fz = (qz - g.lo[2]) * g.inv_cell;
    const int cx = min(max(__float2int_rd(fx), 0), g.nx - 1), cy = min(max(__float2int_rd(fy), 0), g.ny - 1),
              cz = min(max(__float2int_rd(fz), 0), g.nz - 1);
    int h = max(__float2int_ru(sqrtf(R2) * g.inv_cell), 1);
    int x0 = 0, x1 = 0, y0 = 0, y1 = 0, z0 = 0, z1 = 0;
    bool covered = false;
    for (; h <= kThreadMaxH; h++) { /* smallest block whose guaranteed radius covers the ball (as in k_grid_radius) */
      x0 = max(cx - h, 0); x1 = min(cx + h, g.nx - 1);
      y0 = max(cy - h, 0); y1 = min(cy + h, g.ny - 1);
      z0 = max(cz - h, 0); z1 = min(cz + h, g.nz - 1);
      float Gc = INFINITY;
      if (x0 > 0) Gc = fminf(Gc, fx - (float)x0);
      if (x1 < g.nx - 1) Gc = fminf(Gc, (float)(x1 + 1) - fx);
      if (y0 > 0) Gc = fminf(Gc, fy - (float)y0);
      if (y1 < g.ny - 1) Gc = fminf(Gc, (float)(y1 + 1) - fy);
      if (z0 > 0) Gc = fminf(Gc, fz - (float)z0);
      if (z1 < g.nz - 1) Gc = fminf(Gc, (float)(z1 + 1) - fz);
      const float G = Gc * g.cell - g.margin;
      if (G > 0.f && !(G * G < R2)) { covered = true; break; }
    }
    heavy = !covered;
    const int ny_r = y1 - y0 + 1, nrows = heavy ? 0 : ny_r * (z1 - z0 + 1);
    for (int row = row_first; row < nrows; row += row_step) {
      const int rzo = row / ny_r;
      const int ry = y0 + row - rzo * ny_r, rz = z0 + rzo;
      const float zlo = g.lo[2] + (float)rz * g.cell, ylo = g.lo[1] + (float)ry * g.cell;
      const float dz = fmaxf(fmaxf(zlo - qz, qz - (zlo + g.cell)) - g.margin, 0.f);
      const float dy = fmaxf(fmaxf(ylo - qy, qy - (ylo + g.cell)) - g.margin, 0.f);
      const float dyz2 = dy * dy + dz * dz;
      if (!(dyz2 < R2)) continue;
      const float t2 = R2 - dyz2;
      const float w = (t2 < 1e30f ? t2 * rsqrtf(t2) * 1.0001f : INFINITY) + g.margin;
      const int xa = max(cell_idx(qx - w, g.lo[0], g.inv_cell, g.nx), x0);
      const int xb = min(cell_idx(qx + w, g.lo[0], g.inv_cell, g.nx), x1);
      if (xa > xb) continue;
      const size_t base = ((size_t)rz * g.ny + ry) * g.nx;
      const int st = (int)__ldg(g.cell_start + base + xa), en = (int)__ldg(g.cell_start + base + xb + 1);
      if (SAMPLE) scanned += en - st;
      else for (int j = st; j < en; j++) {
        const float4 p = __ldg(g.pts + j);
        const float d2 = pmk::dist2(p.x, p.y, p.z, qx, qy, qz);
        if (d2 < R2) {
          add_flux<PREMUL>(__ldg(color4 + __float_as_int(p.w)), sr, sg, sb);
          cnt++;
        }
      }
    }
  }
  if (SAMPLE) { /* warp-uniform: lit and heavy depend on the hit point only */
    scanned = __reduce_add_sync(0xffffffffu, scanned);
    if (lit && (threadIdx.x & 31) == 0) {
      atomicAdd(&pick->candidates, heavy ? 4 * kRadiusDenseMean : min(scanned, 64 * kRadiusDenseMean));
      atomicAdd(&pick->lit, 1);
    }
    return;
  }
  if (heavy) {
    heavy_list[atomicAdd(&pick->heavy, 1)] = qi;
    return;
  }
  const float fpos = f > 0.f ? f : 0.f;
  out_flux[(size_t)qi * 3 + 0] = __fmul_rn(fpos, sr);
  out_flux[(size_t)qi * 3 + 1] = __fmul_rn(fpos, sg);
  out_flux[(size_t)qi * 3 + 2] = __fmul_rn(fpos, sb);
  out_count[qi] = cnt;
}

/* kd-tree range query, one thread per hit point: the fallback of k_grid_radius and the only path of maps
 * without a grid. Same visiting order as cukd's stack traversal with a fixed cull radius. */
__global__ void __launch_bounds__(128)
k_kd_radius(const float4* __restrict__ nodes, const float4* __restrict__ color4, int n, const float* __restrict__ q,
            const float* __restrict__ brdf, const float* __restrict__ radius2, int nq, const int* __restrict__ list,
            const int* __restrict__ list_count, float* __restrict__ out_flux, int32_t* __restrict__ out_count) {
  const int total = list ? *list_count : nq;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const int qi = list ? list[t] : t;
    const float f = brdf[qi], R2 = radius2[qi];
    float sr = 0.f, sg = 0.f, sb = 0.f;
    int cnt = 0;
    if (f > 0.f && R2 > 0.f) {
      const float qx = q[(size_t)qi * 3 + 0], qy = q[(size_t)qi * 3 + 1], qz = q[(size_t)qi * 3 + 2];
      int stack[48];
      int sp = 0, curr = 0;
      while (true) {
        while (curr < n) {
          const float4 nd = __ldg(nodes + curr);
          const int dim = __float_as_int(nd.w);
          const float d2 = pmk::dist2(nd.x, nd.y, nd.z, qx, qy, qz);
          if (d2 < R2) {
            const float4 c = __ldg(color4 + curr);
            sr = __fadd_rn(sr, __fmul_rn(c.w, c.x));
            sg = __fadd_rn(sg, __fmul_rn(c.w, c.y));
            sb = __fadd_rn(sb, __fmul_rn(c.w, c.z));
            cnt++;
          }
          const float dp = dim == 0 ? qx - nd.x : (dim == 1 ? qy - nd.y : qz - nd.z);
          const int l = 2 * curr + 1;
          const int far_child = dp < 0.f ? l + 1 : l;
          if (__fmul_rn(dp, dp) < R2 && far_child < n && sp < 48) stack[sp++] = far_child;
          curr = dp < 0.f ? l : l + 1;
        }
        if (sp == 0) break;
        curr = stack[--sp];
      }
    }
    out_flux[(size_t)qi * 3 + 0] = __fmul_rn(f > 0.f ? f : 0.f, sr);
    out_flux[(size_t)qi * 3 + 1] = __fmul_rn(f > 0.f ? f : 0.f, sg);
    out_flux[(size_t)qi * 3 + 2] = __fmul_rn(f > 0.f ? f : 0.f, sb);
    out_count[qi] = cnt;
  }
}

/* ---- grid construction ------------------------------------------------------------------ */
__global__ void k_cell_ids(const float4* __restrict__ nodes, int n, GridView g, uint32_t* __restrict__ keys,
                           int* __restrict__ vals) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = __ldg(nodes + i);
  const int cx = cell_of(p.x, g.lo[0], g.inv_cell, g.nx), cy = cell_of(p.y, g.lo[1], g.inv_cell, g.ny),
            cz = cell_of(p.z, g.lo[2], g.inv_cell, g.nz);
  keys[i] = (uint32_t)(((size_t)cz * g.ny + cy) * g.nx + cx);
  vals[i] = i;
}

/* sorted (cell, node) pairs -> cell-ordered records and the offset table (empty cells included) */
__global__ void k_grid_fill(const float4* __restrict__ nodes, int n, const uint32_t* __restrict__ keys,
                            const int* __restrict__ vals, uint32_t n_cells, float4* __restrict__ pts,
                            uint32_t* __restrict__ cell_start) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > n) return;
  const int64_t prev = i == 0 ? -1 : (int64_t)keys[i - 1];
  const int64_t cur = i == n ? (int64_t)n_cells : (int64_t)keys[i];
  for (int64_t c = prev + 1; c <= cur; c++) cell_start[c] = (uint32_t)i;
  if (i < n) {
    const int id = vals[i];
    const float4 p = __ldg(nodes + id);
    pts[i] = make_float4(p.x, p.y, p.z, __int_as_float(id));
  }
}

/* cell_start[(z * ny + y) * nx + x] -> cell_start_t[(x * nz + z) * ny + y] for x in [0, nx]; column nx holds the end
 * of every row (= the start of the next row's first cell) */
__global__ void k_cells_transpose(const uint32_t* __restrict__ cs, int nx, int ny, int nz, uint32_t* __restrict__ out) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t rows = (int64_t)ny * nz;
  if (t >= rows * (nx + 1)) return;
  const int64_t x = t / rows, zy = t - x * rows; /* consecutive threads write consecutive words */
  out[t] = __ldg(cs + zy * nx + x);
}

__global__ void k_sample_queries(const float4* __restrict__ nodes, int n, int ns, float* __restrict__ q) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ns) return;
  /* a strided sample of the level-order array touches every level, i.e. it is spread over the cloud */
  const float4 p = __ldg(nodes + (size_t)i * (n / ns));
  q[3 * i + 0] = p.x; q[3 * i + 1] = p.y; q[3 * i + 2] = p.z;
}

GridView view_of(const pm_map* m) {
  GridView g;
  g.pts = m->grid_pts; g.cell_start = m->cell_start; g.cell_start_t = m->cell_start_t;
  for (int d = 0; d < 3; d++) g.lo[d] = m->bounds[d];
  g.cell = m->cell; g.inv_cell = m->inv_cell; g.margin = m->cell_margin;
  g.nx = m->gdim[0]; g.ny = m->gdim[1]; g.nz = m->gdim[2];
  return g;
}

int host_cell_of(float v, float lo, float inv_cell) {
  const float u = (v - lo) * inv_cell; /* same float arithmetic as cell_of() */
  return (int)std::floor(std::min(std::max(u, 0.f), 1e9f)); /* NaN / inf never reach the cast */
}

}  // namespace

namespace pm {

/* cells-only maps: colours follow the cell order of grid_pts and the id a record carries becomes its own position,
 * so the colour of a candidate sits next to its position (input order is the order photons were traced in: every
 * colour fetch would be a DRAM sector of its own). The colours are stored multiplied by the power (PREMUL kernels). */
__global__ void k_cells_reorder(int n, float4* __restrict__ pts, const float4* __restrict__ color_in, float4* __restrict__ color_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 p = pts[i];
  const float4 c = __ldg(color_in + __float_as_int(p.w));
  /* power * colour is formed here once (same products, same rounding as in the gather loop of a full map) */
  color_out[i] = make_float4(__fmul_rn(c.w, c.x), __fmul_rn(c.w, c.y), __fmul_rn(c.w, c.z), c.w);
  p.w = __int_as_float(i);
  pts[i] = p;
}

int cells_reorder(pm_ctx* ctx, pm_map* m) {
  if (!m->grid_pts) return PM_OK;
  const int n = (int)m->n;
  float4* sorted = nullptr;
  PM_CUDA(pm::map_alloc(&sorted, sizeof(float4) * (size_t)n, ctx->stream));
  k_cells_reorder<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(n, m->grid_pts, m->color4, sorted);
  ctx->launches++;
  PM_CUDA(cudaGetLastError());
  PM_CUDA(cudaFreeAsync(m->color4, ctx->stream));
  m->color4 = sorted;
  return PM_OK;
}

/* candidates expected inside the first radius: 1.25 k + 10. Swept on C2 (B200, ms per 1 M-query step): 1.1 k + 10
 * 0.806, 1.25 k + 10 0.797, 1.25 k + 20 0.803, 1.4 k + 10 0.822, 1.5 k + 16 0.827 — flat around the choice. */
static float fast_expected_count(int k) { return 1.25f * (float)k + 10.f; }

/* dimensions for m->cell (grown until the offset table fits), then: node ids sorted by cell, the cell-ordered records,
 * the offset table and — fast-path maps — its x-slowest copy. Leaves the map without a grid when the memory is not there. */
static int grid_fill_arrays(pm_ctx* ctx, pm_map* m, bool given) {
  const int n = (int)m->n;
  cudaStream_t st = ctx->stream;
  float cell = m->cell;
  float maxabs = 0.f;
  for (int d = 0; d < 3; d++) {
    maxabs = std::max(maxabs, std::max(std::fabs(m->bounds[d]), std::fabs(m->bounds[3 + d])));
  }
  int dim[3];
  /* a cells-only map is rebuilt every pass: keep its offset table (4 B per cell, cleared and filled per build) small */
  const int64_t cell_cap = std::min(kMaxCellCap, std::max(kMinCellCap, (int64_t)(given ? 8 : 32) * n));
  while (true) {
    const float inv = 1.f / cell;
    int64_t cells = 1;
    for (int d = 0; d < 3; d++) {
      dim[d] = host_cell_of(m->bounds[3 + d], m->bounds[d], inv) + 1;
      if (dim[d] < 1) dim[d] = 1;
      cells *= dim[d];
    }
    if (cells <= cell_cap && dim[0] < 32768 && dim[1] < 32768 && dim[2] < 32768) break;
    cell *= 1.26f;
  }
  m->cell = cell; m->inv_cell = 1.f / cell;
  m->cell_margin = 1e-3f * cell + 16.f * FLT_EPSILON * maxabs;
  m->gdim[0] = dim[0]; m->gdim[1] = dim[1]; m->gdim[2] = dim[2];
  const int64_t n_cells = (int64_t)dim[0] * dim[1] * dim[2];

  uint32_t *keys_in, *keys_out; int *vals_in, *vals_out;
  PM_CHECK(ctx->get_scratch(pm::S_KEYS0, (size_t)n, &keys_in));
  PM_CHECK(ctx->get_scratch(pm::S_KEYS1, (size_t)n, &keys_out));
  PM_CHECK(ctx->get_scratch(pm::S_VALS0, (size_t)n, &vals_in));
  PM_CHECK(ctx->get_scratch(pm::S_VALS1, (size_t)n, &vals_out));
  if (pm::map_alloc(&m->grid_pts, sizeof(float4) * ((size_t)n + 32), st) != cudaSuccess ||
      pm::map_alloc(&m->cell_start, sizeof(uint32_t) * (size_t)(n_cells + 1), st) != cudaSuccess) {
    cudaGetLastError();
    pm::map_free(m->grid_pts);
    m->grid_pts = nullptr; m->cell_start = nullptr; /* no grid: the kd-tree kernels answer everything */
    m->fast_path = false;
    if (given) { pm::set_error("pm_map_build_cells: out of device memory for the cell table"); return PM_ERR_NOMEM; }
    return PM_OK;
  }
  PM_CUDA(cudaMemsetAsync(m->grid_pts + n, 0, sizeof(float4) * 32, st)); /* read (and dropped) by lanes past a query's last candidate */
  GridView g = view_of(m);
  k_cell_ids<<<(unsigned)div_up(n, 256), 256, 0, st>>>(m->nodes4, n, g, keys_in, vals_in);
  int bits = 1;
  while (((int64_t)1 << bits) < n_cells) bits++;
  size_t tb = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tb, keys_in, keys_out, vals_in, vals_out, n, 0, bits, st);
  void* tmp = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_CUB, tb, &tmp));
  PM_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, keys_in, keys_out, vals_in, vals_out, n, 0, bits, st));
  k_grid_fill<<<(unsigned)div_up((int64_t)n + 1, 256), 256, 0, st>>>(m->nodes4, n, keys_out, vals_out, (uint32_t)n_cells,
                                                                    m->grid_pts, m->cell_start);
  ctx->launches += 6;
  PM_CUDA(cudaGetLastError());
  if (m->fast_path) {
    const int64_t words = (int64_t)(dim[0] + 1) * dim[1] * dim[2];
    if (pm::map_alloc(&m->cell_start_t, sizeof(uint32_t) * (size_t)words, st) != cudaSuccess) {
      cudaGetLastError();
      m->cell_start_t = nullptr;
      m->fast_path = false; /* the general kernel answers everything */
    } else {
      k_cells_transpose<<<(unsigned)div_up(words, 256), 256, 0, st>>>(m->cell_start, dim[0], dim[1], dim[2], m->cell_start_t);
      ctx->launches++;
      PM_CUDA(cudaGetLastError());
    }
  }
  return PM_OK;
}

__global__ void k_sample_to_sq(const float* __restrict__ q, int ns, float4* __restrict__ sq) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < ns) sq[i] = make_float4(q[3 * i], q[3 * i + 1], q[3 * i + 2], __int_as_float(i));
}

/* share of the sample the fast path may hand to the general kernel and still be the one to use */
constexpr float kFastTrialMaxHandover = 0.10f;

int grid_query(pm_ctx* ctx, const pm_map* map, const pmk::QueryIO& io, bool gather, int* fb_list, int* fb_count);

int grid_build(pm_ctx* ctx, pm_map* m, float cell_size) {
  const bool given = cell_size > 0.f; /* cells-only map: the caller names the cell, nothing is sampled (there is no tree) */
  if (!given && m->n < kMinGridNodes) return PM_OK;
  const int n = (int)m->n;
  cudaStream_t st = ctx->stream;
  if (given) {
    m->cell = cell_size; m->tuned_rk = 2.f * cell_size; m->fast_path = false;
    return grid_fill_arrays(ctx, m, true);
  }
  /* 1. sample of exact k-th-neighbour distances (kd-tree kernel on kTuneSamples node positions) */
  const int ns = kTuneSamples;
  float *sq = nullptr, *sr2 = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_MISC0, (size_t)ns * 3, &sq));
  PM_CHECK(ctx->get_scratch(pm::S_MISC3, (size_t)ns, &sr2));
  k_sample_queries<<<(unsigned)div_up(ns, 256), 256, 0, st>>>(m->nodes4, n, ns, sq);
  ctx->launches++;
  {
    pmk::QueryIO io;
    io.q = sq; io.nq = ns; io.k = kTuneK; io.cutoff2 = INFINITY; io.r2 = sr2;
    PM_CHECK(kd_query(ctx, m, io, false, nullptr, nullptr));
  }
  std::vector<float> r2(ns);
  PM_CUDA(cudaMemcpyAsync(r2.data(), sr2, sizeof(float) * ns, cudaMemcpyDeviceToHost, st));
  PM_CUDA(cudaStreamSynchronize(st));
  std::sort(r2.begin(), r2.end());
  const float r50 = std::sqrt(r2[ns / 2]), r90 = std::sqrt(r2[(size_t)(ns * 0.90)]), r98 = std::sqrt(r2[(size_t)(ns * 0.98)]);
  if (!(r98 > 0.f) || !std::isfinite(r98)) return PM_OK; /* degenerate cloud: kd-tree kernels only */
  m->tuned_r50 = r50;
  float maxabs0 = 0.f;
  for (int d = 0; d < 6; d++) maxabs0 = std::max(maxabs0, std::fabs(m->bounds[d]));
  /* The fast path (k_knn_fast) starts every query at ONE radius — the one expected to hold 1.25 kTuneK + 10 photons at
   * the median density — and needs two cells (less the margins) to cover it: see fast_first_radius2. Tight
   * distributions of the k-th-neighbour distance (uniform clouds: r98 < 1.25 r50) take it unseen. Broader ones — photon
   * maps proper: surfaces lit unevenly — take it if it proves itself: the sample is run through it, and it stays when it
   * hands over no more than a tenth. (Forced on the surface clouds of the C1-like 1024x1024 frame and of the C4 room it
   * answered everything itself and the gather kernels took 11.8 instead of 17.1 ms, the C4-like render 60.8 instead of
   * 84.2 ms: a query off the median density costs a second attempt at a rescaled radius, which is still cheaper than
   * the general kernel's density pass and histogram selection.) */
  const bool uniform_like = r98 < 1.25f * r50;
  const float r_first = r50 * std::cbrt(fast_expected_count(kTuneK) / ((float)kTuneK - 0.33f));
  const float cell_fast = (0.5f * r_first + 2.f * 16.f * FLT_EPSILON * maxabs0) / (1.f - 2.5e-3f);
  if (cell_fast > 0.f && std::isfinite(cell_fast)) {
    m->cell = cell_fast; m->tuned_rk = r98 * 1.03f; m->fast_path = true;
    PM_CHECK(grid_fill_arrays(ctx, m, false));
    if (!m->grid_pts) return PM_OK;
    if (m->fast_path && uniform_like) return PM_OK;
    if (m->fast_path) {
      float4* sq4 = nullptr; int32_t* ids = nullptr; int *fb_list = nullptr, *fb_count = nullptr;
      PM_CHECK(ctx->get_scratch(pm::S_VALS1, (size_t)ns, &sq4));
      PM_CHECK(ctx->get_scratch(pm::S_VALS0, (size_t)ns * kTuneK, &ids));
      PM_CHECK(ctx->get_scratch(pm::S_MISC1, (size_t)ns, &fb_list));
      PM_CHECK(ctx->get_scratch(pm::S_MISC2, (size_t)4, &fb_count));
      PM_CUDA(cudaMemsetAsync(fb_count, 0, sizeof(int), st));
      k_sample_to_sq<<<(unsigned)div_up(ns, 256), 256, 0, st>>>(sq, ns, sq4);
      ctx->launches++;
      pmk::QueryIO io;
      io.q = sq; io.sq = sq4; io.nq = ns; io.k = kTuneK; io.cutoff2 = INFINITY; io.ids = ids;
      PM_CHECK(grid_query(ctx, m, io, false, fb_list, fb_count));
      int handed = ns;
      if (ctx->ho_count_dev) PM_CUDA(cudaMemcpyAsync(&handed, ctx->ho_count_dev, sizeof(int), cudaMemcpyDeviceToHost, st));
      PM_CUDA(cudaStreamSynchronize(st));
      ctx->ho_count_dev = nullptr;
      if ((float)handed <= kFastTrialMaxHandover * (float)ns) return PM_OK;
    }
    /* not this cloud: drop the grid and build the general kernel's */
    pm::map_free(m->grid_pts); pm::map_free(m->cell_start); pm::map_free(m->cell_start_t);
    m->grid_pts = nullptr; m->cell_start = nullptr; m->cell_start_t = nullptr;
  }
  /* G >= 2 cells: two cells must cover the k-th-neighbour distance of nearly every query; a broad distribution would
   * pay for its tail in every dense neighbourhood, so it takes the 90th percentile and leaves the rest to the kd-tree
   * kernel. */
  m->fast_path = false;
  m->tuned_rk = r90;
  m->cell = 0.5f * r90;
  /* degenerate cloud (most sampled neighbourhoods are coincident photons): kd-tree kernels only */
  if (!(m->cell > 0.f) || !std::isfinite(m->cell)) return PM_OK;
  return grid_fill_arrays(ctx, m, false);
}

/* first search radius² of the fast path for k neighbours: the radius expected to hold 1.25 k + 10 photons, from the
 * sampled median distance to the kTuneK-th neighbour (volume ~ count), capped by what a 5x5x5 block guarantees */
static float fast_first_radius2(const pm_map* m, int k, float cutoff2) {
  const float ek = fast_expected_count(k);
  const float r = m->tuned_r50 * std::cbrt(ek / ((float)kTuneK - 0.33f));
  const float rmax = 2.f * m->cell - 4.f * m->cell_margin;
  const float rr = std::min(r, rmax);
  return std::min(rr * rr, cutoff2);
}

int grid_query(pm_ctx* ctx, const pm_map* map, const pmk::QueryIO& io, bool gather, int* fb_list, int* fb_count) {
  const GridView g = view_of(map);
  const int64_t cap = (int64_t)ctx->sm_count * 64;
  if (map->fast_path && io.sq && io.k <= kBuf / 2) {
    /* uniform-like cloud: k_knn_fast, k_grid_query only for the handed-over queries */
    int *ho_list = nullptr, *ho_count = nullptr;
    PM_CHECK(ctx->get_scratch(pm::S_FAST1, (size_t)io.nq, &ho_list));
    PM_CHECK(ctx->get_scratch(pm::S_FAST2, (size_t)4, &ho_count));
    PM_CUDA(cudaMemsetAsync(ho_count, 0, sizeof(int), ctx->stream));
    FastParams fp;
    fp.sq = io.sq;
    fp.R2 = fast_first_radius2(map, io.k, io.cutoff2);
    fp.margin_c = map->cell_margin * map->inv_cell;
    const float rc = std::sqrt(fp.R2) * map->inv_cell + 2.f * fp.margin_c;
    fp.Rc2 = rc * rc;
    for (int n = 0; n < kWinSize; n++)
      fp.win[n] = std::pow(((float)io.k + 0.5f) / (float)std::max(n, io.k + 1), 2.f / 3.f);
    const unsigned fgrid = (unsigned)std::min<int64_t>(div_up(io.nq, kWarps), (int64_t)ctx->sm_count * 8 * 4);
    if (gather) k_knn_fast<true><<<fgrid, kWarps * 32, 0, ctx->stream>>>(g, map->color4, io, fp, ho_list, ho_count);
    else k_knn_fast<false><<<fgrid, kWarps * 32, 0, ctx->stream>>>(g, map->color4, io, fp, ho_list, ho_count);
    pmk::QueryIO rest = io;
    rest.order = ho_list;
    const unsigned grid = (unsigned)std::min<int64_t>(div_up(io.nq, kWarps), (int64_t)ctx->sm_count * 8);
    if (gather) k_grid_query<true><<<grid, kWarps * 32, 0, ctx->stream>>>(g, map->color4, rest, ho_count, fb_list, fb_count);
    else k_grid_query<false><<<grid, kWarps * 32, 0, ctx->stream>>>(g, map->color4, rest, ho_count, fb_list, fb_count);
    ctx->launches += 2;
    ctx->ho_count_dev = ho_count;
    PM_CUDA(cudaGetLastError());
    return PM_OK;
  }
  ctx->ho_count_dev = nullptr;
  int64_t grid = div_up(io.nq, kWarps);
  if (grid > cap) grid = cap;
  if (gather) k_grid_query<true><<<(unsigned)grid, kWarps * 32, 0, ctx->stream>>>(g, map->color4, io, nullptr, fb_list, fb_count);
  else k_grid_query<false><<<(unsigned)grid, kWarps * 32, 0, ctx->stream>>>(g, map->color4, io, nullptr, fb_list, fb_count);
  ctx->launches++;
  PM_CUDA(cudaGetLastError());
  return PM_OK;
}

int radius_query(pm_ctx* ctx, const pm_map* map, const float* q, const float* brdf, const float* radius2, const int* order,
                 int nq, float* out_flux, int32_t* out_count, int* fb_list, int* fb_count) {
  int64_t kd_grid = div_up(nq, 128);
  if (map->grid_pts) {
    const GridView g = view_of(map);
    int64_t grid = div_up(nq, kWarps);
    const int64_t cap = (int64_t)ctx->sm_count * 64;
    if (grid > cap) grid = cap;
    /* a cells-only map has no tree to fall back on: its blocks may grow to the whole grid (then every side is
     * clipped and the block covers any ball), so nothing is ever handed over */
    const int max_h = map->cells_only ? std::max(map->gdim[0], std::max(map->gdim[1], map->gdim[2])) : kMaxHRadius;
    /* a sample of the hit points decides between one thread and one warp per hit point (see k_grid_radius_thread) */
    int* heavy_list = nullptr;
    RadiusPick* pick = nullptr;
    PM_CHECK(ctx->get_scratch(pm::S_FAST1, (size_t)nq, &heavy_list));
    PM_CHECK(ctx->get_scratch(pm::S_FAST2, (size_t)1, &pick));
    PM_CUDA(cudaMemsetAsync(pick, 0, sizeof(RadiusPick), ctx->stream));
    ctx->ho_count_dev = &pick->heavy; /* reported as the call's hand-over count */
    const int stride = std::max(1, nq / kRadiusSamples);
    const unsigned sgrid = (unsigned)div_up(div_up(nq, stride) * 32, 256), tgrid = (unsigned)div_up(nq, 256);
    const unsigned ggrid = (unsigned)std::min<int64_t>(div_up(div_up(nq, kGroup), kWarps), cap);
    if (map->cells_only) {
      k_grid_radius_thread<true, true><<<sgrid, 256, 0, ctx->stream>>>(g, map->color4, q, brdf, radius2, order, nq, stride, nullptr,
                                                                       nullptr, nullptr, pick);
      k_grid_radius_thread<false, true><<<tgrid, 256, 0, ctx->stream>>>(g, map->color4, q, brdf, radius2, order, nq, 1, out_flux,
                                                                        out_count, heavy_list, pick);
      k_grid_radius_group<true><<<ggrid, kWarps * 32, 0, ctx->stream>>>(g, map->color4, q, brdf, radius2, order, nq, max_h, out_flux,
                                                                       out_count, heavy_list, pick);
      k_grid_radius<true><<<(unsigned)grid, kWarps * 32, 0, ctx->stream>>>(g, map->color4, q, brdf, radius2, order, nq, heavy_list, pick,
                                                                          max_h, out_flux, out_count, fb_list, fb_count);
    } else {
      k_grid_radius_thread<true, false><<<sgrid, 256, 0, ctx->stream>>>(g, map->color4, q, brdf, radius2, order, nq, stride, nullptr,
                                                                        nullptr, nullptr, pick);
      k_grid_radius_thread<false, false><<<tgrid, 256, 0, ctx->stream>>>(g, map->color4, q, brdf, radius2, order, nq, 1, out_flux,
                                                                         out_count, heavy_list, pick);
      k_grid_radius_group<false><<<ggrid, kWarps * 32, 0, ctx->stream>>>(g, map->color4, q, brdf, radius2, order, nq, max_h, out_flux,
                                                                        out_count, heavy_list, pick);
      k_grid_radius<false><<<(unsigned)grid, kWarps * 32, 0, ctx->stream>>>(g, map->color4, q, brdf, radius2, order, nq, heavy_list, pick,
                                                                           max_h, out_flux, out_count, fb_list, fb_count);
    }
    ctx->launches += 3;
    if (map->cells_only) { ctx->launches++; PM_CUDA(cudaGetLastError()); return PM_OK; }
    if (kd_grid > ctx->sm_count * 2) kd_grid = ctx->sm_count * 2;
    k_kd_radius<<<(unsigned)kd_grid, 128, 0, ctx->stream>>>(map->nodes4, map->color4, (int)map->n, q, brdf, radius2, nq, fb_list,
                                                           fb_count, out_flux, out_count);
    ctx->launches += 2;
  } else {
    if (kd_grid > ctx->sm_count * 64) kd_grid = ctx->sm_count * 64;
    k_kd_radius<<<(unsigned)kd_grid, 128, 0, ctx->stream>>>(map->nodes4, map->color4, (int)map->n, q, brdf, radius2, nq, nullptr,
                                                           nullptr, out_flux, out_count);
    ctx->launches++;
  }
  PM_CUDA(cudaGetLastError());
  return PM_OK;
}

}  // namespace pm
