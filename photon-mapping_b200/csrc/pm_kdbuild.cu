/* pm_kdbuild.cu — stage 2: photon-map construction.
 *
 * Produces exactly the array cukd::buildTree<Photon,Photon_traits> produces
 * (reference: externals/cudaKDTree/cukd/builder_thrust.h:414-512, called from
 * ray-tracer/src/hostCode.cu:94-95): a left-balanced kd-tree stored in level order (children of
 * node i are 2i+1 / 2i+2), each node's split dimension = widest axis of the node's DOMAIN box
 * (world AABB clipped by its ancestors' planes; ties → highest axis, cukd/box.h:112-125), pivot =
 * element of rank |left subtree| among the node's elements ordered by (coord[dim], input index).
 *
 * The reference re-sorts all N 48-byte (tag, photon) tuples with a comparison sort once per level.
 * This builder never moves photons until the end:
 *   1. three radix sorts (one per axis) of 4-byte element indices by order-preserving float keys;
 *   2. per level, for every live subtree: pick the pivot from the list sorted along the subtree's
 *      split axis (its position is known in closed form), mark every element left/right, and
 *      stable-partition the three index lists inside each subtree's range (prefix sums);
 *   3. one gather writes the 44-byte records (and the 16-byte query-side copies) in node order.
 * All three lists share the same segment geometry: positions are IN-ORDER ranks, so a subtree's
 * range is [begin, begin+size) and its pivot sits at begin + |left subtree|; the slot a pivot
 * leaves behind simply stays dead.
 */
#include <cstdlib>
#include <cstring>
#include <cub/cub.cuh>
#include "pm_internal.h"

namespace {

constexpr int kDead = -1;

__host__ __device__ inline int64_t subtree_size(int64_t root, int64_t n) {
  int64_t total = 0, first = root, width = 1;
  while (first < n) {
    const int64_t last = (first + width < n) ? first + width : n;
    total += last - first;
    first = 2 * first + 1;
    width *= 2;
  }
  return total;
}

__device__ inline uint32_t ordered_key(float f) {
  f = f + 0.0f;  /* -0.0f -> +0.0f so that both zeros tie and fall back to index order */
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

/* AoS -> keys for the three axis sorts, and iota values */
__global__ void k_extract(const pm_photon44* __restrict__ ph, int n, float* __restrict__ cx, float* __restrict__ cy,
                          float* __restrict__ cz, uint32_t* __restrict__ kx, uint32_t* __restrict__ ky,
                          uint32_t* __restrict__ kz, int* __restrict__ iota, uint4* __restrict__ keys4 /* or NULL */) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* p = reinterpret_cast<const float*>(ph) + (size_t)i * 11;
  const float x = p[0], y = p[1], z = p[2];
  cx[i] = x; cy[i] = y; cz[i] = z;
  const uint32_t a = ordered_key(x), b = ordered_key(y), c = ordered_key(z);
  kx[i] = a; ky[i] = b; kz[i] = c;
  iota[i] = i;
  if (keys4) keys4[i] = make_uint4(a, b, c, 0u); /* all three in one record: k_carry_init gathers it once per list entry */
}

/* bounds from the sorted lists: min = first element, max = last element of each axis list */
__global__ void k_root(const int* __restrict__ ox, const int* __restrict__ oy, const int* __restrict__ oz,
                       const float* __restrict__ cx, const float* __restrict__ cy, const float* __restrict__ cz, int n,
                       float* __restrict__ seg_box, int* __restrict__ seg_begin, float* __restrict__ bounds_out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const float lo[3] = {cx[ox[0]], cy[oy[0]], cz[oz[0]]};
  const float hi[3] = {cx[ox[n - 1]], cy[oy[n - 1]], cz[oz[n - 1]]};
  for (int d = 0; d < 3; d++) { seg_box[d] = lo[d]; seg_box[3 + d] = hi[d]; bounds_out[d] = lo[d]; bounds_out[3 + d] = hi[d]; }
  seg_begin[0] = 0;
}

__device__ inline int widest_dim(const float* b) {
  int best = 0; float wbest = 0.f;
#pragma unroll
  for (int d = 0; d < 3; d++) {
    const float w = b[3 + d] - b[d];
    if (w < wbest) continue;
    wbest = w; best = d;
  }
  return best;
}

/* what k_level needs to know about one subtree range of the current level */
struct SegRec {
  int pivot_dim; /* pivot position | split axis << 30 */
  int begin;
  uint32_t pkey, pelem; /* the pivot in the order the lists were sorted in: (ordered key, input index) */
};

/* one thread per subtree of the current level: choose dim + pivot, emit the node, derive children */
__global__ void k_split(int level_first, int j0, int level_count, int n, const int* __restrict__ ord0,
                        const int* __restrict__ ord1, const int* __restrict__ ord2, const float* __restrict__ c0,
                        const float* __restrict__ c1, const float* __restrict__ c2,
                        const float* __restrict__ box_cur, const int* __restrict__ begin_cur,
                        float* __restrict__ box_next, int* __restrict__ begin_next, SegRec* __restrict__ seg_rec,
                        int* __restrict__ elem_of_node, uint8_t* __restrict__ dim_of_node) {
  /* subtrees [j0, level_count) of the level: a rank of a sharded build owns a contiguous range of them */
  const int j = j0 + blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= level_count) return;
  const int s = level_first + j;
  float box[6];
#pragma unroll
  for (int d = 0; d < 6; d++) box[d] = box_cur[(size_t)j * 6 + d];
  const int dim = widest_dim(box);
  const int begin = begin_cur[j];
  const int lchild = 2 * s + 1;
  const int left = (lchild < n) ? (int)subtree_size(lchild, n) : 0;
  const int pivot = begin + left;
  const int* ord = dim == 0 ? ord0 : (dim == 1 ? ord1 : ord2);
  const float* c = dim == 0 ? c0 : (dim == 1 ? c1 : c2);
  const int e = ord[pivot];
  const float split = c[e];
  elem_of_node[s] = e;
  dim_of_node[s] = (uint8_t)dim;
  seg_rec[j] = SegRec{pivot | (dim << 30), begin, ordered_key(split), (uint32_t)e};
  /* children (index in the next level's arrays = 2j, 2j+1) */
  const int next_first = 2 * level_first + 1;
  if (lchild < n) {
    const size_t o = (size_t)(lchild - next_first) * 6;
#pragma unroll
    for (int d = 0; d < 6; d++) box_next[o + d] = box[d];
    box_next[o + 3 + dim] = fminf(split, box[3 + dim]);
    begin_next[lchild - next_first] = begin;
  }
  if (lchild + 1 < n) {
    const size_t o = (size_t)(lchild + 1 - next_first) * 6;
#pragma unroll
    for (int d = 0; d < 6; d++) box_next[o + d] = box[d];
    box_next[o + dim] = fmaxf(split, box[dim]);
    begin_next[lchild + 1 - next_first] = pivot + 1;
  }
}

/* ---- one level in one pass --------------------------------------------------------------------
 * k_level moves every element of the three lists to its place on the next level: a stable partition of each
 * subtree range around its pivot. In the split-axis list the side of an element is its position; in the other
 * two it is compared with the pivot by (ordered coordinate key, input index), the order the lists were sorted
 * in — one 4-byte gather per list. A tile of kTile positions computes the sides, scans them in shared memory
 * and resolves what its first (spilled-over) subtree range
 * accumulated in earlier tiles with a decoupled look-back over one 64-bit status word per tile. The scan is
 * segmented (it restarts at every range begin and at every dead slot), so a tile that holds a range begin
 * publishes a resolved value without waiting for anybody, and the look-back never walks past the previous
 * range begin. Only the two lists that are not sorted along the split axis move (slot 0 = axis dim+1, slot 1 =
 * axis dim+2); per slot the word carries the lefts seen so far in the range (30 bits) and whether the pivot
 * was seen (1 bit): rights = (i - begin) - lefts - pivot_seen. Tiles take their number from a ticket counter,
 * so a tile only ever waits for tiles that are already running. */
/* A tile is kLvItems stripes of kLvThreads consecutive positions: thread t owns position t of every stripe, so all
 * loads and the in-place store are coalesced, a thread has kLvItems independent load chains in flight (position ->
 * subtree record -> two coordinate gathers), and the fixed cost of a tile (ticket, three barriers, look-back) is paid
 * once per 1 024 positions. With one position per thread (round 1 .. r2l) the kernel ran at 22 % of the DRAM rate, 34 % of
 * L2 and IPC 1.5 at 95 % occupancy: bound by those per-tile latencies, not by any unit. */
constexpr int kLvThreads = 256, kLvWarps = kLvThreads / 32;
constexpr int kLvItems = 4;       /* lists of element ids only: 1 024 positions per tile */
/* measured on B200 (tools/time_build.py, whole build, with / without carried keys): 5 M 3.87 / 3.74 ms, 10 M 7.28 / 7.17,
 * 20 M 14.41 / 14.59, 30 M 20.87 / 22.11, 50 M 35.93 / 40.93, 100 M 73.3 / 100.5 */
constexpr int64_t kCarryMinPhotons = (int64_t)1 << 24;
int64_t g_carry_min_photons = kCarryMinPhotons; /* pm_test_build_carry_from */
constexpr int kLvItemsCarry = 2;  /* lists that carry their keys (below): 36 instead of 12 bytes per position in registers */
constexpr uint64_t kStInvalid = 0ull, kStAgg = 1ull << 62, kStPrefix = 2ull << 62, kStMask = 3ull << 62;

__device__ inline uint64_t tile_to_global(uint64_t x) { /* 4 x 11-bit tile-local fields -> L0 | P0<<30 | L1<<31 | P1<<61 */
  return (x & 0x7ffull) | (((x >> 11) & 0x7ffull) << 30) | (((x >> 22) & 0x7ffull) << 31) | (((x >> 33) & 0x7ffull) << 61);
}

/* CARRY: every entry of the list sorted along axis d carries, next to the element id, the element's ordered keys
 * along the other two axes ((d+1)%3 in ka, (d+2)%3 in kb). The side of an element is then decided from what the pass
 * streams anyway; without them it takes a 4-byte gather of the element's coordinate per list — free while the
 * coordinate arrays sit in L2 (10 M photons: 120 MB), a 64-byte DRAM access each once they do not (100 M photons:
 * 181 bytes of DRAM traffic per position and level, 3.5 ms per level; with the keys 80 bytes, all of it streamed). */
struct CarriedKeys {
  const uint32_t* ka_in[3]; const uint32_t* kb_in[3];
  uint32_t* ka_out[3]; uint32_t* kb_out[3];
};

template <bool CARRY, int ITEMS>
__global__ void __launch_bounds__(kLvThreads, CARRY ? 3 : 4)
k_level(int n, int pos_lo, int pos_hi, int tile0, int level_first, const int* __restrict__ seg_of_pos, const SegRec* __restrict__ seg_rec,
        const int* __restrict__ in0, const int* __restrict__ in1, const int* __restrict__ in2,
        const float* __restrict__ c0, const float* __restrict__ c1, const float* __restrict__ c2,
        int* __restrict__ out0, int* __restrict__ out1, int* __restrict__ out2, int* __restrict__ seg_next,
        unsigned long long* __restrict__ status, unsigned int* __restrict__ ticket, const CarriedKeys ck) {
  constexpr int kTileT = kLvThreads * ITEMS;   /* tile-local counts <= 1 024 fit the 11-bit fields */
  constexpr int kParts = ITEMS * kLvWarps;     /* warp totals of a tile in scan order: stripe-major */
  static_assert(kParts <= 32 && kTileT <= 1024, "one warp scans the warp totals; 11-bit fields");
  __shared__ uint64_t scanv[kTileT];
  __shared__ uint64_t warp_tot[32];
  __shared__ uint64_t carry_sh;
  __shared__ int tile_sh, last_head_sh, first_is_head_sh;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) { tile_sh = (int)atomicAdd(ticket, 1u); last_head_sh = -1; }
  __syncthreads();
  /* a sharded build passes only over [pos_lo, pos_hi) — the range of the rank's own subtree, which starts with a
   * range begin, so its first tile never looks back; positions outside are other ranks' and count as dead */
  const int tile = tile_sh;
  const int tile_base = (tile0 + tile) * kTileT;

  int s[ITEMS], e0[ITEMS], e1[ITEMS], e2[ITEMS];
  uint32_t ka[CARRY ? ITEMS : 1][3], kb[CARRY ? ITEMS : 1][3];
#pragma unroll
  for (int it = 0; it < ITEMS; it++) {
    const int i = tile_base + it * kLvThreads + tid;
    s[it] = kDead; e0[it] = e1[it] = e2[it] = 0;
    if (i >= pos_lo && i < pos_hi) {
      s[it] = seg_of_pos[i]; e0[it] = in0[i]; e1[it] = in1[i]; e2[it] = in2[i];
      if (CARRY) {
#pragma unroll
        for (int d = 0; d < 3; d++) { ka[it][d] = ck.ka_in[d][i]; kb[it][d] = ck.kb_in[d][i]; }
      }
    }
  }
  /* per position: side of the element in slot 0 (bits 0-1) and slot 1 (bits 2-3); 2 = it is the pivot */
  uint32_t sides[ITEMS];
  uint64_t f[ITEMS];
  int my_last_head = -1;
#pragma unroll
  for (int it = 0; it < ITEMS; it++) {
    const int i = tile_base + it * kLvThreads + tid;
    f[it] = 0ull; sides[it] = 0xau;
    bool head = true; /* a dead slot restarts the scan like a range begin */
    if (s[it] != kDead) {
      const int4 rec = *reinterpret_cast<const int4*>(seg_rec + (s[it] - level_first));
      const int pivot = rec.x & 0x3fffffff, dim = (int)((uint32_t)rec.x >> 30), begin = rec.y;
      const uint2 pk = make_uint2((uint32_t)rec.z, (uint32_t)rec.w);
      const uint32_t ea = (uint32_t)(dim == 0 ? e1[it] : (dim == 1 ? e2[it] : e0[it]));
      const uint32_t eb = (uint32_t)(dim == 0 ? e2[it] : (dim == 1 ? e0[it] : e1[it]));
      uint32_t key_a, key_b; /* the two elements' keys along the split axis */
      if (CARRY) {
        /* slot 0 is list (dim+1)%3, whose kb is axis dim; slot 1 is list (dim+2)%3, whose ka is axis dim */
        key_a = dim == 0 ? kb[it][1] : (dim == 1 ? kb[it][2] : kb[it][0]);
        key_b = dim == 0 ? ka[it][2] : (dim == 1 ? ka[it][0] : ka[it][1]);
      } else {
        const float* c = dim == 0 ? c0 : (dim == 1 ? c1 : c2);
        key_a = ordered_key(c[ea]); key_b = ordered_key(c[eb]);
      }
      const uint32_t sd_a = ea == pk.y ? 2u : ((key_a < pk.x || (key_a == pk.x && ea < pk.y)) ? 0u : 1u);
      const uint32_t sd_b = eb == pk.y ? 2u : ((key_b < pk.x || (key_b == pk.x && eb < pk.y)) ? 0u : 1u);
      seg_next[i] = i < pivot ? 2 * s[it] + 1 : (i > pivot ? 2 * s[it] + 2 : kDead);
      f[it] = (sd_a == 0u ? 1ull : (sd_a == 2u ? 1ull << 11 : 0ull)) + (sd_b == 0u ? 1ull << 22 : (sd_b == 2u ? 1ull << 33 : 0ull));
      sides[it] = sd_a | (sd_b << 2);
      head = i == begin;
    } else if (i >= pos_lo && i < pos_hi) {
      seg_next[i] = kDead;
    }
    if (head) my_last_head = it * kLvThreads + tid;
    if (it == 0 && tid == 0) first_is_head_sh = head ? 1 : 0;
  }
  {
    const int wl = __reduce_max_sync(0xffffffffu, my_last_head);
    if (lane == 0 && wl >= 0) atomicMax(&last_head_sh, wl);
  }

  /* exclusive scan of f over the tile in position order: stripe by stripe, warp by warp */
  uint64_t inc[ITEMS];
#pragma unroll
  for (int it = 0; it < ITEMS; it++) {
    uint64_t v = f[it];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint64_t u = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += u;
    }
    inc[it] = v;
    if (lane == 31) warp_tot[it * kLvWarps + wid] = v;
  }
  __syncthreads();
  if (wid == 0) {
    uint64_t w = lane < kParts ? warp_tot[lane] : 0ull;
#pragma unroll
    for (int o = 1; o < kParts; o <<= 1) {
      const uint64_t v = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += v;
    }
    if (lane < kParts) warp_tot[lane] = w; /* inclusive over the parts */
  }
  __syncthreads();
  uint64_t ex[ITEMS];
#pragma unroll
  for (int it = 0; it < ITEMS; it++) {
    const int part = it * kLvWarps + wid;
    ex[it] = inc[it] - f[it] + (part ? warp_tot[part - 1] : 0ull);
    scanv[it * kLvThreads + tid] = ex[it];
  }
  const uint64_t total = warp_tot[kParts - 1];
  const int last_head = last_head_sh;
  const bool need_carry = !first_is_head_sh;
  __syncthreads();

  /* publish, and look back if the tile starts inside a range */
  if (wid == 0) {
    unsigned long long* mine = status + tile;
    if (last_head >= 0) {
      if (lane == 0) atomicExch(mine, kStPrefix | tile_to_global(total - scanv[last_head]));
    } else {
      if (lane == 0) atomicExch(mine, kStAgg | tile_to_global(total));
    }
    if (need_carry) {
      uint64_t carry = 0ull;
      int look = tile - 1;
      while (true) {
        const int t = look - lane;
        uint64_t w = kStPrefix;
        if (t >= 0) {
          do { w = *reinterpret_cast<volatile unsigned long long*>(status + t); } while ((w & kStMask) == kStInvalid);
        }
        const unsigned resolved = __ballot_sync(0xffffffffu, (w & kStMask) == kStPrefix);
        const int stop = resolved ? __ffs(resolved) - 1 : 32; /* nearest resolved tile */
        uint64_t v = lane <= stop ? (w & ~kStMask) : 0ull;
#pragma unroll
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        carry += v;
        if (resolved) break;
        look -= 32;
      }
      if (lane == 0) {
        carry_sh = carry;
        if (last_head < 0) atomicExch(mine, kStPrefix | (carry + tile_to_global(total)));
      }
    }
  }
  if (need_carry) __syncthreads();

#pragma unroll
  for (int it = 0; it < ITEMS; it++) {
    if (s[it] == kDead) continue;
    const int i = tile_base + it * kLvThreads + tid;
    const int4 rec = *reinterpret_cast<const int4*>(seg_rec + (s[it] - level_first)); /* again: an L1 hit, four registers less */
    const int pivot = rec.x & 0x3fffffff, dim = (int)((uint32_t)rec.x >> 30), begin = rec.y;
    uint32_t La, Pa, Lb, Pb;
    if (begin >= tile_base) {
      const uint64_t rel = ex[it] - scanv[begin - tile_base]; /* field-wise, no borrows: the fields are monotone */
      La = (uint32_t)(rel & 0x7ffu); Pa = (uint32_t)((rel >> 11) & 0x7ffu);
      Lb = (uint32_t)((rel >> 22) & 0x7ffu); Pb = (uint32_t)((rel >> 33) & 0x7ffu);
    } else {
      const uint64_t g = carry_sh + tile_to_global(ex[it]);
      La = (uint32_t)(g & 0x3fffffffu); Pa = (uint32_t)((g >> 30) & 1u);
      Lb = (uint32_t)((g >> 31) & 0x3fffffffu); Pb = (uint32_t)((g >> 61) & 1u);
    }
    const uint32_t sd_a = sides[it] & 3u, sd_b = sides[it] >> 2;
    const int off = i - begin;
    const int dst_a = sd_a == 0u ? begin + (int)La : (sd_a == 1u ? pivot + 1 + (off - (int)La - (int)Pa) : pivot);
    const int dst_b = sd_b == 0u ? begin + (int)Lb : (sd_b == 1u ? pivot + 1 + (off - (int)Lb - (int)Pb) : pivot);
    /* list dim stays in place, list (dim+1)%3 goes to dst_a, list (dim+2)%3 to dst_b */
    const int d0 = dim == 0 ? i : (dim == 1 ? dst_b : dst_a);
    const int d1 = dim == 1 ? i : (dim == 2 ? dst_b : dst_a);
    const int d2 = dim == 2 ? i : (dim == 0 ? dst_b : dst_a);
    out0[d0] = e0[it]; out1[d1] = e1[it]; out2[d2] = e2[it];
    if (CARRY) {
      ck.ka_out[0][d0] = ka[it][0]; ck.kb_out[0][d0] = kb[it][0];
      ck.ka_out[1][d1] = ka[it][1]; ck.kb_out[1][d1] = kb[it][1];
      ck.ka_out[2][d2] = ka[it][2]; ck.kb_out[2][d2] = kb[it][2];
    }
  }
}

/* the carried keys of the freshly sorted lists: one 16-byte gather per list entry, once */
__global__ void k_carry_init(int n, const int* __restrict__ ord0, const uint4* __restrict__ keys4,
                             uint32_t* __restrict__ ka0, uint32_t* __restrict__ kb0) {
  /* ord0 / ka0 / kb0 are the bases of 3 n entries: list d at d * n */
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 3 * (int64_t)n) return;
  const int d = (int)(t / n);
  const uint4 k = __ldg(keys4 + ord0[t]);
  ka0[t] = d == 0 ? k.y : (d == 1 ? k.z : k.x);
  kb0[t] = d == 0 ? k.z : (d == 1 ? k.x : k.y);
}

/* ---- the last kLocalLevels levels: one CTA per subtree, everything in shared memory ----------
 * Once a subtree holds at most kLocalCap elements its three index lists, coordinates and segment tables
 * fit in shared memory, and the remaining levels (the same split / mark / stable-partition steps as
 * above, on 16-bit local ids) cost no global traffic beyond one read of the lists and the node writes.
 * At 10 M photons this replaces 10 of 24 level passes over the whole array. */
constexpr int kLocalLevels = 10;
constexpr int kLocalCap = (1 << kLocalLevels) - 1; /* elements of a subtree with kLocalLevels levels */
constexpr int kLocalThreads = 256;
constexpr int kLocalItems = 4;                     /* positions per thread: 256 * 4 >= kLocalCap */
constexpr int kLocalSegs = 1 << (kLocalLevels - 1);
constexpr uint16_t kDead16 = 0xffffu;

struct LocalSmem {
  float c[3][kLocalCap + 1];
  int gid[kLocalCap + 1];
  uint64_t scanv[kLocalCap + 1];
  uint16_t lst[2][3][kLocalCap + 1];
  uint16_t segid[kLocalCap + 1];
  uint8_t side_el[kLocalCap + 1];   /* side of every element on the current level: 0 left, 1 right, 2 pivot */
  float box_even[kLocalSegs / 2][6]; /* boxes of even local levels (<= 256 nodes at level 8) */
  float box_odd[kLocalSegs / 4][6];  /* odd local levels (<= 128 at level 7); level 9 keeps only its split axis */
  uint8_t last_dim[kLocalSegs];
  uint16_t seg_begin[2][kLocalSegs];
  uint16_t seg_pivot[kLocalSegs];
  uint8_t seg_dim[kLocalSegs];
};

__device__ __forceinline__ float key_to_float(uint32_t k) { /* inverse of ordered_key (-0 comes back as +0) */
  return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k);
}

/* CARRY: the lists carry their keys (see k_level): the coordinates come out of the list entries — y and z from list 0's
 * entry, x from list 1's — instead of three scattered 4-byte reads per element */
template <bool CARRY>
__global__ void __launch_bounds__(kLocalThreads)
k_finish_subtrees(int level_first, int j0, int level_count, int n, const int* __restrict__ ord0, const int* __restrict__ ord1,
                  const int* __restrict__ ord2, const float* __restrict__ cx, const float* __restrict__ cy,
                  const float* __restrict__ cz, const float* __restrict__ box_cur, const int* __restrict__ begin_cur,
                  int* __restrict__ elem_of_node, uint8_t* __restrict__ dim_of_node, const CarriedKeys ck) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  LocalSmem& sm = *reinterpret_cast<LocalSmem*>(smem_raw);
  using Scan = cub::BlockScan<uint64_t, kLocalThreads>;
  __shared__ typename Scan::TempStorage scan_tmp;
  const int j = j0 + blockIdx.x;
  if (j >= level_count) return;
  const int s0 = level_first + j;
  const int begin0 = begin_cur[j];
  const int m = (int)subtree_size(s0, n);
  const int tid = threadIdx.x;

  /* element id -> local index (its position in list 0) for the other two lists: an open-addressing table in shared
   * memory, keys in the scan buffer and values in the second list buffer, neither of which is in use yet (2048
   * slots for at most 1023 keys). It replaces a 4-byte scatter to and two 4-byte gathers from a global array per
   * element, each a DRAM sector of its own. */
  int* hkeys = reinterpret_cast<int*>(sm.scanv);
  uint16_t* hvals = &sm.lst[1][0][0];
  constexpr int kHashBits = 11, kHashSlots = 1 << kHashBits;
  static_assert(sizeof(sm.scanv) >= kHashSlots * sizeof(int) && sizeof(sm.lst[1]) >= kHashSlots * sizeof(uint16_t), "hash table space");
  for (int i = tid; i < kHashSlots; i += kLocalThreads) hkeys[i] = -1;
  if (tid < 6) sm.box_even[0][tid] = box_cur[(size_t)j * 6 + tid];
  if (tid == 0) sm.seg_begin[0][0] = 0;
  __syncthreads();
  for (int i = tid; i < m; i += kLocalThreads) {
    const int e = ord0[begin0 + i];
    sm.gid[i] = e;
    if (CARRY) { sm.c[1][i] = key_to_float(ck.ka_in[0][begin0 + i]); sm.c[2][i] = key_to_float(ck.kb_in[0][begin0 + i]); }
    else { sm.c[0][i] = cx[e]; sm.c[1][i] = cy[e]; sm.c[2][i] = cz[e]; }
    sm.lst[0][0][i] = (uint16_t)i;
    sm.segid[i] = 0;
    uint32_t h = ((uint32_t)e * 2654435761u) >> (32 - kHashBits);
    while (atomicCAS(&hkeys[h], -1, e) != -1) h = (h + 1) & (kHashSlots - 1);
    hvals[h] = (uint16_t)i;
  }
  __syncthreads();
  for (int i = tid; i < m; i += kLocalThreads) {
#pragma unroll
    for (int d = 1; d < 3; d++) {
      const int e = (d == 1 ? ord1 : ord2)[begin0 + i];
      uint32_t h = ((uint32_t)e * 2654435761u) >> (32 - kHashBits);
      while (hkeys[h] != e) h = (h + 1) & (kHashSlots - 1);
      sm.lst[0][d][i] = hvals[h];
      if (CARRY && d == 1) sm.c[0][hvals[h]] = key_to_float(ck.kb_in[1][begin0 + i]); /* list 1 carries (z, x) */
    }
  }
  __syncthreads();

  int cur = 0;
  for (int t = 0;; t++) {
    const int64_t first = ((int64_t)(s0 + 1) << t) - 1; /* first node of this subtree's level t */
    if (first >= n) break;
    const int nseg = (int)min((int64_t)1 << t, (int64_t)n - first);
    const int tb = t & 1;
    /* split: one thread per node of the level */
    for (int q = tid; q < nseg; q += kLocalThreads) {
      const int s = (int)first + q;
      const bool last = t + 1 == kLocalLevels;
      float box[6];
      int dim;
      if (last) dim = sm.last_dim[q];
      else {
        const float* src = tb ? sm.box_odd[q] : sm.box_even[q];
#pragma unroll
        for (int d = 0; d < 6; d++) box[d] = src[d];
        dim = widest_dim(box);
      }
      const int begin = sm.seg_begin[tb][q];
      const int64_t lchild = 2 * (int64_t)s + 1;
      const int left = lchild < n ? (int)subtree_size(lchild, n) : 0;
      const int pivot = begin + left;
      const int el = sm.lst[cur][dim][pivot];
      const float split = sm.c[dim][el];
      elem_of_node[s] = sm.gid[el];
      dim_of_node[s] = (uint8_t)dim;
      sm.seg_pivot[q] = (uint16_t)pivot;
      sm.seg_dim[q] = (uint8_t)dim;
      if (!last) {
#pragma unroll
        for (int c = 0; c < 2; c++) {
          if (lchild + c >= n) continue;
          float cb[6];
#pragma unroll
          for (int d = 0; d < 6; d++) cb[d] = box[d];
          if (c == 0) cb[3 + dim] = fminf(split, box[3 + dim]);
          else cb[dim] = fmaxf(split, box[dim]);
          if (t + 2 == kLocalLevels) sm.last_dim[2 * q + c] = (uint8_t)widest_dim(cb);
          else {
            float* dst = tb ? sm.box_even[2 * q + c] : sm.box_odd[2 * q + c];
#pragma unroll
            for (int d = 0; d < 6; d++) dst[d] = cb[d];
          }
          sm.seg_begin[tb ^ 1][2 * q + c] = (uint16_t)(c == 0 ? begin : pivot + 1);
        }
      }
    }
    __syncthreads();
    if (2 * first + 1 >= n || t + 1 >= kLocalLevels) break; /* no level below */
    /* stable partition of the three lists inside every segment, one packed scan for all of them: list d keeps
     * its left count in bits [20d, 20d+10) and its right count in bits [20d+10, 20d+20) (totals <= 1023).
     * Position i of the split-axis list goes by position; in the other two lists an element goes by the list
     * order (ordered key, input index) against the pivot, which is the same decision */
    uint64_t f[kLocalItems], ex[kLocalItems];
    uint32_t sides[kLocalItems]; /* 2 bits per list, and the segment id above them */
    /* the side of every ELEMENT, from its position in the list sorted along its segment's split axis (no key is
     * compared: left of the pivot position = smaller in (coordinate, input index) order) ... */
#pragma unroll
    for (int it = 0; it < kLocalItems; it++) {
      const int i = tid * kLocalItems + it;
      if (i >= m) continue;
      const uint16_t sg = sm.segid[i];
      if (sg == kDead16) continue;
      const int piv = sm.seg_pivot[sg];
      sm.side_el[sm.lst[cur][sm.seg_dim[sg]][i]] = (uint8_t)(i < piv ? 0 : (i > piv ? 1 : 2));
    }
    __syncthreads();
    /* ... which the other two lists look up by element */
#pragma unroll
    for (int it = 0; it < kLocalItems; it++) {
      const int i = tid * kLocalItems + it;
      f[it] = 0ull; sides[it] = 0x3fu;
      if (i >= m) continue;
      const uint16_t sg = sm.segid[i]; /* position i is only ever touched by this thread */
      if (sg == kDead16) continue;
      const int piv = sm.seg_pivot[sg], dim = sm.seg_dim[sg];
      sm.segid[i] = i < piv ? (uint16_t)(2 * sg) : (i > piv ? (uint16_t)(2 * sg + 1) : kDead16);
      uint32_t pack = (uint32_t)sg << 8;
#pragma unroll
      for (int d = 0; d < 3; d++) {
        const uint32_t sd = d == dim ? (i < piv ? 0u : (i > piv ? 1u : 2u)) : (uint32_t)sm.side_el[sm.lst[cur][d][i]];
        pack |= sd << (2 * d);
        if (sd < 2u) f[it] += 1ull << (20 * d + 10 * (int)sd);
      }
      sides[it] = pack;
    }
    Scan(scan_tmp).ExclusiveSum(f, ex);
#pragma unroll
    for (int it = 0; it < kLocalItems; it++) {
      const int i = tid * kLocalItems + it;
      if (i < m) sm.scanv[i] = ex[it];
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < kLocalItems; it++) {
      const int i = tid * kLocalItems + it;
      if (sides[it] == 0x3fu) continue;
      const uint32_t sg = sides[it] >> 8;
      const int begin = sm.seg_begin[tb][sg], pivot = sm.seg_pivot[sg];
      const uint64_t rel = ex[it] - sm.scanv[begin]; /* field-wise: every field of ex >= that of the base */
#pragma unroll
      for (int d = 0; d < 3; d++) {
        const uint32_t sd = (sides[it] >> (2 * d)) & 3u;
        int dst;
        if (sd == 0) dst = begin + (int)((rel >> (20 * d)) & 0x3ffu);
        else if (sd == 1) dst = pivot + 1 + (int)((rel >> (20 * d + 10)) & 0x3ffu);
        else dst = pivot;
        sm.lst[cur ^ 1][d][dst] = sm.lst[cur][d][i];
      }
    }
    __syncthreads();
    cur ^= 1;
  }
}

/* final gather: node j <- element elem_of_node[j]; 11 threads per record so that both the random
 * 44-byte read and the streaming write are full-width */
__global__ void k_emit_nodes(int n, const int* __restrict__ elem_of_node, const uint8_t* __restrict__ dim_of_node,
                             const uint32_t* __restrict__ src_words, uint32_t* __restrict__ dst_words) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)n * 11;
  if (t >= total) return;
  const int node = (int)(t / 11);
  const int w = (int)(t - (int64_t)node * 11);
  const int e = elem_of_node[node];
  uint32_t v = src_words[(size_t)e * 11 + w];
  if (w == 10) v = (v & 0x00FFFFFFu) | ((uint32_t)dim_of_node[node] << 24);  /* split_dim is byte 43 */
  dst_words[t] = v;
}

/* query-side copies from a tree-ordered AoS array */
__global__ void k_make_soa(int n, const pm_photon44* __restrict__ ph, float4* __restrict__ nodes4,
                           float4* __restrict__ color4) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* p = reinterpret_cast<const float*>(ph) + (size_t)i * 11;
  const uint32_t last = reinterpret_cast<const uint32_t*>(p)[10];
  nodes4[i] = make_float4(p[0], p[1], p[2], __int_as_float((int)(last >> 24)));
  color4[i] = make_float4(p[6], p[7], p[8], p[9]);
}

int make_soa(pm_ctx* ctx, pm_map* m) {
  const int n = (int)m->n;
  PM_CUDA(pm::map_alloc(&m->nodes4, sizeof(float4) * (size_t)n, ctx->stream));
  PM_CUDA(pm::map_alloc(&m->color4, sizeof(float4) * (size_t)n, ctx->stream));
  k_make_soa<<<(unsigned)pm::div_up(n, 256), 256, 0, ctx->stream>>>(n, m->photons_dev, m->nodes4, m->color4);
  ctx->launches++;
  PM_CUDA(cudaGetLastError());
  return PM_OK;
}

}  // namespace

extern "C" {

/* Sorts + level passes: fills the node tables (element and split axis of every node) from the unsorted records in
 * `src`. world > 1: a sharded build — the first log2(world) levels run on everything (replicated on every rank), below
 * them this rank continues only inside subtree (2^S - 1) + rank; table entries of other ranks' subtrees stay zero,
 * and only rank 0 keeps the replicated top levels, so that a SUM over ranks yields the complete tables. */
static int build_nodes(pm_ctx* ctx, const pm_photon44* src, int n, int rank, int world, int** elem_out, uint8_t** dim_out,
                       float** bounds_dev_out) {
  cudaStream_t st = ctx->stream;
  float *cx, *cy, *cz; uint32_t *keys_in, *keys_out; int *iota;
  int* ord[2][3];
  PM_CHECK(ctx->get_scratch(pm::S_BUILD1, (size_t)n, &cx));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD2, (size_t)n, &cy));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD3, (size_t)n, &cz));
  PM_CHECK(ctx->get_scratch(pm::S_KEYS0, (size_t)n * 3, &keys_in));
  PM_CHECK(ctx->get_scratch(pm::S_KEYS1, (size_t)n, &keys_out));
  PM_CHECK(ctx->get_scratch(pm::S_VALS0, (size_t)n, &iota));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD4, (size_t)n * 3, &ord[0][0]));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD5, (size_t)n * 3, &ord[1][0]));
  for (int b = 0; b < 2; b++) { ord[b][1] = ord[b][0] + n; ord[b][2] = ord[b][0] + 2 * (size_t)n; }

  const unsigned nb256 = (unsigned)pm::div_up(n, 256);
  /* lists that carry their keys from kCarryMinPhotons on (see k_level) */
  const bool carry = n >= g_carry_min_photons;
  uint4* keys4 = nullptr;
  if (carry) PM_CHECK(ctx->get_scratch(pm::S_BUILD19, (size_t)n, &keys4));
  k_extract<<<nb256, 256, 0, st>>>(src, n, cx, cy, cz, keys_in, keys_in + n, keys_in + 2 * (size_t)n, iota, keys4);
  ctx->launches++;

  /* three stable LSD radix sorts: ties keep ascending element index */
  size_t cub_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, keys_in, keys_out, iota, ord[0][0], n, 0, 32, st);
  void* cub_tmp = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_CUB, cub_bytes, &cub_tmp));
  const size_t cub_cap = cub_bytes;
  for (int d = 0; d < 3; d++) {
    size_t tb = cub_cap;
    PM_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp, tb, keys_in + (size_t)d * n, keys_out, iota, ord[0][d], n, 0, 32, st));
    ctx->launches += 5;
  }

  /* per-level state */
  const int num_levels = 32 - __builtin_clz((unsigned)n);  /* floor(log2 n) + 1 */
  const int max_width = (n + 1) / 2 + 1;
  const int kTile = kLvThreads * (carry ? kLvItemsCarry : kLvItems);
  uint32_t *key_a[2] = {nullptr, nullptr}, *key_b[2] = {nullptr, nullptr};
  if (carry) {
    PM_CHECK(ctx->get_scratch(pm::S_BUILD17, (size_t)n * 6, &key_a[0]));
    PM_CHECK(ctx->get_scratch(pm::S_BUILD18, (size_t)n * 6, &key_b[0]));
    key_a[1] = key_a[0] + 3 * (size_t)n; key_b[1] = key_b[0] + 3 * (size_t)n;
    k_carry_init<<<(unsigned)pm::div_up(3 * (int64_t)n, 256), 256, 0, st>>>(n, ord[0][0], keys4, key_a[0], key_b[0]);
    ctx->launches++;
  }
  const unsigned ntiles = (unsigned)pm::div_up(n, kTile);
  const size_t status_words = (size_t)num_levels * ntiles + 64; /* per-level tile status words, then the tickets */
  float* box[2]; int* begin[2]; int *seg[2], *elem_of_node; uint8_t* dim_of_node; SegRec* seg_rec;
  unsigned long long* status;
  PM_CHECK(ctx->get_scratch(pm::S_BUILD6, (size_t)max_width * 6, &box[0]));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD7, (size_t)max_width * 6, &box[1]));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD8, (size_t)max_width, &begin[0]));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD9, (size_t)max_width, &begin[1]));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD11, (size_t)n * 2, &seg[0]));
  seg[1] = seg[0] + n;
  PM_CHECK(ctx->get_scratch(pm::S_BUILD12, (size_t)n, &elem_of_node));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD13, (size_t)n + 4, &dim_of_node));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD16, (size_t)max_width, &seg_rec));
  PM_CHECK(ctx->get_scratch(pm::S_BUILD14, status_words, &status));
  float* bounds_dev = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_BUILD15, (size_t)8, &bounds_dev));

  k_root<<<1, 32, 0, st>>>(ord[0][0], ord[0][1], ord[0][2], cx, cy, cz, n, box[0], begin[0], bounds_dev);
  ctx->launches++;
  PM_CUDA(cudaMemsetAsync(seg[0], 0, sizeof(int) * (size_t)n, st)); /* everything starts in subtree 0 */

  unsigned int* tickets = reinterpret_cast<unsigned int*>(status + (size_t)num_levels * ntiles);
  PM_CUDA(cudaMemsetAsync(status, 0, sizeof(unsigned long long) * status_words, st));

  const int local_from = num_levels > kLocalLevels ? num_levels - kLocalLevels : 0; /* first level finished per CTA */
  /* sharding: S replicated levels, then subtree (2^S - 1) + rank. Too small a tree is built whole by everybody. */
  int S = 0;
  while ((1 << S) < world) S++;
  const bool sharded = world > 1 && (1 << S) == world && S < local_from && (int64_t)n >= ((int64_t)2 << S);
  int pos_lo = 0, pos_hi = n;
  if (sharded) {
    PM_CUDA(cudaMemsetAsync(elem_of_node, 0, sizeof(int) * (size_t)n, st));
    PM_CUDA(cudaMemsetAsync(dim_of_node, 0, (size_t)n + 4, st));
    int64_t node = 0, b = 0; /* walk down to the rank's subtree: in-order ranges, pivot at begin + |left subtree| */
    for (int l = 0; l < S; l++) {
      const int64_t lc = 2 * node + 1;
      const int64_t left = lc < n ? subtree_size(lc, n) : 0;
      if ((rank >> (S - 1 - l)) & 1) { b = b + left + 1; node = lc + 1; } else { node = lc; }
    }
    pos_lo = (int)b;
    pos_hi = (int)(b + subtree_size(node, n));
  }
  int cur = 0;
  for (int L = 0; L < num_levels; L++) {
    const int level_first = (1 << L) - 1;
    const int level_count = (n - level_first < (1 << L)) ? n - level_first : (1 << L);
    /* this rank's subtrees of the level: all of them above the split level, a contiguous share below */
    int j0 = 0, j1 = level_count;
    if (sharded && L >= S) {
      j0 = std::min(level_count, rank << (L - S));
      j1 = std::min(level_count, (rank + 1) << (L - S));
    }
    if (L == local_from) {
      CarriedKeys ck{};
      if (carry)
        for (int d = 0; d < 3; d++) { ck.ka_in[d] = key_a[cur] + (size_t)d * n; ck.kb_in[d] = key_b[cur] + (size_t)d * n; }
      PM_CUDA(cudaFuncSetAttribute(k_finish_subtrees<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LocalSmem)));
      PM_CUDA(cudaFuncSetAttribute(k_finish_subtrees<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LocalSmem)));
      if (j1 > j0) {
        if (carry)
          k_finish_subtrees<true><<<(unsigned)(j1 - j0), kLocalThreads, sizeof(LocalSmem), st>>>(
              level_first, j0, j1, n, ord[cur][0], ord[cur][1], ord[cur][2], cx, cy, cz, box[L & 1], begin[L & 1],
              elem_of_node, dim_of_node, ck);
        else
          k_finish_subtrees<false><<<(unsigned)(j1 - j0), kLocalThreads, sizeof(LocalSmem), st>>>(
              level_first, j0, j1, n, ord[cur][0], ord[cur][1], ord[cur][2], cx, cy, cz, box[L & 1], begin[L & 1],
              elem_of_node, dim_of_node, ck);
      }
      ctx->launches++;
      break;
    }
    if (j1 > j0)
      k_split<<<(unsigned)pm::div_up(j1 - j0, 128), 128, 0, st>>>(
          level_first, j0, j1, n, ord[cur][0], ord[cur][1], ord[cur][2], cx, cy, cz, box[L & 1], begin[L & 1],
          box[(L + 1) & 1], begin[(L + 1) & 1], seg_rec, elem_of_node, dim_of_node);
    ctx->launches++;
    if (L == num_levels - 1) break;
    const bool part = sharded && L >= S;
    const int lo = part ? pos_lo : 0, hi = part ? pos_hi : n;
    const int tile0 = lo / kTile;
    const unsigned tiles = (unsigned)(pm::div_up(hi, kTile) - tile0);
    if (hi > lo) {
      CarriedKeys ck{};
      if (carry) {
        for (int d = 0; d < 3; d++) {
          ck.ka_in[d] = key_a[cur] + (size_t)d * n; ck.kb_in[d] = key_b[cur] + (size_t)d * n;
          ck.ka_out[d] = key_a[cur ^ 1] + (size_t)d * n; ck.kb_out[d] = key_b[cur ^ 1] + (size_t)d * n;
        }
        k_level<true, kLvItemsCarry><<<tiles, kLvThreads, 0, st>>>(
            n, lo, hi, tile0, level_first, seg[L & 1], seg_rec, ord[cur][0], ord[cur][1], ord[cur][2], cx, cy, cz, ord[cur ^ 1][0],
            ord[cur ^ 1][1], ord[cur ^ 1][2], seg[(L + 1) & 1], status + (size_t)L * ntiles, tickets + L, ck);
      } else {
        k_level<false, kLvItems><<<tiles, kLvThreads, 0, st>>>(
            n, lo, hi, tile0, level_first, seg[L & 1], seg_rec, ord[cur][0], ord[cur][1], ord[cur][2], cx, cy, cz, ord[cur ^ 1][0],
            ord[cur ^ 1][1], ord[cur ^ 1][2], seg[(L + 1) & 1], status + (size_t)L * ntiles, tickets + L, ck);
      }
    }
    ctx->launches++;
    cur ^= 1;
  }
  if (sharded && rank != 0) { /* the replicated top levels are rank 0's contribution to the sum */
    const size_t top = ((size_t)1 << S) - 1;
    PM_CUDA(cudaMemsetAsync(elem_of_node, 0, sizeof(int) * top, st));
    PM_CUDA(cudaMemsetAsync(dim_of_node, 0, top, st));
  }
  PM_CUDA(cudaGetLastError());
  *elem_out = elem_of_node; *dim_out = dim_of_node; *bounds_dev_out = bounds_dev;
  return PM_OK;
}

void pm_test_build_carry_from(int64_t n_min) { g_carry_min_photons = n_min > 0 ? n_min : kCarryMinPhotons; }

int pm_map_build(pm_ctx* ctx, pm_photon44* photons, int64_t n64, float* bounds_host, pm_map** out) {
  PM_REQUIRE(ctx && out, "ctx/out is NULL");
  *out = nullptr;
  if (n64 < 1) return PM_OK; /* builder_thrust.h:441: "return gracefully" */
  PM_REQUIRE(photons, "photons is NULL");
  PM_REQUIRE(n64 < (int64_t)1 << 30, "at most 2^30 photons per map");
  PM_CUDA(cudaSetDevice(ctx->device));
  const int n = (int)n64;
  cudaStream_t st = ctx->stream;
  const bool dev_in = pm::is_device_ptr(photons);

  pm_map* m = new pm_map();
  m->ctx = ctx; m->device = ctx->device; m->n = n;
  pm_photon44* tree = nullptr;  /* destination, tree order */
  if (dev_in) {
    tree = photons;
  } else {
    if (pm::map_alloc(&m->owned_photons, sizeof(pm_photon44) * (size_t)n, ctx->stream) != cudaSuccess) {
      cudaGetLastError(); delete m; pm::set_error("pm_map_build: out of device memory"); return PM_ERR_NOMEM;
    }
    tree = m->owned_photons;
  }
  m->photons_dev = tree;

#define PM_B(call) do { int r__ = (call); if (r__ != PM_OK) { pm_map_destroy(m); return r__; } } while (0)
#define PM_BC(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { pm::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); pm_map_destroy(m); return PM_ERR_CUDA; } } while (0)

  /* unsorted copy of the input records (the gather source) */
  pm_photon44* src = nullptr;
  PM_B(ctx->get_scratch(pm::S_BUILD0, (size_t)n, &src));
  PM_BC(cudaMemcpyAsync(src, photons, sizeof(pm_photon44) * (size_t)n, dev_in ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
  int* elem_of_node = nullptr; uint8_t* dim_of_node = nullptr; float* bounds_dev = nullptr;
  PM_B(build_nodes(ctx, src, n, 0, 1, &elem_of_node, &dim_of_node, &bounds_dev));

  k_emit_nodes<<<(unsigned)pm::div_up((int64_t)n * 11, 256), 256, 0, st>>>(
      n, elem_of_node, dim_of_node, reinterpret_cast<const uint32_t*>(src), reinterpret_cast<uint32_t*>(tree));
  ctx->launches++;
  PM_BC(cudaGetLastError());
  PM_B(make_soa(ctx, m));
  PM_BC(cudaMemcpyAsync(m->bounds, bounds_dev, sizeof(float) * 6, cudaMemcpyDeviceToHost, st));
  if (!dev_in) PM_BC(cudaMemcpyAsync(photons, tree, sizeof(pm_photon44) * (size_t)n, cudaMemcpyDeviceToHost, st));
  PM_BC(cudaStreamSynchronize(st));
  if (bounds_host) memcpy(bounds_host, m->bounds, sizeof(float) * 6);
  PM_B(pm::grid_build(ctx, m));
  PM_BC(cudaStreamSynchronize(st));
  *out = m;
  return PM_OK;
}

/* ---- sharded build (multi-GPU): node tables per rank, summed by the caller, then the common tail ---- */
int pm_map_build_shard(pm_ctx* ctx, const pm_photon44* photons, int64_t n64, int rank, int world, int32_t* node_elem,
                       uint8_t* node_dim, float* bounds_host) {
  PM_REQUIRE(ctx && photons && node_elem && node_dim, "ctx/photons/node_elem/node_dim is NULL");
  PM_REQUIRE(n64 >= 1 && n64 < (int64_t)1 << 30, "1 <= n < 2^30 photons per map");
  PM_REQUIRE(world >= 1 && rank >= 0 && rank < world && (world & (world - 1)) == 0, "world must be a power of two, 0 <= rank < world");
  PM_REQUIRE(pm::is_device_ptr(photons) && pm::is_device_ptr(node_elem) && pm::is_device_ptr(node_dim),
             "photons / node_elem / node_dim must be device memory");
  PM_CUDA(cudaSetDevice(ctx->device));
  const int n = (int)n64;
  cudaStream_t st = ctx->stream;
  int* elem = nullptr; uint8_t* dim = nullptr; float* bounds_dev = nullptr;
  PM_CHECK(build_nodes(ctx, photons, n, rank, world, &elem, &dim, &bounds_dev));
  /* a tree too small to shard was built whole: keep rank 0's copy only, so that the sum still is the table */
  int S = 0;
  while ((1 << S) < world) S++;
  const int num_levels = 32 - __builtin_clz((unsigned)n);
  const int local_from = num_levels > kLocalLevels ? num_levels - kLocalLevels : 0;
  const bool sharded = world > 1 && S < local_from && n64 >= ((int64_t)2 << S);
  if (world > 1 && !sharded && rank != 0) {
    PM_CUDA(cudaMemsetAsync(node_elem, 0, sizeof(int32_t) * (size_t)n, st));
    PM_CUDA(cudaMemsetAsync(node_dim, 0, (size_t)n, st));
  } else {
    PM_CUDA(cudaMemcpyAsync(node_elem, elem, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
    PM_CUDA(cudaMemcpyAsync(node_dim, dim, (size_t)n, cudaMemcpyDeviceToDevice, st));
  }
  float b[6];
  PM_CUDA(cudaMemcpyAsync(b, bounds_dev, sizeof(float) * 6, cudaMemcpyDeviceToHost, st));
  PM_CUDA(cudaStreamSynchronize(st));
  if (bounds_host) memcpy(bounds_host, b, sizeof(float) * 6);
  return PM_OK;
}

int pm_map_build_finish(pm_ctx* ctx, pm_photon44* photons, int64_t n64, const int32_t* node_elem, const uint8_t* node_dim,
                        const float* bounds_host, pm_map** out) {
  PM_REQUIRE(ctx && out, "ctx/out is NULL");
  *out = nullptr;
  PM_REQUIRE(photons && node_elem && node_dim && bounds_host, "photons/node_elem/node_dim/bounds is NULL");
  PM_REQUIRE(n64 >= 1 && n64 < (int64_t)1 << 30, "1 <= n < 2^30 photons per map");
  PM_REQUIRE(pm::is_device_ptr(photons) && pm::is_device_ptr(node_elem) && pm::is_device_ptr(node_dim),
             "photons / node_elem / node_dim must be device memory");
  PM_CUDA(cudaSetDevice(ctx->device));
  const int n = (int)n64;
  cudaStream_t st = ctx->stream;
  pm_map* m = new pm_map();
  m->ctx = ctx; m->device = ctx->device; m->n = n; m->photons_dev = photons;
  pm_photon44* src = nullptr;
  PM_B(ctx->get_scratch(pm::S_BUILD0, (size_t)n, &src));
  PM_BC(cudaMemcpyAsync(src, photons, sizeof(pm_photon44) * (size_t)n, cudaMemcpyDeviceToDevice, st));
  k_emit_nodes<<<(unsigned)pm::div_up((int64_t)n * 11, 256), 256, 0, st>>>(
      n, node_elem, node_dim, reinterpret_cast<const uint32_t*>(src), reinterpret_cast<uint32_t*>(photons));
  ctx->launches++;
  PM_BC(cudaGetLastError());
  PM_B(make_soa(ctx, m));
  memcpy(m->bounds, bounds_host, sizeof(float) * 6);
  PM_BC(cudaStreamSynchronize(st));
  PM_B(pm::grid_build(ctx, m));
  PM_BC(cudaStreamSynchronize(st));
#undef PM_B
#undef PM_BC
  *out = m;
  return PM_OK;
}

int pm_map_from_tree(pm_ctx* ctx, const pm_photon44* photons, int64_t n, pm_map** out) {
  PM_REQUIRE(ctx && out, "ctx/out is NULL");
  *out = nullptr;
  if (n < 1) return PM_OK;
  PM_REQUIRE(photons, "photons is NULL");
  PM_REQUIRE(n < (int64_t)1 << 30, "at most 2^30 photons per map");
  PM_CUDA(cudaSetDevice(ctx->device));
  pm_map* m = new pm_map();
  m->ctx = ctx; m->device = ctx->device; m->n = n;
  if (pm::is_device_ptr(photons)) {
    m->photons_dev = photons;
  } else {
    if (pm::map_alloc(&m->owned_photons, sizeof(pm_photon44) * (size_t)n, ctx->stream) != cudaSuccess) {
      cudaGetLastError(); delete m; pm::set_error("pm_map_from_tree: out of device memory"); return PM_ERR_NOMEM;
    }
    cudaMemcpyAsync(m->owned_photons, photons, sizeof(pm_photon44) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream);
    m->photons_dev = m->owned_photons;
  }
  int r = make_soa(ctx, m);
  if (r == PM_OK) r = pm::map_bounds(ctx, m);
  if (r == PM_OK) r = pm::grid_build(ctx, m);
  if (r == PM_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) { pm::set_error("pm_map_from_tree: sync failed"); r = PM_ERR_CUDA; }
  if (r != PM_OK) { pm_map_destroy(m); return r; }
  *out = m;
  return PM_OK;
}

int pm_map_build_cells(pm_ctx* ctx, const pm_photon44* photons, int64_t n, float cell_size, pm_map** out) {
  PM_REQUIRE(ctx && out, "ctx/out is NULL");
  *out = nullptr;
  if (n < 1) return PM_OK;
  PM_REQUIRE(photons, "photons is NULL");
  PM_REQUIRE(n < (int64_t)1 << 30, "at most 2^30 photons per map");
  PM_REQUIRE(cell_size > 0.f && cell_size < INFINITY, "cell_size must be positive and finite");
  PM_CUDA(cudaSetDevice(ctx->device));
  pm_map* m = new pm_map();
  m->ctx = ctx; m->device = ctx->device; m->n = n; m->cells_only = true;
  if (pm::is_device_ptr(photons)) {
    m->photons_dev = photons;
  } else {
    if (pm::map_alloc(&m->owned_photons, sizeof(pm_photon44) * (size_t)n, ctx->stream) != cudaSuccess) {
      cudaGetLastError(); delete m; pm::set_error("pm_map_build_cells: out of device memory"); return PM_ERR_NOMEM;
    }
    cudaMemcpyAsync(m->owned_photons, photons, sizeof(pm_photon44) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream);
    m->photons_dev = m->owned_photons;
  }
  int r = make_soa(ctx, m); /* input order: node id = input index; the split_dim byte is not used */
  if (r == PM_OK) r = pm::map_bounds(ctx, m);
  if (r == PM_OK) r = pm::grid_build(ctx, m, cell_size);
  if (r == PM_OK) r = pm::cells_reorder(ctx, m);
  if (r == PM_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) { pm::set_error("pm_map_build_cells: sync failed"); r = PM_ERR_CUDA; }
  if (r != PM_OK) { pm_map_destroy(m); return r; }
  *out = m;
  return PM_OK;
}

void pm_map_destroy(pm_map* m) {
  if (!m) return;
  cudaSetDevice(m->device);
  cudaDeviceSynchronize();
  pm::map_free(m->owned_photons);
  pm::map_free(m->nodes4);
  pm::map_free(m->color4);
  pm::map_free(m->grid_pts);
  pm::map_free(m->cell_start);
  pm::map_free(m->cell_start_t);
  cudaStreamSynchronize((cudaStream_t)0); /* the frees have happened: the next build may take the memory on its own stream */
  delete m;
}

int64_t pm_map_size(const pm_map* m) { return m ? m->n : 0; }
int pm_map_query_path(const pm_map* m) { return !m || !m->grid_pts ? 0 : (m->fast_path ? 2 : 1); }

}  // extern "C"
