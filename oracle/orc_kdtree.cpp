/* orc_kdtree.cpp — CPU oracle for stage 2 (photon-map build) and the kNN / gather part of
 * stage 3. TEST INFRASTRUCTURE ONLY (see orc_common.h).
 *
 * Parity status: PINNED for the tree (tests compare orc_kd_build with cukd::buildTree_host
 * compiled from the reference into oracle/_ref, on unique-coordinate inputs as the reference's
 * own testBuildersSameResult.cu does) and for kNN on the GPU box (tests compare against the
 * reference's cukd::stackBased::knn kernel compiled from the reference into oracle/_ref).
 */
#include <algorithm>
#include <vector>
#include <numeric>
#include <limits>
#include <omp.h>
#include "orc_common.h"

using namespace orc;

namespace {

/* number of nodes of an n-node left-balanced (level-order, children 2i+1 / 2i+2) tree that lie
 * in the subtree rooted at `root`. Same quantity as ArbitraryBinaryTree::numNodesInSubtree
 * (cukd/helpers.h:162-178), computed here by summing the level ranges directly. */
int64_t subtree_size(int64_t root, int64_t n) {
  int64_t total = 0;
  int64_t first = root, width = 1;
  while (first < n) {
    total += std::min(n, first + width) - first;
    first = 2 * first + 1;
    width *= 2;
  }
  return total;
}

/* box_t::widestDimension (cukd/box.h:112-125): ties go to the HIGHEST axis */
int widest_dim(const float lo[3], const float hi[3]) {
  int best = 0; float wbest = 0.f;
  for (int d = 0; d < 3; d++) {
    const float w = hi[d] - lo[d];
    if (w < wbest) continue;
    wbest = w; best = d;
  }
  return best;
}

struct Builder {
  const pm_photon44* in;
  pm_photon44* out;
  int64_t n;
  /* recursive statement of the level-synchronous build (builder_thrust.h:259-314):
   * node `node` owns idx[b..e); its split dim is the widest axis of its DOMAIN box (world box
   * clipped by the ancestors' planes, builder_common.h:179-217); the pivot is the element of rank
   * |left subtree| in ascending (coord[dim], input index) order. */
  void build(int32_t* idx, int64_t b, int64_t e, int64_t node, const float lo[3], const float hi[3]) {
    if (e <= b) return;
    const int dim = widest_dim(lo, hi);
    const int64_t left = (2 * node + 1 < n) ? subtree_size(2 * node + 1, n) : 0;
    auto less = [&](int32_t a, int32_t c) {
      const float ca = in[a].pos[dim], cc = in[c].pos[dim];
      return ca < cc || (ca == cc && a < c);
    };
    std::nth_element(idx + b, idx + b + left, idx + e, less);
    const int32_t piv = idx[b + left];
    out[node] = in[piv];
    out[node].split_dim = (uint8_t)dim;
    const float split = in[piv].pos[dim];
    /* left child: upper[dim] = min(split, upper[dim]); right: lower[dim] = max(split, lower[dim])
     * (findBounds, builder_common.h:203-212) */
    float lhi[3] = {hi[0], hi[1], hi[2]}; lhi[dim] = std::min(split, hi[dim]);
    float rlo[3] = {lo[0], lo[1], lo[2]}; rlo[dim] = std::max(split, lo[dim]);
    float llo[3] = {lo[0], lo[1], lo[2]};
    float rhi[3] = {hi[0], hi[1], hi[2]};
    const int64_t lb = b, le = b + left, rb = b + left + 1, re = e;
    if (le - lb > 4096) {
#pragma omp task firstprivate(lb, le, node, llo, lhi)
      build(idx, lb, le, 2 * node + 1, llo, lhi);
      build(idx, rb, re, 2 * node + 2, rlo, rhi);
#pragma omp taskwait
    } else {
      build(idx, lb, le, 2 * node + 1, llo, lhi);
      build(idx, rb, re, 2 * node + 2, rlo, rhi);
    }
  }
};

/* ---- HeapCandidateList<k> (cukd/knn.h:267-364) ---------------------------------------- */
struct Heap {
  uint64_t* e; int k;
  static uint64_t encode(float d, int id) { return ((uint64_t)f2u(d) << 32) | (uint32_t)id; }
  void init(float cutoff) { for (int i = 0; i < k; i++) e[i] = encode(cutoff * cutoff, -1); }
  float maxRadius2() const { return u2f((uint32_t)(e[0] >> 32)); }
  void push(float dist, int id) {
    const uint64_t v = encode(dist, id);
    if (v >= e[0]) return;
    int pos = 0;
    while (true) {
      uint64_t largestChildValue = ~0ull;
      const int first = 2 * pos + 1;
      int largest = k;
      if (first < k) { largest = first; largestChildValue = e[first]; }
      const int second = first + 1;
      if (second < k && e[second] > largestChildValue) { largest = second; largestChildValue = e[second]; }
      if (largest == k || largestChildValue < v) { e[pos] = v; break; }
      e[pos] = largestChildValue; pos = largest;
    }
  }
};

/* canonical squared distance, d = node - query (cukd/math.h:208-211). The contraction order is
 * the one nvcc 12.9 emits for cukd's dot(d,d) = (dx*dx + dy*dy) + dz*dz in the reference kernel
 * (SASS of oracle/_ref: FMUL t,dy,dy ; FFMA t,dx,dx,t ; FFMA d2,dz,dz,t). */
inline float sqr_dist(const float* p, const float* q) {
  const float dx = p[0] - q[0], dy = p[1] - q[1], dz = p[2] - q[2];
  return fmaf(dz, dz, fmaf(dx, dx, dy * dy));
}

/* traverse_default (cukd/traverse-default-stack-based.h:24-90) with explicit split dims */
float knn_traverse(const pm_photon44* nodes, int n, const float q[3], int k, float cutoff, uint64_t* heap_mem,
                   int64_t* visited) {
  Heap heap{heap_mem, k};
  heap.init(cutoff);
  float cull = heap.maxRadius2();
  struct Entry { int node; float sqrDist; };
  Entry stack[64];
  int sp = 0;
  int curr = 0;
  int64_t nvis = 0;
  while (true) {
    while (curr < n) {
      const pm_photon44& nd = nodes[curr];
      const int dim = nd.split_dim;
      const float d2 = sqr_dist(nd.pos, q);
      nvis++;
      heap.push(d2, curr);
      cull = heap.maxRadius2();
      const float nc = nd.pos[dim], qc = q[dim];
      const bool leftIsClose = qc < nc;
      const int l = 2 * curr + 1, r = l + 1;
      const int closeChild = leftIsClose ? l : r;
      const int farChild = leftIsClose ? r : l;
      const float dp = qc - nc;
      const float sqrDistToPlane = dp * dp;
      if (sqrDistToPlane < cull && farChild < n) { stack[sp].node = farChild; stack[sp].sqrDist = sqrDistToPlane; sp++; }
      curr = closeChild;
    }
    bool found = false;
    while (sp > 0) {
      --sp;
      if (stack[sp].sqrDist >= cull) continue;
      curr = stack[sp].node; found = true; break;
    }
    if (!found) break;
  }
  if (visited) *visited = nvis;
  return heap.maxRadius2();
}

}  // namespace

namespace orc {
/* gatherPhotons (ray-tracer/cuda/shading.h:93-121), slots visited in heap-array order.
 * Also used by orc_render.cpp. */
void gather_one(const pm_photon44* nodes, int n, const float hit[3], float brdf, int k, float cutoff,
                uint64_t* heap_mem, float out[3]) {
  const float r2 = knn_traverse(nodes, n, hit, k, cutoff, heap_mem, nullptr);
  vec3 flux = v3(0.f, 0.f, 0.f);
  const vec3 hp = v3(hit);
  for (int p = 0; p < k; p++) {
    const int id = (int)(uint32_t)heap_mem[p];
    if (id < 0 || id > n) continue;                 /* sic: '>' as in shading.h:105 */
    const pm_photon44& ph = nodes[id];
    const float power = ph.power;
    const float dist = norm(v3(ph.pos) - hp);
    const float w = 1 - (dist / sqrtf(r2) * 1.1f);
    flux = flux + ((brdf * power) * w) * v3(ph.color);
  }
  const float denom = (1 - (2.f / 3.f) * (1.f / 1.1f)) * 2 * ORC_PI * r2;
  const vec3 res = flux / denom;
  out[0] = res.x; out[1] = res.y; out[2] = res.z;
}
}  // namespace orc

extern "C" {

int64_t orc_subtree_size(int64_t root, int64_t n) { return subtree_size(root, n); }

/* stage 2 oracle: in-place reorder into left-balanced kd-tree order + split dims + bounds */
void orc_kd_build(pm_photon44* photons, int64_t n, float bounds[6]) {
  if (n < 1) return;
  float lo[3] = {std::numeric_limits<float>::infinity(), std::numeric_limits<float>::infinity(),
                 std::numeric_limits<float>::infinity()};
  float hi[3] = {-lo[0], -lo[1], -lo[2]};
  for (int64_t i = 0; i < n; i++)
    for (int d = 0; d < 3; d++) {
      lo[d] = std::min(lo[d], photons[i].pos[d]);
      hi[d] = std::max(hi[d], photons[i].pos[d]);
    }
  if (bounds) { for (int d = 0; d < 3; d++) { bounds[d] = lo[d]; bounds[3 + d] = hi[d]; } }
  std::vector<pm_photon44> in(photons, photons + n);
  std::vector<int32_t> idx(n);
  std::iota(idx.begin(), idx.end(), 0);
  Builder b{in.data(), photons, n};
#pragma omp parallel
#pragma omp single
  b.build(idx.data(), 0, n, 0, lo, hi);
}

/* literal traversal + heap; heap_out receives the k raw heap entries (heap order) */
float orc_knn_traverse(const pm_photon44* nodes, int n, const float q[3], int k, float cutoff,
                       uint64_t* heap_out, int64_t* visited) {
  return knn_traverse(nodes, n, q, k, cutoff, heap_out, visited);
}

/* brute force: k smallest (dist² bits, id) with dist² < cutoff², ascending; unfilled = (cutoff², -1).
 * *tie = 1 when the k-th and (k+1)-th candidates have bit-equal dist² (set is then not unique
 * under pruned traversal). Returns r² as knn() would. */
float orc_knn_brute(const pm_photon44* nodes, int n, const float q[3], int k, float cutoff,
                    uint64_t* out_sorted, int* tie) {
  const uint64_t sentinel = Heap::encode(cutoff * cutoff, -1);
  std::vector<uint64_t> best;
  best.reserve(n);
  for (int i = 0; i < n; i++) {
    const uint64_t v = Heap::encode(sqr_dist(nodes[i].pos, q), i);
    if (v < sentinel) best.push_back(v);
  }
  const size_t kk = std::min<size_t>(k + 1, best.size());
  std::partial_sort(best.begin(), best.begin() + kk, best.end());
  if (tie) *tie = (best.size() > (size_t)k && (best[k] >> 32) == (best[k - 1] >> 32)) ? 1 : 0;
  for (int i = 0; i < k; i++) out_sorted[i] = (size_t)i < best.size() ? best[i] : sentinel;
  return u2f((uint32_t)(out_sorted[k - 1] >> 32));
}

/* batch kNN through the literal traversal, slots sorted ascending; OpenMP over queries.
 * This is the timed CPU baseline of the gather microbench. */
void orc_knn_batch(const pm_photon44* nodes, int n, const float* queries, int64_t nq, int k, float cutoff,
                   int32_t* out_ids, float* out_dist2, float* out_r2, int nthreads) {
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel num_threads(nthreads)
  {
    std::vector<uint64_t> heap(k);
#pragma omp for schedule(dynamic, 256)
    for (int64_t i = 0; i < nq; i++) {
      const float r2 = knn_traverse(nodes, n, queries + 3 * i, k, cutoff, heap.data(), nullptr);
      std::sort(heap.begin(), heap.end());
      for (int j = 0; j < k; j++) {
        if (out_ids) out_ids[i * k + j] = (int32_t)(uint32_t)heap[j];
        if (out_dist2) out_dist2[i * k + j] = u2f((uint32_t)(heap[j] >> 32));
      }
      if (out_r2) out_r2[i] = r2;
    }
  }
}

void orc_gather_batch(const pm_photon44* nodes, int n, const float* hit, const float* brdf, int64_t nq, int k,
                      float cutoff, float* out_rgb, int nthreads) {
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel num_threads(nthreads)
  {
    std::vector<uint64_t> heap(k);
#pragma omp for schedule(dynamic, 256)
    for (int64_t i = 0; i < nq; i++) {
      if (n < 1) { out_rgb[3 * i] = out_rgb[3 * i + 1] = out_rgb[3 * i + 2] = 0.f; continue; }
      gather_one(nodes, n, hit + 3 * i, brdf[i], k, cutoff, heap.data(), out_rgb + 3 * i);
    }
  }
}

/* checkRec-style validity predicate (cudaKDTree/testing/floatN-knn-and-fcp.cu:337-368):
 * every node's coordinate in its split dim bounds its subtrees. Returns 1 if valid. */
static bool check_rec(const pm_photon44* nodes, int n, int node, float lo[3], float hi[3]) {
  if (node >= n) return true;
  const pm_photon44& nd = nodes[node];
  for (int d = 0; d < 3; d++)
    if (nd.pos[d] < lo[d] || nd.pos[d] > hi[d]) return false;
  const int dim = nd.split_dim;
  if (dim > 2) return false;
  const float s = nd.pos[dim];
  float lhi[3] = {hi[0], hi[1], hi[2]}; lhi[dim] = s;
  float rlo[3] = {lo[0], lo[1], lo[2]}; rlo[dim] = s;
  return check_rec(nodes, n, 2 * node + 1, lo, lhi) && check_rec(nodes, n, 2 * node + 2, rlo, hi);
}
int orc_kd_check(const pm_photon44* nodes, int n) {
  float lo[3] = {-INFINITY, -INFINITY, -INFINITY}, hi[3] = {INFINITY, INFINITY, INFINITY};
  return check_rec(nodes, n, 0, lo, hi) ? 1 : 0;
}

/* Fixed-radius gather (the progressive pass of BASELINE config 5; no reference code exists, this is the
 * definition libpmhot implements): flux = brdf * sum(power * colour) over nodes with dist² < radius²,
 * dist² as in the kNN contract; brute force, float accumulation in node order. */
void orc_gather_radius(const pm_photon44* nodes, int n, const float* hit, const float* brdf, const float* radius2,
                       int64_t nq, float* out_flux, int32_t* out_count) {
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t i = 0; i < nq; i++) {
    float s[3] = {0.f, 0.f, 0.f};
    int cnt = 0;
    if (brdf[i] > 0.f && radius2[i] > 0.f) {
      for (int j = 0; j < n; j++) {
        if (sqr_dist(nodes[j].pos, hit + 3 * i) < radius2[i]) {
          for (int c = 0; c < 3; c++) s[c] = s[c] + nodes[j].power * nodes[j].color[c];
          cnt++;
        }
      }
    }
    const float f = brdf[i] > 0.f ? brdf[i] : 0.f;
    for (int c = 0; c < 3; c++) out_flux[3 * i + c] = f * s[c];
    out_count[i] = cnt;
  }
}

int orc_omp_threads(void) { return omp_get_max_threads(); }
/* launchers (torchrun) export OMP_NUM_THREADS=1; the timed CPU arm uses every host thread */
void orc_omp_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }

}  // extern "C"
