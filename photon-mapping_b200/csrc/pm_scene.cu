/* pm_scene.cu — scene upload and LBVH construction (replaces loadGeometry + owlGroupBuildAccel,
 * common/src/world.cpp:3-58) and the two ray-cast test hooks.
 *
 * Build, all on the device:
 *   1. k_tri_setup   per triangle: A, B-A, C-A records, bbox, 30-bit Morton code of the bbox centre
 *                    inside the scene AABB; sort key = morton << 32 | triangle id (unique)
 *   2. cub radix sort of the 64-bit keys
 *   3. k_hierarchy   Karras 2012: one thread per inner node finds its key range and split
 *   4. k_refit       leaves walk up; the second arrival at a node merges the child boxes
 *   5. k_pack        inner node -> 64-byte record holding both (padded) child boxes
 * Scenes of at most kSahMaxTris triangles (the reference's Cornell box has 58) get a full-sweep SAH hierarchy built
 * on the host instead of steps 1-4 of the Morton build: same node records, same one-triangle leaves, fewer box and
 * triangle tests per ray. The hierarchy only decides which triangles are skipped, never the hit that is returned.
 */
#include <cub/cub.cuh>
#include <algorithm>
#include <cfloat>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "pm_scene.cuh"

namespace {

using namespace pmd;

struct Aabb { float lo[3], hi[3]; };

__device__ __forceinline__ uint32_t expand10(uint32_t v) {
  v &= 0x3FFu;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}

__global__ void k_tri_setup(const float* __restrict__ verts, const int* __restrict__ tris, const int* __restrict__ tri_mesh,
                            int nt, Aabb scene, float4* __restrict__ rec /*3 per tri, input order*/,
                            float* __restrict__ box /*6 per tri*/, uint64_t* __restrict__ keys) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nt) return;
  const int a = tris[3 * i], b = tris[3 * i + 1], c = tris[3 * i + 2];
  const vec3 A = v3(verts + 3 * (size_t)a), B = v3(verts + 3 * (size_t)b), C = v3(verts + 3 * (size_t)c);
  const vec3 e1 = B - A, e2 = C - A;
  rec[3 * (size_t)i + 0] = make_float4(A.x, A.y, A.z, __int_as_float(i));
  rec[3 * (size_t)i + 1] = make_float4(e1.x, e1.y, e1.z, __int_as_float(tri_mesh ? tri_mesh[i] : 0));
  rec[3 * (size_t)i + 2] = make_float4(e2.x, e2.y, e2.z, 0.f);
  const float lo[3] = {fminf(A.x, fminf(B.x, C.x)), fminf(A.y, fminf(B.y, C.y)), fminf(A.z, fminf(B.z, C.z))};
  const float hi[3] = {fmaxf(A.x, fmaxf(B.x, C.x)), fmaxf(A.y, fmaxf(B.y, C.y)), fmaxf(A.z, fmaxf(B.z, C.z))};
  uint32_t code = 0;
#pragma unroll
  for (int d = 0; d < 3; d++) {
    box[6 * (size_t)i + d] = lo[d];
    box[6 * (size_t)i + 3 + d] = hi[d];
    const float w = scene.hi[d] - scene.lo[d];
    float t = w > 0.f ? (0.5f * (lo[d] + hi[d]) - scene.lo[d]) / w * 1024.f : 0.f;
    t = fminf(fmaxf(t, 0.f), 1023.f);
    code |= expand10((uint32_t)t) << d;
  }
  keys[i] = ((uint64_t)code << 32) | (uint32_t)i;
}

__device__ __forceinline__ int delta(const uint64_t* __restrict__ k, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  return __clzll(k[i] ^ k[j]);
}

/* inner node i in [0, n-1): children encoded as inner index >= 0 or ~leaf; parent links for the refit */
__global__ void k_hierarchy(const uint64_t* __restrict__ keys, int n, int* __restrict__ child0, int* __restrict__ child1,
                            int* __restrict__ parent /*[2n-1]: inner 0..n-2, leaf n-1+j*/) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
  const int dmin = delta(keys, n, i, i - d);
  int lmax = 2;
  while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int t = lmax / 2; t >= 1; t /= 2)
    if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = delta(keys, n, i, j);
  int s = 0;
  int t = l;
  do {
    t = (t + 1) / 2;
    if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  const int gamma = i + s * d + min(d, 0);
  const int lo = min(i, j), hi = max(i, j);
  const int c0 = (lo == gamma) ? ~gamma : gamma;
  const int c1 = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
  child0[i] = c0; child1[i] = c1;
  parent[c0 < 0 ? n - 1 + ~c0 : c0] = i;
  parent[c1 < 0 ? n - 1 + ~c1 : c1] = i;
  if (i == 0) parent[0] = -1;
}

__global__ void k_refit(int n, const uint64_t* __restrict__ keys, const float* __restrict__ tri_box,
                        const int* __restrict__ child0, const int* __restrict__ child1, const int* __restrict__ parent,
                        float* __restrict__ node_box /*6 per node, 2n-1*/, int* __restrict__ visits) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const int tri = (int)(uint32_t)keys[j];
  float b[6];
#pragma unroll
  for (int d = 0; d < 6; d++) { b[d] = tri_box[6 * (size_t)tri + d]; node_box[6 * (size_t)(n - 1 + j) + d] = b[d]; }
  int node = parent[n - 1 + j];
  while (node >= 0) {
    __threadfence();
    if (atomicAdd(&visits[node], 1) == 0) return; /* first arrival: the sibling finishes the node */
    const int c0 = child0[node], c1 = child1[node];
    const size_t o0 = 6 * (size_t)(c0 < 0 ? n - 1 + ~c0 : c0), o1 = 6 * (size_t)(c1 < 0 ? n - 1 + ~c1 : c1);
#pragma unroll
    for (int d = 0; d < 3; d++) {
      b[d] = fminf(__ldcg(node_box + o0 + d), __ldcg(node_box + o1 + d));
      b[3 + d] = fmaxf(__ldcg(node_box + o0 + 3 + d), __ldcg(node_box + o1 + 3 + d));
    }
#pragma unroll
    for (int d = 0; d < 6; d++) node_box[6 * (size_t)node + d] = b[d];
    node = parent[node];
  }
}

__global__ void k_pack(int n, float pad, const int* __restrict__ child0, const int* __restrict__ child1,
                       const float* __restrict__ node_box, float4* __restrict__ nodes) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int c0 = child0[i], c1 = child1[i];
  const float* b0 = node_box + 6 * (size_t)(c0 < 0 ? n - 1 + ~c0 : c0);
  const float* b1 = node_box + 6 * (size_t)(c1 < 0 ? n - 1 + ~c1 : c1);
  nodes[4 * (size_t)i + 0] = make_float4(b0[0] - pad, b0[3] + pad, b0[1] - pad, b0[4] + pad);
  nodes[4 * (size_t)i + 1] = make_float4(b1[0] - pad, b1[3] + pad, b1[1] - pad, b1[4] + pad);
  nodes[4 * (size_t)i + 2] = make_float4(b0[2] - pad, b0[5] + pad, b1[2] - pad, b1[5] + pad);
  nodes[4 * (size_t)i + 3] = make_float4(__int_as_float(c0), __int_as_float(c1), 0.f, 0.f);
}

/* triangle records into leaf (sorted) order */
__global__ void k_permute_tris(int n, const uint64_t* __restrict__ keys, const float4* __restrict__ rec,
                               float4* __restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 3 * n) return;
  const int j = t / 3, w = t - 3 * j;
  out[t] = rec[3 * (size_t)(uint32_t)keys[j] + w];
}

/* single triangle: one inner node whose second child does not exist */
__global__ void k_single(const float* __restrict__ box, float pad, float4* __restrict__ nodes) {
  nodes[0] = make_float4(box[0] - pad, box[3] + pad, box[1] - pad, box[4] + pad);
  nodes[1] = make_float4(0.f, 0.f, 0.f, 0.f);
  nodes[2] = make_float4(box[2] - pad, box[5] + pad, 0.f, 0.f);
  nodes[3] = make_float4(__int_as_float(~0), __int_as_float(kNoChild), 0.f, 0.f);
}

__global__ void k_trace_closest(SceneView sc, const pm_ray* __restrict__ rays, int64_t n, pm_hit* __restrict__ hits) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const pm_ray r = rays[i];
  const Hit h = trace<false>(sc, v3(r.org), v3(r.dir), r.tmin, r.tmax);
  hits[i].prim = h.prim;
  hits[i].t = h.prim >= 0 ? h.t : r.tmax;
}

__global__ void k_trace_occluded(SceneView sc, const pm_ray* __restrict__ rays, int64_t n, uint8_t* __restrict__ occ) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const pm_ray r = rays[i];
  occ[i] = trace<true>(sc, v3(r.org), v3(r.dir), r.tmin, r.tmax).prim >= 0 ? 1 : 0;
}

/* ---- host-side SAH build for small scenes ---- */
constexpr int kSahMaxTris = 4096;
constexpr int kSahMaxDepth = 44; /* below this depth: median splits (<= 12 more levels), so the 64-entry ray stack holds */

struct SahBuilder {
  const float* tb;            /* 6 floats per triangle: lo, hi */
  std::vector<int> order;     /* leaf slot -> triangle */
  std::vector<int> c0, c1;    /* inner node -> child (inner index or ~slot) */
  std::vector<float> cbox;    /* inner node -> the two child boxes, 12 floats */
  std::vector<float> lcost;   /* sweep scratch */

  static float half_area(const float* b) {
    const float dx = b[3] - b[0], dy = b[4] - b[1], dz = b[5] - b[2];
    return dx * dy + dy * dz + dz * dx;
  }
  static void grow(float* b, const float* t) {
    for (int d = 0; d < 3; d++) { b[d] = std::min(b[d], t[d]); b[3 + d] = std::max(b[3 + d], t[3 + d]); }
  }
  static void empty(float* b) { for (int d = 0; d < 3; d++) { b[d] = FLT_MAX; b[3 + d] = -FLT_MAX; } }

  void sort_axis(int* ids, int cnt, int axis) const {
    const float* t = tb;
    std::sort(ids, ids + cnt, [t, axis](int a, int b) {
      const float ca = t[6 * (size_t)a + axis] + t[6 * (size_t)a + 3 + axis], cb = t[6 * (size_t)b + axis] + t[6 * (size_t)b + 3 + axis];
      return ca < cb || (ca == cb && a < b);
    });
  }

  /* builds the subtree over ids[0, cnt); returns its encoded reference and its box */
  int build(int* ids, int cnt, int depth, float* box) {
    if (cnt == 1) {
      const int slot = (int)order.size();
      order.push_back(ids[0]);
      for (int d = 0; d < 6; d++) box[d] = tb[6 * (size_t)ids[0] + d];
      return ~slot;
    }
    const int me = (int)c0.size();
    c0.push_back(0); c1.push_back(0); cbox.resize(cbox.size() + 12);
    int best_axis = 0, best_i = cnt / 2;
    if (depth < kSahMaxDepth) {
      float best = FLT_MAX;
      lcost.resize((size_t)cnt);
      for (int axis = 0; axis < 3; axis++) {
        sort_axis(ids, cnt, axis);
        float b[6];
        empty(b);
        for (int i = 0; i < cnt - 1; i++) { grow(b, tb + 6 * (size_t)ids[i]); lcost[i + 1] = half_area(b) * (float)(i + 1); }
        empty(b);
        for (int i = cnt - 1; i >= 1; i--) {
          grow(b, tb + 6 * (size_t)ids[i]);
          const float cost = lcost[i] + half_area(b) * (float)(cnt - i);
          if (cost < best) { best = cost; best_axis = axis; best_i = i; }
        }
      }
    }
    sort_axis(ids, cnt, best_axis);
    float b0[6], b1[6];
    const int r0 = build(ids, best_i, depth + 1, b0);
    const int r1 = build(ids + best_i, cnt - best_i, depth + 1, b1);
    c0[me] = r0; c1[me] = r1;
    for (int d = 0; d < 6; d++) { cbox[12 * (size_t)me + d] = b0[d]; cbox[12 * (size_t)me + 6 + d] = b1[d]; }
    for (int d = 0; d < 3; d++) { box[d] = std::min(b0[d], b1[d]); box[3 + d] = std::max(b0[3 + d], b1[3 + d]); }
    return me;
  }
};

}  // namespace

extern "C" {

int pm_scene_create(pm_ctx* ctx, const float* verts_host, int nv, const int32_t* tris_host, int nt,
                    const int32_t* tri_mesh_host, const pm_material* mats_host, int n_mesh, pm_scene** out) {
  PM_REQUIRE(ctx && out, "ctx/out is NULL");
  *out = nullptr;
  PM_REQUIRE(nv >= 0 && nt >= 0 && n_mesh >= 0, "negative size");
  PM_REQUIRE(nt == 0 || (verts_host && tris_host && nv > 0), "verts/tris is NULL");
  for (int64_t i = 0; i < 3 * (int64_t)nt; i++)
    PM_REQUIRE(tris_host[i] >= 0 && tris_host[i] < nv, "triangle vertex index out of range");
  if (tri_mesh_host)
    for (int i = 0; i < nt; i++) PM_REQUIRE(tri_mesh_host[i] >= 0 && tri_mesh_host[i] < n_mesh, "mesh id out of range");
  PM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  pm_scene* s = new pm_scene();
  s->ctx = ctx; s->device = ctx->device; s->nt = nt; s->nv = nv; s->n_mesh = n_mesh;
#define PM_S(call) do { int r__ = (call); if (r__ != PM_OK) { pm_scene_destroy(s); return r__; } } while (0)
#define PM_SC(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { pm::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); pm_scene_destroy(s); return PM_ERR_CUDA; } } while (0)
  if (n_mesh > 0 && mats_host) {
    PM_SC(cudaMalloc(&s->mats, sizeof(pm_material) * (size_t)n_mesh));
    PM_SC(cudaMemcpyAsync(s->mats, mats_host, sizeof(pm_material) * (size_t)n_mesh, cudaMemcpyHostToDevice, st));
  }
  if (nt == 0) { PM_SC(cudaStreamSynchronize(st)); *out = s; return PM_OK; }

  /* scene AABB over the referenced vertices (host; the arrays arrive on the host anyway) */
  Aabb sb;
  for (int d = 0; d < 3; d++) { sb.lo[d] = FLT_MAX; sb.hi[d] = -FLT_MAX; }
  for (int64_t i = 0; i < 3 * (int64_t)nt; i++)
    for (int d = 0; d < 3; d++) {
      const float v = verts_host[3 * (size_t)tris_host[i] + d];
      if (v < sb.lo[d]) sb.lo[d] = v;
      if (v > sb.hi[d]) sb.hi[d] = v;
    }
  float ext = 0.f;
  for (int d = 0; d < 3; d++) {
    ext = fmaxf(ext, sb.hi[d] - sb.lo[d]);
    ext = fmaxf(ext, fmaxf(fabsf(sb.lo[d]), fabsf(sb.hi[d])));
  }
  const float pad = ext * 2e-5f + 1e-30f;

  float *d_verts, *tri_box, *node_box; int *d_tris, *d_mesh = nullptr, *child0, *child1, *parent, *visits;
  float4* rec; uint64_t *keys_in, *keys_out;
  PM_S(ctx->get_scratch(pm::S_BUILD0, (size_t)nv * 3, &d_verts));
  PM_S(ctx->get_scratch(pm::S_BUILD1, (size_t)nt * 3, &d_tris));
  PM_S(ctx->get_scratch(pm::S_BUILD2, (size_t)nt, &d_mesh));
  PM_S(ctx->get_scratch(pm::S_BUILD3, (size_t)nt * 3, &rec));
  PM_S(ctx->get_scratch(pm::S_BUILD4, (size_t)nt * 6, &tri_box));
  PM_S(ctx->get_scratch(pm::S_BUILD5, (size_t)nt, &keys_in));
  PM_S(ctx->get_scratch(pm::S_BUILD6, (size_t)nt, &keys_out));
  PM_S(ctx->get_scratch(pm::S_BUILD7, (size_t)nt * 2 * 6, &node_box));
  PM_S(ctx->get_scratch(pm::S_BUILD8, (size_t)nt, &child0));
  PM_S(ctx->get_scratch(pm::S_BUILD9, (size_t)nt, &child1));
  PM_S(ctx->get_scratch(pm::S_BUILD10, (size_t)nt * 2, &parent));
  PM_S(ctx->get_scratch(pm::S_BUILD11, (size_t)nt, &visits));
  PM_SC(cudaMemcpyAsync(d_verts, verts_host, sizeof(float) * 3 * (size_t)nv, cudaMemcpyHostToDevice, st));
  PM_SC(cudaMemcpyAsync(d_tris, tris_host, sizeof(int) * 3 * (size_t)nt, cudaMemcpyHostToDevice, st));
  if (tri_mesh_host) PM_SC(cudaMemcpyAsync(d_mesh, tri_mesh_host, sizeof(int) * (size_t)nt, cudaMemcpyHostToDevice, st));
  PM_SC(cudaMalloc(&s->tris, sizeof(float4) * 3 * (size_t)nt));
  PM_SC(cudaMalloc(&s->nodes, sizeof(float4) * 4 * (size_t)(nt > 1 ? nt - 1 : 1)));

  const unsigned nb = (unsigned)pm::div_up(nt, 256);
  k_tri_setup<<<nb, 256, 0, st>>>(d_verts, d_tris, tri_mesh_host ? d_mesh : nullptr, nt, sb, rec, tri_box, keys_in);
  ctx->launches++;
  if (nt == 1) {
    PM_SC(cudaMemcpyAsync(s->tris, rec, sizeof(float4) * 3, cudaMemcpyDeviceToDevice, st));
    k_single<<<1, 1, 0, st>>>(tri_box, pad, s->nodes);
    ctx->launches++;
  } else if (nt <= kSahMaxTris) {
    std::vector<float> tbox((size_t)nt * 6);
    for (int i = 0; i < nt; i++)
      for (int d = 0; d < 3; d++) {
        const float a = verts_host[3 * (size_t)tris_host[3 * i] + d], b = verts_host[3 * (size_t)tris_host[3 * i + 1] + d],
                    c = verts_host[3 * (size_t)tris_host[3 * i + 2] + d];
        tbox[6 * (size_t)i + d] = fminf(a, fminf(b, c));
        tbox[6 * (size_t)i + 3 + d] = fmaxf(a, fmaxf(b, c));
      }
    SahBuilder sah;
    sah.tb = tbox.data();
    std::vector<int> ids((size_t)nt);
    for (int i = 0; i < nt; i++) ids[i] = i;
    float root_box[6];
    sah.build(ids.data(), nt, 0, root_box);
    std::vector<float4> hnodes((size_t)(nt - 1) * 4);
    for (int i = 0; i < nt - 1; i++) {
      const float* b0 = &sah.cbox[12 * (size_t)i];
      const float* b1 = b0 + 6;
      hnodes[4 * (size_t)i + 0] = make_float4(b0[0] - pad, b0[3] + pad, b0[1] - pad, b0[4] + pad);
      hnodes[4 * (size_t)i + 1] = make_float4(b1[0] - pad, b1[3] + pad, b1[1] - pad, b1[4] + pad);
      hnodes[4 * (size_t)i + 2] = make_float4(b0[2] - pad, b0[5] + pad, b1[2] - pad, b1[5] + pad);
      int c0 = sah.c0[i], c1 = sah.c1[i];
      float4 q;
      memcpy(&q.x, &c0, 4); memcpy(&q.y, &c1, 4); q.z = q.w = 0.f;
      hnodes[4 * (size_t)i + 3] = q;
    }
    std::vector<uint64_t> hkeys((size_t)nt);
    for (int j = 0; j < nt; j++) hkeys[j] = (uint64_t)(uint32_t)sah.order[j];
    PM_SC(cudaMemcpyAsync(s->nodes, hnodes.data(), sizeof(float4) * hnodes.size(), cudaMemcpyHostToDevice, st));
    PM_SC(cudaMemcpyAsync(keys_out, hkeys.data(), sizeof(uint64_t) * hkeys.size(), cudaMemcpyHostToDevice, st));
    k_permute_tris<<<(unsigned)pm::div_up(3 * (int64_t)nt, 256), 256, 0, st>>>(nt, keys_out, rec, s->tris);
    ctx->launches++;
    PM_SC(cudaStreamSynchronize(st)); /* the host vectors go out of scope */
  } else {
    size_t tb = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, tb, keys_in, keys_out, nt, 0, 62, st);
    void* tmp = nullptr;
    PM_S(ctx->get_scratch(pm::S_CUB, tb, &tmp));
    PM_SC(cub::DeviceRadixSort::SortKeys(tmp, tb, keys_in, keys_out, nt, 0, 62, st));
    PM_SC(cudaMemsetAsync(visits, 0, sizeof(int) * (size_t)nt, st));
    k_hierarchy<<<nb, 256, 0, st>>>(keys_out, nt, child0, child1, parent);
    k_refit<<<nb, 256, 0, st>>>(nt, keys_out, tri_box, child0, child1, parent, node_box, visits);
    k_pack<<<nb, 256, 0, st>>>(nt, pad, child0, child1, node_box, s->nodes);
    k_permute_tris<<<(unsigned)pm::div_up(3 * (int64_t)nt, 256), 256, 0, st>>>(nt, keys_out, rec, s->tris);
    ctx->launches += 9;
  }
  PM_SC(cudaGetLastError());
  PM_SC(cudaStreamSynchronize(st));
#undef PM_S
#undef PM_SC
  *out = s;
  return PM_OK;
}

void pm_scene_destroy(pm_scene* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  cudaDeviceSynchronize();
  if (s->nodes) cudaFree(s->nodes);
  if (s->tris) cudaFree(s->tris);
  if (s->mats) cudaFree(s->mats);
  delete s;
}

int pm_trace_closest(pm_ctx* ctx, const pm_scene* scene, const pm_ray* rays, int64_t n, pm_hit* hits) {
  PM_REQUIRE(ctx && scene, "ctx/scene is NULL");
  PM_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return PM_OK;
  PM_REQUIRE(rays && hits, "rays/hits is NULL");
  PM_CUDA(cudaSetDevice(ctx->device));
  const void* r_dev = nullptr;
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN0, rays, sizeof(pm_ray) * (size_t)n, &r_dev));
  pm::OutStage o;
  PM_CHECK(pm::stage_out(ctx, pm::S_STAGE_OUT0, hits, sizeof(pm_hit) * (size_t)n, &o));
  PM_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  k_trace_closest<<<(unsigned)pm::div_up(n, 128), 128, 0, ctx->stream>>>(scene->view(), (const pm_ray*)r_dev, n, (pm_hit*)o.dev);
  ctx->launches++;
  PM_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
  PM_CUDA(cudaGetLastError());
  PM_CHECK(pm::stage_out_finish(ctx, o));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  PM_CUDA(cudaEventElapsedTime(&ctx->last_kernel_ms, ctx->ev0, ctx->ev1));
  return PM_OK;
}

int pm_trace_occluded(pm_ctx* ctx, const pm_scene* scene, const pm_ray* rays, int64_t n, uint8_t* occluded) {
  PM_REQUIRE(ctx && scene, "ctx/scene is NULL");
  PM_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return PM_OK;
  PM_REQUIRE(rays && occluded, "rays/occluded is NULL");
  PM_CUDA(cudaSetDevice(ctx->device));
  const void* r_dev = nullptr;
  PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN0, rays, sizeof(pm_ray) * (size_t)n, &r_dev));
  pm::OutStage o;
  PM_CHECK(pm::stage_out(ctx, pm::S_STAGE_OUT0, occluded, (size_t)n, &o));
  k_trace_occluded<<<(unsigned)pm::div_up(n, 128), 128, 0, ctx->stream>>>(scene->view(), (const pm_ray*)r_dev, n, (uint8_t*)o.dev);
  ctx->launches++;
  PM_CUDA(cudaGetLastError());
  PM_CHECK(pm::stage_out_finish(ctx, o));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  return PM_OK;
}

}  // extern "C"
