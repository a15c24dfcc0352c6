/* ref_shim.cu — thin extern "C" entry points around the REFERENCE's own code, compiled from the
 * sources where they lie under /root/reference (see Makefile target `ref`; output goes to
 * oracle/_ref/, which is git-ignored). TEST INFRASTRUCTURE ONLY: used to pin the oracle and as the
 * "reference on the same box" comparator; never linked into the product.
 *
 * What is reference code here: cukd::buildTree_host / cukd::buildTree (builder_thrust.h),
 * cukd::stackBased::knn + HeapCandidateList (knn.h, traverse-default-stack-based.h),
 * owl::common::LCG<4> (random.h) and owl's normalize (vec.h). What is ours: only the glue.
 *
 * The Photon/Photon_traits pair mirrors ray-tracer/include/photon.h:11-40; the reference marks
 * get_dim/set_dim __device__-only, so (as SURVEY §8c notes) the host builder needs them __host__
 * __device__. ray-tracer/include/photon.h itself cannot be included: it pulls in <owl/owl.h>,
 * which needs the OptiX SDK headers that are not on this machine.
 */
#include <cstdint>
#include <cfloat>
#include <cuda_runtime.h>
#include <cukd/builder.h>
#include <cukd/knn.h>
#include <owl/common/math/vec.h>
#include <owl/common/math/random.h>

struct Photon {
  float3 pos;
  float3 dir;
  float3 color;
  float power;
  uint8_t quantized_normal[3];
  uint8_t split_dim;
};
static_assert(sizeof(Photon) == 44, "Photon must be 44 bytes");

struct Photon_traits : public cukd::default_data_traits<float3> {
  using point_t = float3;
  enum { has_explicit_dim = true };
  static inline __device__ __host__ float3 get_point(const Photon& d) { return d.pos; }
  static inline __device__ __host__ float get_coord(const Photon& d, int dim) { return cukd::get_coord(get_point(d), dim); }
  static inline __device__ __host__ int get_dim(const Photon& p) { return p.split_dim; }
  static inline __device__ __host__ void set_dim(Photon& p, int dim) { p.split_dim = dim; }
};

template <int K>
__global__ void ref_knn_kernel(const Photon* nodes, int n, const float3* queries, int nq, float cutoff,
                               uint64_t* out_entries, float* out_r2) {
  const int tid = threadIdx.x + blockIdx.x * blockDim.x;
  if (tid >= nq) return;
  cukd::HeapCandidateList<K> result(cutoff);
  const float r2 = cukd::stackBased::knn<cukd::HeapCandidateList<K>, Photon, Photon_traits>(result, queries[tid], nodes, n);
  for (int i = 0; i < K; i++)
    out_entries[(size_t)tid * K + i] =
        (uint64_t(__float_as_uint(result.get_dist2(i))) << 32) | uint32_t(result.get_pointID(i));
  if (out_r2) out_r2[tid] = r2;
}

extern "C" {

/* CPU: the reference's host builder, in place; bounds_out = {lo.xyz, hi.xyz} */
void ref_build_tree_host(void* photons, int n, float* bounds_out) {
  cukd::box_t<float3> box;
  cukd::buildTree_host<Photon, Photon_traits>((Photon*)photons, n, &box);
  if (bounds_out && n > 0) {
    bounds_out[0] = box.lower.x; bounds_out[1] = box.lower.y; bounds_out[2] = box.lower.z;
    bounds_out[3] = box.upper.x; bounds_out[4] = box.upper.y; bounds_out[5] = box.upper.z;
  }
}

/* GPU: the builder the reference calls (ray-tracer/src/hostCode.cu:94); device pointers.
 * bounds_dev: 6 floats of device (or managed) memory. Returns a cudaError_t. */
int ref_build_tree_gpu(void* photons_dev, int n, void* bounds_dev) {
  try {
    cukd::buildTree<Photon, Photon_traits>((Photon*)photons_dev, n, (cukd::box_t<float3>*)bounds_dev);
  } catch (...) { return -1; }
  return (int)cudaDeviceSynchronize();
}

/* GPU: one thread per query, block 128, as cudaKDTree/testing/floatN-knn-and-fcp.cu:140-172.
 * out_entries: nq*k raw heap entries (dist² bits << 32 | id), heap order. Asynchronous on the
 * default stream; returns the launch error. */
int ref_knn_gpu(const void* nodes_dev, int n, const void* queries_dev, int nq, int k, float cutoff,
                uint64_t* out_entries_dev, float* out_r2_dev) {
  const int bs = 128, nb = (nq + bs - 1) / bs;
  if (nq <= 0) return 0;
  if (k == 50)
    ref_knn_kernel<50><<<nb, bs>>>((const Photon*)nodes_dev, n, (const float3*)queries_dev, nq, cutoff, out_entries_dev, out_r2_dev);
  else if (k == 64)
    ref_knn_kernel<64><<<nb, bs>>>((const Photon*)nodes_dev, n, (const float3*)queries_dev, nq, cutoff, out_entries_dev, out_r2_dev);
  else if (k == 8)
    ref_knn_kernel<8><<<nb, bs>>>((const Photon*)nodes_dev, n, (const float3*)queries_dev, nq, cutoff, out_entries_dev, out_r2_dev);
  else
    return -1;
  return (int)cudaGetLastError();
}

/* owl::LCG<4>: init(v0,v1), then n draws; returns the final state */
uint32_t ref_lcg(uint32_t v0, uint32_t v1, int n, float* out) {
  owl::common::LCG<4> rng;
  rng.init(v0, v1);
  for (int i = 0; i < n; i++) out[i] = rng();
  return rng.state;
}

void ref_normalize(float x, float y, float z, float* out) {
  const owl::common::vec3f v = owl::common::normalize(owl::common::vec3f(x, y, z));
  out[0] = v.x; out[1] = v.y; out[2] = v.z;
}

}  // extern "C"
