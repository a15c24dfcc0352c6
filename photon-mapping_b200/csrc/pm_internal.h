/* pm_internal.h — shared internals of libpmhot (not part of the ABI). */
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/pmhot.h"

namespace pm {

void set_error(const char* fmt, ...);

#define PM_CUDA(call)                                                                       \
  do {                                                                                      \
    cudaError_t e__ = (call);                                                               \
    if (e__ != cudaSuccess) {                                                               \
      pm::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
      cudaGetLastError(); /* reported: do not leave it for an unrelated later check */     \
      return PM_ERR_CUDA;                                                                   \
    }                                                                                       \
  } while (0)

#define PM_CHECK(call)            \
  do {                            \
    int r__ = (call);             \
    if (r__ != PM_OK) return r__; \
  } while (0)

#define PM_REQUIRE(cond, msg)                       \
  do {                                              \
    if (!(cond)) {                                  \
      pm::set_error("%s: %s", __func__, msg);       \
      return PM_ERR_INVALID;                        \
    }                                               \
  } while (0)

inline int64_t div_up(int64_t a, int64_t b) { return (a + b - 1) / b; }

/* limits of the multi-GPU entry points (pm_comm.cu): lights as in pm_emit_photons, ranks of one node */
constexpr int kMaxLightsComm = 64, kMaxRanksComm = 16;

/* Buffers a photon map owns (query-side copies, grid copy) come from the device's stream-ordered pool: a render
 * loop that builds and drops one map per pass (progressive mode: 64 of them) then reuses the same memory without
 * the device-wide synchronisation and the page mapping cost of cudaMalloc / cudaFree. pm_ctx_create lets the pool
 * keep up to a third of the device's memory; anything beyond goes back to the driver at the next synchronisation. */
template <typename T>
inline cudaError_t map_alloc(T** p, size_t bytes, cudaStream_t st) { return cudaMallocAsync(reinterpret_cast<void**>(p), bytes, st); }
inline void map_free(void* p) { if (p) cudaFreeAsync(p, (cudaStream_t)0); }

/* growable device scratch buffers owned by the context, reused across calls so the hot calls do
 * not cudaMalloc/cudaFree (both synchronise the device) */
struct Scratch {
  void* ptr = nullptr;
  size_t cap = 0;
};

enum ScratchSlot {
  S_STAGE_IN0 = 0, S_STAGE_IN1, S_STAGE_IN2, S_STAGE_OUT0, S_STAGE_OUT1, S_STAGE_OUT2,
  S_KEYS0, S_KEYS1, S_VALS0, S_VALS1, S_CUB, S_PERM, S_MISC0, S_MISC1, S_MISC2, S_MISC3,
  S_BUILD0, S_BUILD1, S_BUILD2, S_BUILD3, S_BUILD4, S_BUILD5, S_BUILD6, S_BUILD7, S_BUILD8,
  S_BUILD9, S_BUILD10, S_BUILD11, S_BUILD12, S_BUILD13, S_BUILD14, S_BUILD15, S_BUILD16, S_BUILD17, S_BUILD18, S_BUILD19,
  S_RENDER0, S_RENDER1, S_RENDER2, S_RENDER3, S_RENDER4, S_RENDER5, S_RENDER6, S_RENDER7,
  S_RENDER8, S_RENDER9, S_RENDER10, S_RENDER11,
  S_FAST0, S_FAST1, S_FAST2,
  S_COMM0, S_COMM1, S_COMM2, S_COMM3, S_COMM4, S_COMM5, S_COMM6, S_COMM7, /* pm_comm.cu: never shared with a build */
  S_COUNT
};

}  // namespace pm

struct pm_ctx {
  int device = 0;
  int sm_count = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaEvent_t ev_order = nullptr;   /* pm_ctx_wait_stream */
  int* host_counts = nullptr;       /* pinned: [0] kd-tree fallbacks, [1] hand-overs of the last query call */
  /* second stream + events for the chunked host-buffer pipeline of pm_knn (created on first use) */
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_kernel[2] = {nullptr, nullptr}, ev_copy[2] = {nullptr, nullptr};
  cudaStream_t in_stream = nullptr; /* host queries of that pipeline: the other PCIe direction */
  cudaEvent_t ev_in = nullptr;
  std::vector<cudaEvent_t> pipe_events; /* pm_gather's host-buffer pipeline: two per chunk */
  int64_t launches = 0;
  float last_kernel_ms = 0.f;
  int64_t last_fallbacks = 0;      /* queries of the last pm_knn / pm_gather answered by the kd-tree kernel */
  int* fb_count_dev = nullptr;     /* where the last query batch counted them (scratch) */
  int* ho_count_dev = nullptr;     /* queries the descriptor fast path handed to k_grid_query (scratch) */
  int64_t last_handovers = 0;
  pm::Scratch scratch[pm::S_COUNT];

  /* returns a device buffer of at least `bytes` in slot `slot` (contents undefined) */
  int get_scratch(int slot, size_t bytes, void** out);
  template <class T> int get_scratch(int slot, size_t count, T** out) {
    void* p = nullptr;
    int r = get_scratch(slot, count * sizeof(T), &p);
    *out = (T*)p;
    return r;
  }
};

namespace pm {

/* true if `p` is device-accessible global memory (device or managed allocation) */
bool is_device_ptr(const void* p);

/* Input staging: if `p` is a host pointer copy `bytes` to scratch slot and return the device
 * alias, else return p unchanged. */
int stage_in(pm_ctx* ctx, int slot, const void* p, size_t bytes, const void** dev);
/* Output staging: if `p` is host memory return a scratch device buffer to write to and remember to
 * copy back with stage_out_finish. */
struct OutStage { void* host = nullptr; void* dev = nullptr; size_t bytes = 0; };
int stage_out(pm_ctx* ctx, int slot, void* p, size_t bytes, OutStage* st);
int stage_out_finish(pm_ctx* ctx, const OutStage& st);

}  // namespace pm

/* pm_hit_order_create: a stable spatial order of a fixed set of hit points, reused by every progressive pass */
struct pm_hit_order {
  int device = 0;
  int64_t nq = 0;
  int* perm = nullptr; /* device, nq indices */
};

/* the photon map: tree-ordered AoS records (what the reference hands to its kernels) plus the
 * query-side SoA copies this library traverses */
struct pm_map {
  pm_ctx* ctx = nullptr;   /* creator; never dereferenced on destroy (it may be gone already) */
  int device = 0;
  int64_t n = 0;
  const pm_photon44* photons_dev = nullptr;  /* tree order; may alias caller memory */
  pm_photon44* owned_photons = nullptr;      /* set if the map owns the AoS copy */
  float4* nodes4 = nullptr;                  /* (x, y, z, bits(split_dim)) per node, tree order */
  float4* color4 = nullptr;                  /* (r, g, b, power) per node */
  float bounds[6] = {0, 0, 0, 0, 0, 0};
  /* uniform-grid copy for the warp-cooperative query kernels (pm_grid.cu); absent for small maps */
  float4* grid_pts = nullptr;                /* (x, y, z, bits(node id)) sorted by cell, x fastest */
  uint32_t* cell_start = nullptr;            /* n_cells + 1 offsets into grid_pts */
  uint32_t* cell_start_t = nullptr;          /* fast-path maps: the offsets again with x slowest ((nx + 1) * ny * nz words) */
  float cell = 0.f, inv_cell = 0.f, cell_margin = 0.f;
  int gdim[3] = {0, 0, 0};
  float tuned_rk = 0.f;                      /* sampled k-th neighbour distance the cell size came from */
  float tuned_r50 = 0.f;                     /* median of the sampled distances to the 64th neighbour */
  bool fast_path = false;                    /* uniform-like cloud: k_knn_fast answers most queries */
  bool cells_only = false;                   /* pm_map_build_cells: input order, no kd-tree; fixed-radius gathers only */
};

namespace pm {
/* pm_grid.cu: builds the uniform-grid copy of a map whose nodes4 and bounds are valid; maps it
 * leaves without a grid are answered by the kd-tree kernels alone. */
int grid_build(pm_ctx* ctx, pm_map* m, float cell_size = 0.f);
/* pm_grid.cu: cells-only maps — colours into cell order, record ids = positions */
int cells_reorder(pm_ctx* ctx, pm_map* m);
/* pm_knn.cu: axis-aligned bounds of nodes4 into m->bounds (maps wrapped with pm_map_from_tree) */
int map_bounds(pm_ctx* ctx, pm_map* m);
/* pm_knn.cu: fused kNN + radiance estimate on device pointers (used by pm_gather and pm_render).
 * Queries whose brdf is not > 0 are not evaluated and yield (0,0,0). */
int gather_device(pm_ctx* ctx, const pm_map* map, const float* hit_dev, const float* brdf_dev, int64_t nq, int k,
                  float cutoff, float* rgb_dev, bool time_it);
}  // namespace pm
