/* pm_comm.cu — the multi-GPU data path of the hot path, in the library (one process per GPU; NCCL over NVLink).
 *
 * The reference is single-GPU (owlContextCreate(nullptr, 1), photon-mapping/src/hostCode.cu:145); everything here
 * is SURVEY §8(e) ground: photons sharded by launch index, ONE exchange into a replicated photon array, a sharded
 * build of the same kd-tree, and the per-pass sum of a progressive render.
 *
 *  - NCCL is bound at run time (dlopen("libnccl.so.2")): the library carries no link-time dependency, a process
 *    that already loaded NCCL (torch) shares that copy, and a single-GPU host never touches it.
 *  - pm_exchange_photons needs no sort and no keys. Rank r traced launch indices r, r + G, ... of every light and
 *    holds its records in (light, launch index, bounce) order with one byte per path saying how many it stored.
 *    The single-GPU order is the same triple over ALL paths, so the final position of a path's records is the
 *    exclusive prefix sum of the per-path counts taken in global path order — two all-gathers (counts: 1 B per path,
 *    records: 40 B each), two single-pass scans (global order for the destination, rank-major for the source) and
 *    one placement kernel that also widens the 40-byte record to the 44-byte map record.
 *  - pm_ppm_update is the Hachisuka radius/flux update of a progressive pass as one kernel; pm_comm_allreduce_sum
 *    sums the (flux.rgb, count) planes in place on the context's stream.
 */
#include <dlfcn.h>
#include <nccl.h>
#include "pm_internal.h"

namespace {

struct NcclApi {
  void* so = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi* nccl() {
  static NcclApi api;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!so) so = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (so) {
      api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(so, "ncclGetUniqueId");
      api.CommInitRank = (decltype(api.CommInitRank))dlsym(so, "ncclCommInitRank");
      api.CommDestroy = (decltype(api.CommDestroy))dlsym(so, "ncclCommDestroy");
      api.AllGather = (decltype(api.AllGather))dlsym(so, "ncclAllGather");
      api.AllReduce = (decltype(api.AllReduce))dlsym(so, "ncclAllReduce");
      api.GetErrorString = (decltype(api.GetErrorString))dlsym(so, "ncclGetErrorString");
      if (api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.AllReduce && api.GetErrorString)
        api.so = so;
    }
  }
  return api.so ? &api : nullptr;
}

#define PM_NCCL(call)                                                                             \
  do {                                                                                            \
    ncclResult_t r__ = (call);                                                                    \
    if (r__ != ncclSuccess) {                                                                     \
      pm::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, nccl()->GetErrorString(r__));   \
      return PM_ERR_CUDA;                                                                         \
    }                                                                                             \
  } while (0)

}  // namespace

struct pm_comm {
  pm_ctx* ctx = nullptr;
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
};

namespace {

/* ---- exclusive prefix sum of 64-bit values, single pass (decoupled look-back) ----------------------------
 * Input through a functor (index -> value); 2048 elements per CTA; tile numbers come from a counter so that a tile
 * never waits for one that has not started. status[t] = flag << 62 | value, flag 1 = the tile's own sum,
 * 2 = the inclusive sum up to and including the tile. */
constexpr int kTile = 2048;
constexpr unsigned long long kValMask = (1ull << 62) - 1;

template <class In>
__global__ void __launch_bounds__(256)
k_scan_i64(In in, int64_t n, int64_t* __restrict__ out, unsigned long long* __restrict__ status, int* __restrict__ tile_counter,
           int64_t* __restrict__ total_out) {
  __shared__ int s_tile;
  __shared__ unsigned long long s_warp[8];
  __shared__ unsigned long long s_prefix;
  if (threadIdx.x == 0) s_tile = atomicAdd(tile_counter, 1);
  __syncthreads();
  const int tile = s_tile;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t base = (int64_t)tile * kTile + threadIdx.x * 8;
  unsigned long long v[8], sum = 0;
#pragma unroll
  for (int e = 0; e < 8; e++) { v[e] = base + e < n ? (unsigned long long)in(base + e) : 0ull; sum += v[e]; }
  unsigned long long incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long u = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += u;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  unsigned long long wbase = 0, total = 0;
#pragma unroll
  for (int w = 0; w < 8; w++) {
    if (w < warp) wbase += s_warp[w];
    total += s_warp[w];
  }
  if (warp == 0) {
    unsigned long long prefix = 0;
    if (tile == 0) {
      if (lane == 0) atomicExch(status, (2ull << 62) | total);
    } else {
      if (lane == 0) atomicExch(status + tile, (1ull << 62) | total);
      int first = tile - 1;
      while (true) {
        const int tt = first - lane;
        unsigned long long st = 0;
        bool ready;
        do {
          st = tt >= 0 ? *((volatile unsigned long long*)(status + tt)) : (2ull << 62);
          ready = (st >> 62) != 0;
        } while (!__all_sync(0xffffffffu, ready));
        const unsigned full = __ballot_sync(0xffffffffu, (st >> 62) == 2);
        const int stop = full ? __ffs(full) - 1 : 32;
        unsigned long long c = lane <= stop ? (st & kValMask) : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        prefix += c;
        if (full) break;
        first -= 32;
      }
      if (lane == 0) atomicExch(status + tile, (2ull << 62) | (prefix + total));
    }
    if (lane == 0) {
      s_prefix = prefix;
      if (total_out && (int64_t)(tile + 1) * kTile >= n) *total_out = (int64_t)(prefix + total);
    }
  }
  __syncthreads();
  unsigned long long run = s_prefix + wbase + incl - sum;
#pragma unroll
  for (int e = 0; e < 8; e++) {
    if (base + e < n) out[base + e] = (int64_t)run;
    run += v[e];
  }
}

/* how the paths of a sharded emission are laid out: light l launches n[l] paths; launch index i belongs to rank
 * i % world and is that rank's local path lf[rank][l] + i / world */
struct ShardLayout {
  int n_lights, world;
  int64_t light_base[pm::kMaxLightsComm + 1];                 /* global path index of (l, 0) */
  int64_t lf[pm::kMaxRanksComm][pm::kMaxLightsComm + 1];      /* local path index of rank r's first path of light l */
  int64_t pcap;                                               /* row pitch of the gathered count table */
};

__device__ __forceinline__ void locate(const ShardLayout& L, int64_t g, int* r, int64_t* p) {
  int l = 0;
  while (l + 1 < L.n_lights && g >= L.light_base[l + 1]) l++;
  const int64_t i = g - L.light_base[l];
  *r = (int)(i % L.world);
  *p = L.lf[*r][l] + i / L.world;
}

struct GlobalOrderCounts {
  ShardLayout L;
  const uint8_t* cnt;
  __device__ int64_t operator()(int64_t g) const {
    int r; int64_t p;
    locate(L, g, &r, &p);
    return (int64_t)cnt[(size_t)r * L.pcap + p];
  }
};

/* rank-major order with the running sum restarting at every rank: value = count, but the scan is made per rank by
 * running it over one rank's row at a time (world launches would cost world tails; instead the row index is folded
 * into the element index and the row's base is subtracted afterwards by the consumer) */
struct RowCounts {
  const uint8_t* cnt;
  __device__ int64_t operator()(int64_t i) const { return (int64_t)cnt[i]; }
};

/* one record word per thread: 11 threads per path walk its records. src = staging[rank][local offset + k] (40 B),
 * dst = out[global offset + k] (44 B: pos, dir, colour, power, 4 zero bytes). */
__global__ void k_place(ShardLayout L, int64_t n_paths, const uint8_t* __restrict__ cnt, const int64_t* __restrict__ goff,
                        const int64_t* __restrict__ roff /* scan over the whole [world][pcap] table */,
                        const pm_photon40* __restrict__ staging, int64_t cap, float power, pm_photon44* __restrict__ out,
                        int64_t limit) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t g = t / 11;
  const int w = (int)(t - g * 11);
  if (g >= n_paths) return;
  int r; int64_t p;
  locate(L, g, &r, &p);
  const size_t ci = (size_t)r * L.pcap + p;
  const int c = cnt[ci];
  if (c == 0) return;
  const int64_t lo = roff[ci] - roff[(size_t)r * L.pcap]; /* offset inside rank r's shard */
  const int64_t o = goff[g];
  const uint32_t* src = reinterpret_cast<const uint32_t*>(staging + (size_t)r * cap + lo);
  uint32_t* dst = reinterpret_cast<uint32_t*>(out + o);
  for (int k = 0; k < c; k++) {
    if (o + k >= limit) break;
    uint32_t v;
    if (w < 6) v = src[(size_t)k * 10 + w];               /* pos, dir */
    else if (w < 9) v = src[(size_t)k * 10 + w + 1];      /* colour sits after the int power */
    else if (w == 9) v = __float_as_uint(power);
    else v = 0u;                                          /* quantized_normal, split_dim */
    dst[(size_t)k * 11 + w] = v;
  }
}

template <class In>
int scan_i64(pm_ctx* ctx, In in, int64_t n, int64_t* out, int64_t* total_dev) {
  const int64_t tiles = pm::div_up(n, kTile);
  unsigned long long* status = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_MISC3, (size_t)tiles + 2, &status));
  PM_CUDA(cudaMemsetAsync(status, 0, sizeof(unsigned long long) * ((size_t)tiles + 2), ctx->stream));
  int* counter = reinterpret_cast<int*>(status + tiles);
  k_scan_i64<<<(unsigned)tiles, 256, 0, ctx->stream>>>(in, n, out, status, counter, total_dev);
  ctx->launches++;
  PM_CUDA(cudaGetLastError());
  return PM_OK;
}

int make_layout(const pm_light* lights, int n_lights, int64_t casted, int world, ShardLayout* L, int64_t* total_paths) {
  PM_REQUIRE(n_lights >= 0 && n_lights <= pm::kMaxLightsComm, "n_lights out of range (max 64)");
  PM_REQUIRE(world >= 1 && world <= pm::kMaxRanksComm, "world out of range (max 16 ranks)");
  memset(L, 0, sizeof *L);
  L->n_lights = n_lights; L->world = world;
  double total_power = 0;
  for (int i = 0; i < n_lights; i++) total_power += lights[i].power;
  /* the launch sizes pm_emit_photons computes (photon-mapping/src/hostCode.cu:86,102-110) */
  const int per_watt = n_lights > 0 && total_power > 0 ? (int)((double)(int)casted / total_power) : 0;
  int64_t pmax = 0;
  for (int l = 0; l < n_lights; l++) {
    const int64_t n = (int64_t)(int)(lights[l].power * per_watt);
    L->light_base[l + 1] = L->light_base[l] + n;
    for (int r = 0; r < world; r++) L->lf[r][l + 1] = L->lf[r][l] + (n > r ? (n - r + world - 1) / world : 0);
  }
  for (int r = 0; r < world; r++) pmax = std::max(pmax, L->lf[r][n_lights]);
  L->pcap = (pmax + 15) & ~(int64_t)15; /* 16-byte rows for the all-gather */
  *total_paths = L->light_base[n_lights];
  return PM_OK;
}

}  // namespace

namespace pm {

/* the single-GPU part of the exchange: counts [world][pcap] and records [world][cap] already gathered */
int exchange_place(pm_ctx* ctx, const ShardLayout& L, int64_t total_paths, const uint8_t* cnt_all, const pm_photon40* staging,
                   int64_t cap, float power, pm_photon44* out, int64_t capacity, int64_t* total) {
  cudaStream_t st = ctx->stream;
  int64_t *goff = nullptr, *roff = nullptr, *tot_dev = nullptr;
  const int64_t table = (int64_t)L.world * L.pcap;
  PM_CHECK(ctx->get_scratch(pm::S_COMM0, (size_t)total_paths + 1, &goff));
  PM_CHECK(ctx->get_scratch(pm::S_COMM1, (size_t)table + 1, &roff));
  PM_CHECK(ctx->get_scratch(pm::S_MISC0, (size_t)4, &tot_dev));
  GlobalOrderCounts go{L, cnt_all};
  PM_CHECK(scan_i64(ctx, go, total_paths, goff, tot_dev));
  RowCounts rc{cnt_all};
  PM_CHECK(scan_i64(ctx, rc, table, roff, nullptr));
  int64_t tot = 0;
  PM_CUDA(cudaMemcpyAsync(&tot, tot_dev, sizeof tot, cudaMemcpyDeviceToHost, st));
  PM_CUDA(cudaStreamSynchronize(st));
  *total = tot;
  if (tot > capacity) {
    pm::set_error("pm_exchange_photons: %lld photons over all ranks but capacity is %lld", (long long)tot, (long long)capacity);
    return PM_ERR_CAPACITY;
  }
  if (tot > 0) {
    k_place<<<(unsigned)div_up(total_paths * 11, 256), 256, 0, st>>>(L, total_paths, cnt_all, goff, roff, staging, cap, power, out,
                                                                    capacity);
    ctx->launches++;
    PM_CUDA(cudaGetLastError());
  }
  PM_CUDA(cudaStreamSynchronize(st));
  return PM_OK;
}

}  // namespace pm

/* ---- progressive pass update (Hachisuka et al.; ppm.py:ppm_update, operation for operation) ---------------- */
__global__ void k_ppm_update(int64_t npx, const float* __restrict__ flux, const int32_t* __restrict__ count_i,
                             const float* __restrict__ fc /* (flux.rgb, count) planes summed over ranks, or NULL */, float alpha,
                             float* __restrict__ N, float* __restrict__ r2, float* __restrict__ tau) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npx) return;
  float M, fr, fg, fb;
  if (fc) { fr = fc[4 * i]; fg = fc[4 * i + 1]; fb = fc[4 * i + 2]; M = fc[4 * i + 3]; }
  else { fr = flux[3 * i]; fg = flux[3 * i + 1]; fb = flux[3 * i + 2]; M = (float)count_i[i]; }
  const float n_old = N[i];
  const bool hit = M > 0.f;
  const float n_new = __fadd_rn(n_old, __fmul_rn(alpha, M));
  const float ratio = hit ? __fdiv_rn(n_new, fmaxf(__fadd_rn(n_old, M), 1e-30f)) : 1.f;
  r2[i] = __fmul_rn(r2[i], ratio);
  tau[3 * i + 0] = __fmul_rn(__fadd_rn(tau[3 * i + 0], fr), ratio);
  tau[3 * i + 1] = __fmul_rn(__fadd_rn(tau[3 * i + 1], fg), ratio);
  tau[3 * i + 2] = __fmul_rn(__fadd_rn(tau[3 * i + 2], fb), ratio);
  if (hit) N[i] = n_new;
}

__global__ void k_pack_fc(int64_t npx, const float* __restrict__ flux, const int32_t* __restrict__ count, float* __restrict__ fc) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npx) return;
  reinterpret_cast<float4*>(fc)[i] = make_float4(flux[3 * i], flux[3 * i + 1], flux[3 * i + 2], (float)count[i]);
}

__global__ void k_counts_add_dim(int64_t n, const int32_t* __restrict__ a, int32_t* __restrict__ b) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) b[i] += a[i];
}

extern "C" {

int pm_comm_unique_id(void* id_out, int bytes) {
  PM_REQUIRE(id_out && bytes >= (int)sizeof(ncclUniqueId), "id buffer must hold PM_COMM_ID_BYTES bytes");
  if (!nccl()) { pm::set_error("pm_comm_unique_id: libnccl.so.2 could not be loaded"); return PM_ERR_UNSUPPORTED; }
  ncclUniqueId id;
  PM_NCCL(nccl()->GetUniqueId(&id));
  memset(id_out, 0, (size_t)bytes);
  memcpy(id_out, &id, sizeof id);
  return PM_OK;
}

int pm_comm_create(pm_ctx* ctx, const void* id_bytes, int rank, int world, pm_comm** out) {
  PM_REQUIRE(ctx && id_bytes && out, "ctx/id/out is NULL");
  *out = nullptr;
  PM_REQUIRE(world >= 1 && world <= pm::kMaxRanksComm && rank >= 0 && rank < world, "0 <= rank < world <= 16");
  if (!nccl()) { pm::set_error("pm_comm_create: libnccl.so.2 could not be loaded"); return PM_ERR_UNSUPPORTED; }
  PM_CUDA(cudaSetDevice(ctx->device));
  ncclUniqueId id;
  memcpy(&id, id_bytes, sizeof id);
  pm_comm* c = new pm_comm();
  c->ctx = ctx; c->rank = rank; c->world = world;
  ncclResult_t r = nccl()->CommInitRank(&c->comm, world, id, rank);
  if (r != ncclSuccess) {
    pm::set_error("pm_comm_create: ncclCommInitRank -> %s", nccl()->GetErrorString(r));
    delete c;
    return PM_ERR_CUDA;
  }
  *out = c;
  return PM_OK;
}

void pm_comm_destroy(pm_comm* c) {
  if (!c) return;
  if (c->comm && nccl()) nccl()->CommDestroy(c->comm);
  delete c;
}

int pm_comm_rank(const pm_comm* c) { return c ? c->rank : 0; }
int pm_comm_world(const pm_comm* c) { return c ? c->world : 1; }

int pm_comm_allreduce_sum(pm_ctx* ctx, pm_comm* c, void* buf_dev, int64_t count, int dtype) {
  PM_REQUIRE(ctx && buf_dev && count >= 0, "ctx/buf is NULL");
  PM_REQUIRE(dtype == PM_DTYPE_F32 || dtype == PM_DTYPE_I32 || dtype == PM_DTYPE_I64, "dtype must be PM_DTYPE_F32 / I32 / I64");
  if (!c || c->world == 1 || count == 0) return PM_OK;
  PM_CUDA(cudaSetDevice(ctx->device));
  const ncclDataType_t t = dtype == PM_DTYPE_F32 ? ncclFloat32 : (dtype == PM_DTYPE_I32 ? ncclInt32 : ncclInt64);
  PM_NCCL(nccl()->AllReduce(buf_dev, buf_dev, (size_t)count, t, ncclSum, c->comm, ctx->stream));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  return PM_OK;
}

int pm_comm_allgather(pm_ctx* ctx, pm_comm* c, const void* send_dev, void* recv_dev, int64_t bytes_per_rank) {
  PM_REQUIRE(ctx && send_dev && recv_dev && bytes_per_rank >= 0, "bad arguments");
  PM_CUDA(cudaSetDevice(ctx->device));
  if (!c || c->world == 1) {
    if (send_dev != recv_dev) PM_CUDA(cudaMemcpyAsync(recv_dev, send_dev, (size_t)bytes_per_rank, cudaMemcpyDeviceToDevice, ctx->stream));
  } else {
    PM_NCCL(nccl()->AllGather(send_dev, recv_dev, (size_t)bytes_per_rank, ncclInt8, c->comm, ctx->stream));
  }
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  return PM_OK;
}

int pm_exchange_place(pm_ctx* ctx, int world, const pm_light* lights, int n_lights, int64_t casted,
                      const uint8_t* counts_all_dev, int64_t counts_pitch, const pm_photon40* shards_dev, int64_t shard_pitch,
                      float power, pm_photon44* out_dev, int64_t capacity, int64_t* total) {
  PM_REQUIRE(ctx && counts_all_dev && out_dev && total, "ctx/counts/out/total is NULL");
  PM_REQUIRE(shards_dev || shard_pitch == 0, "shards is NULL");
  PM_CUDA(cudaSetDevice(ctx->device));
  ShardLayout L;
  int64_t total_paths = 0;
  PM_CHECK(make_layout(lights, n_lights, casted, world, &L, &total_paths));
  PM_REQUIRE(counts_pitch >= L.lf[0][n_lights], "counts_pitch smaller than a rank's path count");
  L.pcap = counts_pitch;
  *total = 0;
  if (total_paths == 0) return PM_OK;
  return pm::exchange_place(ctx, L, total_paths, counts_all_dev, shards_dev, shard_pitch, power, out_dev, capacity, total);
}

int pm_exchange_photons(pm_ctx* ctx, pm_comm* c, const pm_light* lights, int n_lights, int64_t casted,
                        const pm_photon40* shard_dev, int64_t n_shard, const uint8_t* path_counts_dev, float power,
                        pm_photon44* out_dev, int64_t capacity, int64_t* total, float* collective_ms) {
  PM_REQUIRE(ctx && out_dev && total, "ctx/out/total is NULL");
  PM_REQUIRE(n_shard >= 0 && (n_shard == 0 || shard_dev), "shard is NULL");
  PM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const int world = c ? c->world : 1, rank = c ? c->rank : 0;
  ShardLayout L;
  int64_t total_paths = 0;
  PM_CHECK(make_layout(lights, n_lights, casted, world, &L, &total_paths));
  *total = 0;
  if (collective_ms) *collective_ms = 0.f;
  if (total_paths == 0) return PM_OK;
  PM_REQUIRE(path_counts_dev, "path_counts is NULL");
  const int64_t my_paths = L.lf[rank][n_lights];
  /* 1. counts: one byte per path, rows padded to pcap */
  uint8_t *cnt_mine = nullptr, *cnt_all = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_COMM2, (size_t)L.pcap, &cnt_mine));
  PM_CHECK(ctx->get_scratch(pm::S_COMM3, (size_t)L.pcap * world, &cnt_all));
  PM_CUDA(cudaMemsetAsync(cnt_mine, 0, (size_t)L.pcap, st));
  PM_CUDA(cudaMemcpyAsync(cnt_mine, path_counts_dev, (size_t)my_paths, cudaMemcpyDeviceToDevice, st));
  /* 2. shard sizes -> common pitch of the record all-gather */
  int64_t* sizes_dev = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_MISC0, (size_t)(2 * pm::kMaxRanksComm), &sizes_dev));
  int64_t sizes[pm::kMaxRanksComm] = {0};
  sizes[rank] = n_shard;
  if (world > 1) {
    PM_CUDA(cudaMemcpyAsync(sizes_dev + pm::kMaxRanksComm + rank, &sizes[rank], sizeof(int64_t), cudaMemcpyHostToDevice, st));
    PM_NCCL(nccl()->AllGather(sizes_dev + pm::kMaxRanksComm + rank, sizes_dev, 1, ncclInt64, c->comm, st));
    PM_CUDA(cudaMemcpyAsync(sizes, sizes_dev, sizeof(int64_t) * world, cudaMemcpyDeviceToHost, st));
    PM_CUDA(cudaStreamSynchronize(st));
  }
  int64_t cap = 1;
  for (int r = 0; r < world; r++) cap = std::max(cap, sizes[r]);
  PM_CUDA(cudaEventRecord(ctx->ev0, st));
  const pm_photon40* staging = shard_dev;
  if (world > 1) {
    PM_NCCL(nccl()->AllGather(cnt_mine, cnt_all, (size_t)L.pcap, ncclInt8, c->comm, st));
    pm_photon40* stage = nullptr;
    PM_CHECK(ctx->get_scratch(pm::S_COMM4, (size_t)cap * world, &stage));
    /* NCCL needs equal contributions: every rank sends `cap` records (the tail of a shorter shard is whatever follows
     * it in the caller's buffer or, if that is shorter than cap, a padded copy) */
    const pm_photon40* send = shard_dev;
    if (n_shard < cap) {
      pm_photon40* padded = stage + (size_t)rank * cap; /* in-place all-gather: own slot of the receive buffer */
      if (n_shard > 0) PM_CUDA(cudaMemcpyAsync(padded, shard_dev, sizeof(pm_photon40) * (size_t)n_shard, cudaMemcpyDeviceToDevice, st));
      send = padded;
    }
    PM_NCCL(nccl()->AllGather(send, stage, sizeof(pm_photon40) * (size_t)cap, ncclInt8, c->comm, st));
    staging = stage;
  } else {
    cnt_all = cnt_mine;
  }
  PM_CUDA(cudaEventRecord(ctx->ev1, st));
  const int rc = pm::exchange_place(ctx, L, total_paths, cnt_all, staging, cap, power, out_dev, capacity, total);
  if (collective_ms) PM_CUDA(cudaEventElapsedTime(collective_ms, ctx->ev0, ctx->ev1));
  return rc;
}

int pm_ppm_update(pm_ctx* ctx, pm_comm* c, int64_t npx, const float* flux_dev, const int32_t* count_dev, float alpha,
                  float* N_dev, float* r2_dev, float* tau_dev, float* collective_ms) {
  PM_REQUIRE(ctx && flux_dev && count_dev && N_dev && r2_dev && tau_dev && npx >= 0, "bad arguments");
  PM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  if (collective_ms) *collective_ms = 0.f;
  if (npx == 0) return PM_OK;
  const unsigned grid = (unsigned)pm::div_up(npx, 256);
  if (c && c->world > 1) {
    float* fc = nullptr;
    PM_CHECK(ctx->get_scratch(pm::S_COMM5, (size_t)npx * 4, &fc));
    k_pack_fc<<<grid, 256, 0, st>>>(npx, flux_dev, count_dev, fc);
    PM_CUDA(cudaEventRecord(ctx->ev0, st));
    PM_NCCL(nccl()->AllReduce(fc, fc, (size_t)npx * 4, ncclFloat32, ncclSum, c->comm, st));
    PM_CUDA(cudaEventRecord(ctx->ev1, st));
    k_ppm_update<<<grid, 256, 0, st>>>(npx, nullptr, nullptr, fc, alpha, N_dev, r2_dev, tau_dev);
    ctx->launches += 2;
  } else {
    k_ppm_update<<<grid, 256, 0, st>>>(npx, flux_dev, count_dev, nullptr, alpha, N_dev, r2_dev, tau_dev);
    ctx->launches++;
  }
  PM_CUDA(cudaGetLastError());
  PM_CUDA(cudaStreamSynchronize(st));
  if (collective_ms && c && c->world > 1) PM_CUDA(cudaEventElapsedTime(collective_ms, ctx->ev0, ctx->ev1));
  return PM_OK;
}

int pm_map_build_sharded(pm_ctx* ctx, pm_comm* c, pm_photon44* photons_dev, int64_t n, float* bounds_host, pm_map** out) {
  PM_REQUIRE(ctx && out, "ctx/out is NULL");
  const int world = c ? c->world : 1, rank = c ? c->rank : 0;
  if (world == 1 || (world & (world - 1)) != 0 || n < 1) return pm_map_build(ctx, photons_dev, n, bounds_host, out);
  PM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  int32_t* elem = nullptr;
  uint8_t* dim = nullptr;
  const size_t dim_words = ((size_t)n + 3) / 4;
  PM_CHECK(ctx->get_scratch(pm::S_COMM6, (size_t)n, &elem));
  PM_CHECK(ctx->get_scratch(pm::S_COMM7, dim_words * 4, &dim));
  PM_CUDA(cudaMemsetAsync(dim, 0, dim_words * 4, st));
  float b[6];
  PM_CHECK(pm_map_build_shard(ctx, photons_dev, n, rank, world, elem, dim, b));
  /* one rank contributes each entry (zero elsewhere): the sums ARE the tables; dim bytes summed as int32 words */
  PM_NCCL(nccl()->AllReduce(elem, elem, (size_t)n, ncclInt32, ncclSum, c->comm, st));
  PM_NCCL(nccl()->AllReduce(dim, dim, dim_words, ncclInt32, ncclSum, c->comm, st));
  PM_CUDA(cudaStreamSynchronize(st));
  if (bounds_host) memcpy(bounds_host, b, sizeof b);
  return pm_map_build_finish(ctx, photons_dev, n, elem, dim, b, out);
}

}  // extern "C"
