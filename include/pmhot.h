/* pmhot.h — C ABI of libpmhot, the B200-native photon-mapping hot path.
 *
 * This is the drop-in boundary for the three device stages of
 * nico-mayora/photon-mapping (citations are paths inside the reference tree):
 *
 *   stage 1  photon emission + multi-bounce tracing
 *            replaces owlRayGenLaunch2D(pointLightRayGen)      photon-mapping/src/hostCode.cu:72-90
 *            and the device programs                           photon-mapping/cuda/deviceCode.cu:10-136
 *   stage 2  photon-map (left-balanced kd-tree) construction
 *            replaces cukd::buildTree<Photon,Photon_traits>    ray-tracer/src/hostCode.cu:94-95
 *   stage 3  per-pixel kNN radiance gather / render
 *            replaces owlRayGenLaunch2D(simpleRayGen)          ray-tracer/src/hostCode.cu:231-237
 *            and cukd::stackBased::knn + gatherPhotons         ray-tracer/cuda/shading.h:11-18,93-121
 *   scene    replaces loadGeometry (OptiX BLAS/TLAS)           common/src/world.cpp:3-58
 *
 * Conventions
 *   - every function returns 0 on success or a negative PM_ERR_* code; pm_last_error()
 *     returns a thread-local message for the last failure. No exceptions cross the ABI.
 *   - plain pointers and sizes only. Pointers named *_host must be host memory, *_dev device
 *     memory; pointers without a suffix may be either (the library asks the CUDA runtime which
 *     it is and stages host buffers through device memory itself).
 *   - a pm_ctx is bound to one CUDA device and owns one stream. Calls are synchronous on
 *     return unless documented otherwise. A context is not thread-safe; use one per thread.
 *   - there is no CPU fallback: without a usable CUDA device pm_ctx_create fails with
 *     PM_ERR_CUDA and nothing else can be called.
 */
#ifndef PMHOT_H
#define PMHOT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PM_OK              0
#define PM_ERR_INVALID    -1   /* bad argument */
#define PM_ERR_CUDA       -2   /* CUDA runtime error (message in pm_last_error) */
#define PM_ERR_NOMEM      -3
#define PM_ERR_IO         -4
#define PM_ERR_CAPACITY   -5   /* output buffer too small; *count holds the required size */
#define PM_ERR_UNSUPPORTED -6

/* ---- record layouts (byte-for-byte the reference structs) ---------------------------- */

/* pass-1 photon, photon-mapping/include/photon.h:5-11 — 40 bytes */
typedef struct pm_photon40 {
  float   pos[3];
  float   dir[3];      /* OUTGOING scattered direction (deviceCode.cu:15) */
  int32_t power;       /* never written by the reference; we write 0 */
  float   color[3];    /* colour carried BEFORE the albedo of this hit */
} pm_photon40;

/* pass-2 photon = kd-tree node, ray-tracer/include/photon.h:11-22 — 44 bytes */
typedef struct pm_photon44 {
  float   pos[3];
  float   dir[3];
  float   color[3];
  float   power;
  uint8_t quantized_normal[3];   /* unused by the reference */
  uint8_t split_dim;             /* filled by pm_map_build */
} pm_photon44;

/* common/src/mesh.h:6-12 — 28 bytes */
typedef struct pm_material {
  float albedo[3];
  float diffuse;
  float specular;
  float transmission;
  float refraction_idx;
} pm_material;

/* point light: the fields of LightSource (common/src/world.h:13-24) the device code reads */
typedef struct pm_light {
  float  pos[3];
  float  rgb[3];
  double power;
} pm_light;

/* emitter shapes: LightSource::source_type / normal / side_length (common/src/world.h:8-20). The reference declares
 * them and parses every light as POINT_LIGHT (common/src/assetImporter.cxx:122); the square emitter is ours
 * (SURVEY §8f rank 4): uniform point on the square centred at pm_light::pos, cosine-distributed direction. */
#define PM_LIGHT_POINT  0
#define PM_LIGHT_SQUARE 1
typedef struct pm_light_shape {
  int32_t type;
  float   normal[3];
  float   side_length;
} pm_light_shape;

/* ray-tracer/include/deviceCode.h:29-34 (already set up as in hostCode.cu:100-108) */
typedef struct pm_camera {
  float pos[3];
  float dir_00[3];
  float dir_du[3];
  float dir_dv[3];
} pm_camera;

typedef struct pm_ray {
  float org[3];
  float tmin;
  float dir[3];
  float tmax;
} pm_ray;

typedef struct pm_hit {
  int32_t prim;     /* global triangle id, -1 on miss */
  float   t;        /* hit distance (ray parameter) */
} pm_hit;

/* RNG modes for pm_emit_photons / pm_render */
#define PM_RNG_LCG_PARITY 0   /* owl::LCG<4> seeded exactly like the reference */
#define PM_RNG_COUNTER    1   /* counter-based hash RNG keyed by (seed, id, draw) */

typedef struct pm_ctx   pm_ctx;
typedef struct pm_scene pm_scene;
typedef struct pm_map   pm_map;

/* ---- context ------------------------------------------------------------------------- */
const char* pm_last_error(void);
int  pm_version(void);
int  pm_ctx_create(int device, pm_ctx** out);
void pm_ctx_destroy(pm_ctx* ctx);
/* cudaStream_t of the context, as an opaque pointer (for event timing by the caller) */
void* pm_ctx_stream(pm_ctx* ctx);
int  pm_ctx_sync(pm_ctx* ctx);
/* Input readiness: the context's stream is created non-blocking, so it is NOT ordered after work the caller queued
 * on other streams (the legacy default stream included). Device buffers handed to a pm_* call must be complete when
 * the call is made — or the caller orders the context's stream after the producing stream first:
 * pm_ctx_wait_stream records an event on `other_stream` (a cudaStream_t; NULL = the legacy default stream) and makes
 * the context's stream wait for it, without blocking the host. Results need no such step: every pm_* call returns
 * after its work on the context's stream has completed. */
int  pm_ctx_wait_stream(pm_ctx* ctx, void* other_stream);
/* device memory for hosts that have no CUDA runtime of their own (the shipped C++ mains link libpmhot only):
 * zero-filled allocation on the context's device, release, and a copy in any direction (host or device pointers on
 * either side), synchronous on return. pm_device_count: CUDA devices visible to the process (0 without a driver). */
int  pm_dev_alloc(pm_ctx* ctx, int64_t bytes, void** out);
void pm_dev_free(pm_ctx* ctx, void* ptr);
int  pm_dev_copy(pm_ctx* ctx, void* dst, const void* src, int64_t bytes);
int  pm_device_count(void);
/* counters: number of kernels launched by this context since creation / last reset */
int64_t pm_ctx_launch_count(pm_ctx* ctx);
void pm_ctx_reset_launch_count(pm_ctx* ctx);
/* device time in ms of the most recent call's dominant kernel(s) (CUDA events on ctx stream) */
float pm_ctx_last_kernel_ms(pm_ctx* ctx);
/* queries of the most recent pm_knn / pm_gather that the uniform-grid kernel handed to the kd-tree
 * kernel (sparse or very dense neighbourhoods); -1 if the map has no grid (small maps) */
int64_t pm_ctx_last_fallback_count(pm_ctx* ctx);
/* queries of the most recent pm_knn / pm_gather that the fast path (uniform-like clouds) handed to the general
 * uniform-grid kernel; 0 if the fast path was not used */
int64_t pm_ctx_last_handover_count(pm_ctx* ctx);

/* ---- scene: LBVH build + traversal  (≙ loadGeometry, common/src/world.cpp:3-58) ------- */
/* verts: 3*nv floats; tris: 3*nt vertex indices; tri_mesh: nt mesh ids; mats: n_mesh records */
int  pm_scene_create(pm_ctx* ctx, const float* verts_host, int nv,
                     const int32_t* tris_host, int nt, const int32_t* tri_mesh_host,
                     const pm_material* mats_host, int n_mesh, pm_scene** out);
void pm_scene_destroy(pm_scene* scene);
/* test hook: closest hit for n rays (≙ optixTrace with closest-hit, both face orientations) */
int  pm_trace_closest(pm_ctx* ctx, const pm_scene* scene, const pm_ray* rays, int64_t n, pm_hit* hits);
/* test hook: occlusion (≙ the TERMINATE_ON_FIRST_HIT shadow ray, ray-tracer/cuda/deviceCode.cu:70-85) */
int  pm_trace_occluded(pm_ctx* ctx, const pm_scene* scene, const pm_ray* rays, int64_t n, uint8_t* occluded);

/* ---- stage 1 (≙ runNormal/runCaustics → runPointLightRayGen loop) -------------------- */
/* Emits for every light int(power * int(casted / sum(power))) photon paths
 * (photon-mapping/src/hostCode.cu:86,102-110), walks them for at most max_depth segments and
 * appends stored photons to `out` (host or device, capacity records). This rank traces launch
 * indices first_photon, first_photon+photon_stride, ... of every light.
 * *count receives the number stored; *n_paths and *n_segments (may be NULL) the work done. */
int  pm_emit_photons(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights_host, int n_lights,
                     int64_t casted, int max_depth, int caustics_mode,
                     uint64_t seed, int rng_mode, int64_t first_photon, int64_t photon_stride,
                     pm_photon40* out, int64_t capacity, int64_t* count,
                     int64_t* n_paths, int64_t* n_segments);

/* Same, and also writes one 64-bit key per stored photon: light << 56 | launch index << 8 | ordinal of
 * the store on its path. Ranks that trace disjoint launch indices (first_photon = rank, photon_stride =
 * ranks) gather their shards and sort by key to obtain exactly the single-GPU output order.
 * out and out_keys must both be host or both be device memory. */
int  pm_emit_photons_keyed(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights_host, int n_lights,
                           int64_t casted, int max_depth, int caustics_mode,
                           uint64_t seed, int rng_mode, int64_t first_photon, int64_t photon_stride,
                           pm_photon40* out, uint64_t* out_keys, int64_t capacity, int64_t* count,
                           int64_t* n_paths, int64_t* n_segments);

/* pm_emit_photons into device memory that also writes ONE BYTE PER PATH of this rank (path_counts_dev, in the rank's
 * path order: light by light, launch indices first_photon, first_photon + photon_stride, ...): how many photons the
 * path stored. That table is all pm_exchange_photons needs to merge the shards of several ranks without keys or a sort. */
int  pm_emit_photons_counted(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights_host,
                             const pm_light_shape* shapes_host /* NULL: point lights */, int n_lights,
                             int64_t casted, int max_depth, int caustics_mode,
                             uint64_t seed, int rng_mode, int64_t first_photon, int64_t photon_stride,
                             pm_photon40* out_dev, uint8_t* path_counts_dev, int64_t capacity, int64_t* count,
                             int64_t* n_paths, int64_t* n_segments);

/* pm_emit_photons_keyed with one pm_light_shape per light (shapes == NULL: all point lights; out_keys may be NULL) */
int  pm_emit_photons_shaped(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights_host,
                            const pm_light_shape* shapes_host, int n_lights,
                            int64_t casted, int max_depth, int caustics_mode,
                            uint64_t seed, int rng_mode, int64_t first_photon, int64_t photon_stride,
                            pm_photon40* out, uint64_t* out_keys, int64_t capacity, int64_t* count,
                            int64_t* n_paths, int64_t* n_segments);

/* pass-1 → pass-2 record conversion done by ray-tracer/src/hostCode.cu:65-83:
 * global map = non-caustic (power 1.0) followed by caustic (power 0.5); caustic map = caustic (0.5).
 * `out_global` needs n_global + n_caustic records, `out_caustic` n_caustic. Device or host pointers. */
int  pm_photons_to_maps(pm_ctx* ctx, const pm_photon40* global_in, int64_t n_global,
                        const pm_photon40* caustic_in, int64_t n_caustic,
                        pm_photon44* out_global, pm_photon44* out_caustic);

/* ---- stage 2 (≙ cukd::buildTree<Photon,Photon_traits>) ------------------------------- */
/* Reorders photons[0..n) IN PLACE into left-balanced kd-tree order (array index = node id),
 * fills split_dim, writes the AABB to bounds[6] = {lo.xyz, hi.xyz} (may be NULL), and returns a
 * map handle that owns the query-side acceleration copy. n < 1 → *out = NULL, returns PM_OK
 * (builder_thrust.h:441). photons may be host or device memory. The map keeps a pointer to
 * `photons` only if it is device memory; otherwise it owns a device copy. */
int  pm_map_build(pm_ctx* ctx, pm_photon44* photons, int64_t n, float* bounds_host, pm_map** out);
/* Sharded build of the SAME tree over `world` (a power of two) GPUs that all hold the same photons[0..n) (after the
 * all-gather of SURVEY §8e): the first log2(world) levels are computed by everybody, below them rank r continues only
 * inside subtree (world - 1) + r. pm_map_build_shard fills node_elem[n] / node_dim[n] (device memory of the caller:
 * input index and split axis of every tree node) with this rank's part and ZERO elsewhere (the common top levels
 * are rank 0's part); the caller sums the tables over the ranks (one all-reduce each, viewing node_dim as bytes
 * padded to int32) and hands them to pm_map_build_finish, which reorders photons in place, fills split_dim and
 * returns the map — byte-identical to pm_map_build's. photons is not modified by pm_map_build_shard. Trees too small
 * to shard are built whole on every rank (same calling sequence). world == 1 equals pm_map_build. */
int  pm_map_build_shard(pm_ctx* ctx, const pm_photon44* photons_dev, int64_t n, int rank, int world,
                        int32_t* node_elem_dev, uint8_t* node_dim_dev, float* bounds_host);
int  pm_map_build_finish(pm_ctx* ctx, pm_photon44* photons_dev, int64_t n, const int32_t* node_elem_dev,
                         const uint8_t* node_dim_dev, const float* bounds_host, pm_map** out);
/* test hook: pm_map_build and its sharded variants let the per-axis index lists carry their sort keys from n_min photons
 * on (the variant for maps too large for L2, 2^24 by default; same tree either way). n_min <= 0 restores the default. */
void pm_test_build_carry_from(int64_t n_min);
/* Wrap an array that is ALREADY in tree order (e.g. built by the reference builder). */
int  pm_map_from_tree(pm_ctx* ctx, const pm_photon44* photons, int64_t n, pm_map** out);
/* Cells-only map for fixed-radius gathers (progressive mode, BASELINE configs[4]; no reference counterpart): the
 * photons stay in input order, no kd-tree is built — only the query-side copies and the uniform-grid copy with the
 * cell size the caller names (about half the gather radius works well). pm_gather_radius answers every ball from the
 * cells (blocks grow to the whole grid if they must); pm_knn / pm_gather / pm_render reject such a map
 * (PM_ERR_INVALID). Builds in a fraction of pm_map_build's time: a progressive pass needs a new map per pass. */
int  pm_map_build_cells(pm_ctx* ctx, const pm_photon44* photons, int64_t n, float cell_size, pm_map** out);
void pm_map_destroy(pm_map* map);
int64_t pm_map_size(const pm_map* map);
/* test hook: which query kernels the map chose when it was built — 0: kd-tree kernels only (small or degenerate maps),
 * 1: uniform-grid general kernel, 2: fast path (uniform-like clouds, and clouds whose tuning sample passed its trial) */
int  pm_map_query_path(const pm_map* map);

/* ---- stage 3 ------------------------------------------------------------------------- */
/* kNN (≙ cukd::stackBased::knn<HeapCandidateList<k>> with cutoff radius):
 * for each query the k smallest (dist² bits, node id) among nodes with dist² < cutoff².
 * out_ids: nq*k int32, slots not filled are -1; out_dist2 (nq*k, may be NULL) holds the matching
 * squared distances (cutoff² in unfilled slots); out_r2 (nq, may be NULL) is the return value of
 * knn(): the largest dist² if k were found, else cutoff². The order of the slots within a row is unspecified (the
 * reference returns heap order); unfilled slots come last.
 * k must be one of 8, 16, 32, 50, 64. All pointers host or device. */
int  pm_knn(pm_ctx* ctx, const pm_map* map, const float* queries, int64_t nq, int k,
            float cutoff_radius, int32_t* out_ids, float* out_dist2, float* out_r2);

/* pm_knn for callers that cannot allocate. NO COUNTERPART IN THE REFERENCE: added
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). Same results as
 * pm_knn, but the call allocates nothing — no cudaMalloc / cudaFree / pool allocation, no staging: the kernels'
 * scratch (query order, bin counters, hand-over lists) is a file-scope device array of the library, 12 MiB,
 * that every GPU which loads the library holds until the process exits. Consequences:
 *   - queries_dev and every output must be DEVICE memory (host pointers: PM_ERR_INVALID);
 *   - nq <= PM_KNN_STATIC_MAX_QUERIES, else PM_ERR_CAPACITY (split the batch, or call pm_knn);
 *   - all calls on a device share the one array: calls that overlap on the GPU (two contexts, two threads) race —
 *     serialise them. The array keeps the last call's contents; each call clears what it needs at zero itself.
 * Synchronous on return like pm_knn. The first call on a device looks the array's address up (cudaGetSymbolAddress)
 * and, like any first launch, loads the kernels: make one call before a region that must not allocate. */
#define PM_KNN_STATIC_MAX_QUERIES 262144
int  pm_knn_static(pm_ctx* ctx, const pm_map* map, const float* queries_dev, int64_t nq, int k,
                   float cutoff_radius, int32_t* out_ids_dev, float* out_dist2_dev, float* out_r2_dev);

/* fused kNN + radiance estimate (≙ gatherPhotons, ray-tracer/cuda/shading.h:93-121).
 * hit: 3*nq, brdf: nq, out_rgb: 3*nq. */
int  pm_gather(pm_ctx* ctx, const pm_map* map, const float* hit, const float* brdf, int64_t nq,
               int k, float cutoff_radius, float* out_rgb);

/* fixed-radius gather: one progressive-photon-mapping pass (BASELINE config 5; the reference has only the
 * cutoff radius of its candidate list, cukd/knn.h:298-306 — this entry point is new). For every hit point i:
 * out_flux[3i..] = brdf[i] * sum(power * colour) over the nodes with dist² < radius2[i] (dist² as in pm_knn),
 * out_count[i] = their number. Hit points with brdf <= 0 or radius2 <= 0 yield zeros. Host or device pointers. */
int  pm_gather_radius(pm_ctx* ctx, const pm_map* map, const float* hit, const float* brdf, const float* radius2,
                      int64_t nq, float* out_flux, int32_t* out_count);
/* A progressive render gathers at the SAME hit points every pass. pm_gather_radius orders the hit points spatially
 * on every call (a stable sort by Morton bin: neighbouring pixels end up in one warp); pm_hit_order_create does it
 * once — bins inside the bounding box of the hit points with brdf > 0 — and pm_gather_radius_ordered reuses the
 * order, with any map and any radii, as long as hit/brdf/nq are the arrays the order was made from. Same results as
 * pm_gather_radius up to the summation order of the flux (counts are identical); the same order gives the same
 * bits on every call. Host or device pointers. */
typedef struct pm_hit_order pm_hit_order;
int  pm_hit_order_create(pm_ctx* ctx, const float* hit, const float* brdf, int64_t nq, pm_hit_order** out);
void pm_hit_order_destroy(pm_hit_order* order);
int  pm_gather_radius_ordered(pm_ctx* ctx, const pm_map* map, const pm_hit_order* order, const float* hit,
                              const float* brdf, const float* radius2, int64_t nq, float* out_flux, int32_t* out_count);

/* render (≙ simpleRayGen launch): fb is W*H RGBA8, rows as the reference writes them
 * (row H - py; row 0 stays 0, the reference's out-of-bounds row H is not written).
 * maps may be NULL (empty). Pixels [pixel_begin, pixel_end) of the launch grid (index = px + W*py)
 * are rendered by this call (tile sharding); pass 0, W*H for everything. */
int  pm_render(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights_host, int n_lights,
               const pm_map* global_map, const pm_map* caustic_map, const pm_camera* camera,
               int width, int height, int spp, int depth, const float sky[3],
               int64_t pixel_begin, int64_t pixel_end, uint32_t* fb_rgba8);

/* pm_render that also returns the colour BEFORE make_rgba quantises it (rgb_linear: 3*W*H floats indexed by launch
 * index px + W*py, only the pixels of this call are written; may be NULL): what an image RMSE against the reference
 * is computed on. Not in the reference (its only output is the 8-bit framebuffer). */
int  pm_render_rgb(pm_ctx* ctx, const pm_scene* scene, const pm_light* lights_host, int n_lights,
                   const pm_map* global_map, const pm_map* caustic_map, const pm_camera* camera,
                   int width, int height, int spp, int depth, const float sky[3],
                   int64_t pixel_begin, int64_t pixel_end, uint32_t* fb_rgba8, float* rgb_linear);

/* ---- photon viewer (≙ photonViewer: photon-viewer/src/hostCode.cu:53-75 + photon-viewer/cuda/deviceCode.cu:10-49) --
 * Splats photons into a W*H RGBA8 image: each photon is projected with the reference's glm::lookAt /
 * glm::perspective(fovy, W/H, 0.1, 1000) pipeline (pixel = int((ndc + 1) * 0.5 * size), y flipped as H - int(..);
 * clip z < 0 is dropped), tested for occlusion with a ray from the camera (tmin 1e-3, tmax = |photon - camera| - 1e-4)
 * and, if visible, its pixel receives make_rgba(color). Pixels no photon reaches keep 0xFF000000. The reference
 * leaves the winner among several photons of one pixel to a data race; here the photon with the HIGHEST index
 * wins (what a sequential run of the reference's launch would leave). photons / fb: host or device. */
int  pm_view_photons(pm_ctx* ctx, const pm_scene* scene, const pm_photon40* photons, int64_t n,
                     const float look_from[3], const float look_at[3], const float look_up[3], float fovy,
                     int width, int height, uint32_t* fb_rgba8);
/* the 4x4 projection * view matrix (column-major, glm layout) pm_view_photons uses: test hook */
void pm_view_matrix(const float look_from[3], const float look_at[3], const float look_up[3], float fovy,
                    int width, int height, float out16[16]);

/* ---- multi-GPU (the reference is single-GPU: owlContextCreate(nullptr, 1),
 * photon-mapping/src/hostCode.cu:145). One process per GPU; the collectives are NCCL calls made by the library on
 * the context's stream. libnccl.so.2 is bound at run time (PM_ERR_UNSUPPORTED if it cannot be loaded). ---------- */
typedef struct pm_comm pm_comm;
#define PM_COMM_ID_BYTES 128
#define PM_DTYPE_F32 0
#define PM_DTYPE_I32 1
#define PM_DTYPE_I64 2
/* rank 0 calls pm_comm_unique_id and hands the 128 bytes to the other ranks by whatever channel the host has (a
 * file, MPI, torch.distributed); then every rank calls pm_comm_create (ncclCommInitRank; collective). */
int  pm_comm_unique_id(void* id_out, int bytes);
int  pm_comm_create(pm_ctx* ctx, const void* id_bytes, int rank, int world, pm_comm** out);
void pm_comm_destroy(pm_comm* comm);
int  pm_comm_rank(const pm_comm* comm);
int  pm_comm_world(const pm_comm* comm);
/* in-place sum over the ranks (comm == NULL or world 1: nothing to do); synchronous on return */
int  pm_comm_allreduce_sum(pm_ctx* ctx, pm_comm* comm, void* buf_dev, int64_t count, int dtype);
/* recv_dev[r * bytes_per_rank ...] = rank r's send_dev[0 .. bytes_per_rank) */
int  pm_comm_allgather(pm_ctx* ctx, pm_comm* comm, const void* send_dev, void* recv_dev, int64_t bytes_per_rank);

/* The path's one exchange. Every rank traced its share of the SAME emission (pm_emit_photons_counted with
 * first_photon = rank, photon_stride = world, same lights / casted) and holds shard_dev[0 .. n_shard) plus one count
 * byte per path. On return out_dev[0 .. *total) holds the photons of ALL ranks as 44-byte map records (power as
 * given, split_dim 0) in exactly the order a single GPU would have produced them — (light, launch index, bounce) —
 * on every rank. Two all-gathers (counts, padded records), two prefix sums, one placement kernel; no keys, no sort.
 * collective_ms (may be NULL): device time of the all-gathers. comm == NULL: one rank (a plain 40 -> 44 conversion). */
int  pm_exchange_photons(pm_ctx* ctx, pm_comm* comm, const pm_light* lights_host, int n_lights, int64_t casted,
                         const pm_photon40* shard_dev, int64_t n_shard, const uint8_t* path_counts_dev, float power,
                         pm_photon44* out_dev, int64_t capacity, int64_t* total, float* collective_ms);
/* the placement step alone, on tables that are already gathered (test hook; also what a host with its own transport
 * calls): counts_all_dev[r * counts_pitch + p], shards_dev[r * shard_pitch + i] */
int  pm_exchange_place(pm_ctx* ctx, int world, const pm_light* lights_host, int n_lights, int64_t casted,
                       const uint8_t* counts_all_dev, int64_t counts_pitch, const pm_photon40* shards_dev,
                       int64_t shard_pitch, float power, pm_photon44* out_dev, int64_t capacity, int64_t* total);
/* pm_map_build_shard + the two node-table all-reduces + pm_map_build_finish in one call: the same tree, array and
 * map as pm_map_build on every rank (world must be a power of two, else every rank builds the whole tree) */
int  pm_map_build_sharded(pm_ctx* ctx, pm_comm* comm, pm_photon44* photons_dev, int64_t n, float* bounds_host, pm_map** out);
/* one progressive pass's state update (Hachisuka et al.): sums (flux.rgb, count) over the ranks (one all-reduce of
 * 16 bytes per hit point; skipped for one rank), then per hit point M = count, N' = N + alpha M,
 * ratio = N' / (N + M), r2 *= ratio, tau = (tau + flux) * ratio. All device pointers. */
int  pm_ppm_update(pm_ctx* ctx, pm_comm* comm, int64_t n_points, const float* flux_dev, const int32_t* count_dev,
                   float alpha, float* N_dev, float* r2_dev, float* tau_dev, float* collective_ms);

/* test hook: the gather kernels divide and take square roots with the fast-path instruction sequences of the IEEE
 * intrinsics (no slow-path call inside the warp-cooperative loops) on operands two predicates accept; this runs n
 * hashed operand pairs through both and counts accepted operands whose results differ in any bit (must be 0). */
int  pm_selftest_math(pm_ctx* ctx, int64_t n, uint32_t seed, int64_t* mismatches, int64_t* accepted, int64_t* rejected);

/* ---- photon wire format (≙ writeAlivePhotons / readPhotonsFromFile) ------------------ */
/* text: 9 floats "pos dir color", std::fixed, 6 decimals (photon-mapping/src/hostCode.cu:31-49) */
int  pm_photons_write_text(const char* path, const pm_photon40* photons_host, int64_t n);
/* reads at most capacity records; *count = number of records in the file */
int  pm_photons_read_text(const char* path, pm_photon40* photons_host, int64_t capacity, int64_t* count);

/* binary variant of the same hand-off (not in the reference: SURVEY §8f rank 1): "PMPH", version 1, int64
 * count, then the 40-byte records; lossless where the text format rounds to 6 decimals */
int  pm_photons_write_bin(const char* path, const pm_photon40* photons_host, int64_t n);
int  pm_photons_read_bin(const char* path, pm_photon40* photons_host, int64_t capacity, int64_t* count);

#ifdef __cplusplus
}
#endif
#endif /* PMHOT_H */
