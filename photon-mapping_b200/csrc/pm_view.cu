/* pm_view.cu — the photon viewer as a CUDA splat (no graphics API involved).
 *
 * Replaces photonViewer: photon-viewer/src/hostCode.cu:53-75 (loadPhotons: glm::lookAt, glm::perspective,
 * pixel coordinates per photon) and photon-viewer/cuda/deviceCode.cu:10-49 (photonViewerRayGen: bounds check,
 * occlusion ray from the camera, frameBuffer[pixel] = make_rgba(color)).
 *
 * The projection matrix is computed on the host in float, operation by operation as glm does
 * (externals/glm/glm/ext/matrix_transform.inl:153-173, ext/matrix_clip_space.inl:249-262,
 * detail/type_mat4x4.inl:660-705 for the product); the per-photon matrix * vector product
 * (type_mat4x4.inl:562-573: (m0 x + m1 y) + (m2 z + m3 w)) and everything after it run on the device, where
 * the reference did them on the host — one thread per photon, the occlusion test through the LBVH.
 *
 * Several photons can land on one pixel; the reference lets the stores race. Here a 64-bit atomicMax on
 * (photon index + 1) << 32 | rgba makes the highest photon index win, which is what the reference's launch
 * leaves when it runs in index order; a second kernel unpacks the colours (0xFF000000 where nothing landed).
 */
#include <cmath>
#include "pm_scene.cuh"

namespace {

using namespace pmd;

struct Mat4 { float c[4][4]; }; /* c[column][row], glm layout */

constexpr float kViewEps = 1e-4f; /* photon-viewer/cuda/deviceCode.cu:8 */
constexpr uint32_t kBlack = 0xFF000000u;

__device__ __forceinline__ uint32_t make_8bit(float f) { return (uint32_t)min(255, max(0, __float2int_rz(f * 256.f))); }

__global__ void k_view_splat(SceneView sc, const pm_photon40* __restrict__ photons, int64_t n, Mat4 m, vec3 cam, int W, int H,
                             unsigned long long* __restrict__ fb64) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* p = reinterpret_cast<const float*>(photons + i);
  const float x = p[0], y = p[1], z = p[2];
  /* clip = M * (x, y, z, 1): (m0 x + m1 y) + (m2 z + m3 1) per component */
  float clip[4];
#pragma unroll
  for (int r = 0; r < 4; r++) clip[r] = (m.c[0][r] * x + m.c[1][r] * y) + (m.c[2][r] * z + m.c[3][r] * 1.f);
  if (clip[2] < 0.f) return; /* hostCode.cu:65-67: pixel = (-1, -1), dropped by the bounds check */
  /* static_cast<int> of the reference is cvt.rzi here (saturating, NaN -> 0) */
  const int px = __float2int_rz((__fdiv_rn(clip[0], clip[3]) + 1.f) * 0.5f * (float)W);
  const int py = H - __float2int_rz((__fdiv_rn(clip[1], clip[3]) + 1.f) * 0.5f * (float)H);
  if (px < 0 || px >= W || py < 0 || py >= H) return;
  const vec3 d = v3(x, y, z) - cam;
  const float tmax = norm3df(d.x, d.y, d.z) - kViewEps;
  if (trace<true>(sc, cam, normalize(d), 1e-3f, tmax).prim >= 0) return;
  const uint32_t rgba = (make_8bit(p[7]) << 0) + (make_8bit(p[8]) << 8) + (make_8bit(p[9]) << 16) + (0xffu << 24);
  atomicMax(fb64 + ((size_t)px + (size_t)W * py), ((unsigned long long)(i + 1) << 32) | rgba);
}

__global__ void k_view_resolve(const unsigned long long* __restrict__ fb64, int64_t n, uint32_t* __restrict__ fb) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long v = fb64[i];
  fb[i] = v ? (uint32_t)v : kBlack;
}

/* ---- glm restated (host, float, one rounding per written operation) ---- */
struct H3 { float x, y, z; };
H3 h_sub(H3 a, H3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
float h_dot(H3 a, H3 b) { const float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z; return tx + ty + tz; }
H3 h_cross(H3 a, H3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
H3 h_normalize(H3 v) { const float s = 1.f / std::sqrt(h_dot(v, v)); return {v.x * s, v.y * s, v.z * s}; }

void view_matrix(const float from[3], const float at[3], const float up[3], float fovy, int W, int H, Mat4* out) {
  const H3 eye{from[0], from[1], from[2]}, center{at[0], at[1], at[2]}, upv{up[0], up[1], up[2]};
  const H3 f = h_normalize(h_sub(center, eye));
  const H3 s = h_normalize(h_cross(f, upv));
  const H3 u = h_cross(s, f);
  float V[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
  V[0][0] = s.x; V[1][0] = s.y; V[2][0] = s.z;
  V[0][1] = u.x; V[1][1] = u.y; V[2][1] = u.z;
  V[0][2] = -f.x; V[1][2] = -f.y; V[2][2] = -f.z;
  V[3][0] = -h_dot(s, eye); V[3][1] = -h_dot(u, eye); V[3][2] = h_dot(f, eye);
  const float aspect = (float)W / static_cast<float>(H), zNear = 0.1f, zFar = 1000.f;
  const float tanHalf = std::tan(fovy / 2.f);
  float P[4][4] = {};
  P[0][0] = 1.f / (aspect * tanHalf);
  P[1][1] = 1.f / tanHalf;
  P[2][2] = -(zFar + zNear) / (zFar - zNear);
  P[2][3] = -1.f;
  P[3][2] = -(2.f * zFar * zNear) / (zFar - zNear);
  /* P * V, column by column: tmp = A0 b.x; tmp += A1 b.y; tmp += A2 b.z; tmp += A3 b.w */
  for (int c = 0; c < 4; c++)
    for (int r = 0; r < 4; r++) {
      float t = P[0][r] * V[c][0];
      t += P[1][r] * V[c][1];
      t += P[2][r] * V[c][2];
      t += P[3][r] * V[c][3];
      out->c[c][r] = t;
    }
}

}  // namespace

extern "C" {

void pm_view_matrix(const float look_from[3], const float look_at[3], const float look_up[3], float fovy, int width,
                    int height, float out16[16]) {
  Mat4 m;
  view_matrix(look_from, look_at, look_up, fovy, width, height, &m);
  memcpy(out16, m.c, sizeof(float) * 16);
}

int pm_view_photons(pm_ctx* ctx, const pm_scene* scene, const pm_photon40* photons, int64_t n, const float look_from[3],
                    const float look_at[3], const float look_up[3], float fovy, int width, int height, uint32_t* fb) {
  PM_REQUIRE(ctx && scene && fb, "ctx/scene/fb is NULL");
  PM_REQUIRE(look_from && look_at && look_up, "camera vectors are NULL");
  PM_REQUIRE(n >= 0 && n < ((int64_t)1 << 32) - 1, "n out of range");
  PM_REQUIRE(width > 0 && height > 0 && (int64_t)width * height < ((int64_t)1 << 31), "bad framebuffer size");
  PM_REQUIRE(n == 0 || photons, "photons is NULL");
  PM_CUDA(cudaSetDevice(ctx->device));
  const int64_t npix = (int64_t)width * height;
  Mat4 m;
  view_matrix(look_from, look_at, look_up, fovy, width, height, &m);
  const void* ph_dev = nullptr;
  if (n > 0) PM_CHECK(pm::stage_in(ctx, pm::S_STAGE_IN0, photons, sizeof(pm_photon40) * (size_t)n, &ph_dev));
  pm::OutStage o;
  PM_CHECK(pm::stage_out(ctx, pm::S_STAGE_OUT0, fb, sizeof(uint32_t) * (size_t)npix, &o));
  unsigned long long* fb64 = nullptr;
  PM_CHECK(ctx->get_scratch(pm::S_MISC0, (size_t)npix, &fb64));
  PM_CUDA(cudaMemsetAsync(fb64, 0, sizeof(unsigned long long) * (size_t)npix, ctx->stream));
  PM_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  if (n > 0) {
    k_view_splat<<<(unsigned)pm::div_up(n, 128), 128, 0, ctx->stream>>>(scene->view(), (const pm_photon40*)ph_dev, n, m,
                                                                      vec3{look_from[0], look_from[1], look_from[2]}, width, height, fb64);
    ctx->launches++;
  }
  k_view_resolve<<<(unsigned)pm::div_up(npix, 256), 256, 0, ctx->stream>>>(fb64, npix, (uint32_t*)o.dev);
  ctx->launches++;
  PM_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
  PM_CUDA(cudaGetLastError());
  PM_CHECK(pm::stage_out_finish(ctx, o));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  PM_CUDA(cudaEventElapsedTime(&ctx->last_kernel_ms, ctx->ev0, ctx->ev1));
  return PM_OK;
}

}  // extern "C"
