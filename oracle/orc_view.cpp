/* orc_view.cpp — CPU oracle of the photon viewer. TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench cpu_baseline).
 *
 * Follows photon-viewer/src/hostCode.cu:53-75 (loadPhotons: view = glm::lookAt(from, at, up), proj =
 * glm::perspective(fovy, W / float(H), 0.1f, 1000.f), screen = (proj * view) * vec4(pos, 1), pixel.x =
 * int((sx / sw + 1) * 0.5 * W), pixel.y = H - int((sy / sw + 1) * 0.5 * H), clip z < 0 -> (-1, -1)) and
 * photon-viewer/cuda/deviceCode.cu:10-49 (bounds check, ray camera -> photon with tmax = norm3d(d) - 1e-4,
 * default tmin 1e-3 (owl_device.h ray defaults), any hit -> skip, else frameBuffer[x + W * y] = make_rgba(color)).
 * glm's arithmetic is restated from externals/glm/glm: ext/matrix_transform.inl:153-173 (lookAtRH),
 * ext/matrix_clip_space.inl:249-262 (perspectiveRH_NO), detail/type_mat4x4.inl:660-705 (mat4 * mat4, unaligned),
 * :562-573 (mat4 * vec4), detail/func_geometric.inl:48-55,73-83,98-105 (dot, cross, normalize).
 * Pinned against real glm through oracle/_ref/libglm_ref.so (tests/golden/view_ref.npz).
 * Positions on undefined behaviour: (i) the reference's stores race when several photons share a pixel — the
 * oracle lets the highest photon index win; (ii) static_cast<int> of an out-of-range float is cvt.rzi
 * (saturating, NaN -> 0); (iii) norm3d is CUDA libdevice (<= 1 ulp): the oracle rounds the double-precision
 * length, so tmax can differ by one ulp (tests allow the handful of photons that flips). */
#include <cmath>
#include <cstring>
#include <vector>
#include "orc_common.h"

namespace {
using namespace orc;

int f2i_rz(float f) {
  if (f != f) return 0;
  if (f >= 2147483648.f) return 2147483647;
  if (f <= -2147483648.f) return (-2147483647 - 1);
  return (int)f;
}
uint32_t m8(float f) { const int v = f2i_rz(f * 256.f); return (uint32_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

float gdot(vec3 a, vec3 b) { const float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z; return tx + ty + tz; }
vec3 gcross(vec3 x, vec3 y) { return v3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
vec3 gnormalize(vec3 v) { const float s = 1.f / sqrtf(gdot(v, v)); return v3(v.x * s, v.y * s, v.z * s); }

void matrix(const float* from, const float* at, const float* up, float fovy, int W, int H, float M[4][4]) {
  const vec3 eye = v3(from), f = gnormalize(v3(at) - eye), s = gnormalize(gcross(f, v3(up))), u = gcross(s, f);
  float V[4][4] = {{s.x, u.x, -f.x, 0.f}, {s.y, u.y, -f.y, 0.f}, {s.z, u.z, -f.z, 0.f},
                   {-gdot(s, eye), -gdot(u, eye), gdot(f, eye), 1.f}};
  const float aspect = W / static_cast<float>(H), zn = 0.1f, zf = 1000.f, th = tanf(fovy / 2.f);
  float P[4][4];
  memset(P, 0, sizeof P);
  P[0][0] = 1.f / (aspect * th); P[1][1] = 1.f / th; P[2][2] = -(zf + zn) / (zf - zn); P[2][3] = -1.f;
  P[3][2] = -(2.f * zf * zn) / (zf - zn);
  for (int c = 0; c < 4; c++)
    for (int r = 0; r < 4; r++) {
      float t = P[0][r] * V[c][0];
      t += P[1][r] * V[c][1]; t += P[2][r] * V[c][2]; t += P[3][r] * V[c][3];
      M[c][r] = t;
    }
}

void project(const float M[4][4], const float* pos, int W, int H, int* px, int* py) {
  float c[4];
  for (int r = 0; r < 4; r++) c[r] = (M[0][r] * pos[0] + M[1][r] * pos[1]) + (M[2][r] * pos[2] + M[3][r] * 1.f);
  if (c[2] < 0.f) { *px = -1; *py = -1; return; }
  *px = f2i_rz((c[0] / c[3] + 1.f) * 0.5f * (float)W);
  *py = H - f2i_rz((c[1] / c[3] + 1.f) * 0.5f * (float)H);
}
}  // namespace

extern "C" {

void orc_view_matrix(const float* from, const float* at, const float* up, float fovy, int W, int H, float* out16) {
  float M[4][4];
  matrix(from, at, up, fovy, W, H, M);
  memcpy(out16, M, sizeof M);
}

void orc_view_project(const float* from, const float* at, const float* up, float fovy, int W, int H, const float* pos,
                      int64_t n, int32_t* pix /* 2 n */) {
  float M[4][4];
  matrix(from, at, up, fovy, W, H, M);
  for (int64_t i = 0; i < n; i++) project(M, pos + 3 * i, W, H, &pix[2 * i], &pix[2 * i + 1]);
}

void orc_view_photons(const float* verts, int nv, const int32_t* tris, int nt, const pm_photon40* photons, int64_t n,
                      const float* from, const float* at, const float* up, float fovy, int W, int H, uint32_t* fb) {
  Scene sc{verts, nv, tris, nt, nullptr, nullptr, 0};
  float M[4][4];
  matrix(from, at, up, fovy, W, H, M);
  for (int64_t i = 0; i < (int64_t)W * H; i++) fb[i] = 0xFF000000u;
  std::vector<uint8_t> vis((size_t)n, 0);
  std::vector<int> pxs((size_t)n), pys((size_t)n);
  const vec3 cam = v3(from);
#pragma omp parallel for schedule(dynamic, 256)
  for (int64_t i = 0; i < n; i++) {
    int px, py;
    project(M, photons[i].pos, W, H, &px, &py);
    pxs[i] = px; pys[i] = py;
    if (px < 0 || px >= W || py < 0 || py >= H) continue;
    const vec3 d = v3(photons[i].pos) - cam;
    const float len = (float)std::sqrt((double)d.x * d.x + (double)d.y * d.y + (double)d.z * d.z);
    if (any_hit(sc, cam, normalize(d), 1e-3f, len - 1e-4f)) continue;
    vis[i] = 1;
  }
  for (int64_t i = 0; i < n; i++) /* index order: the last (highest) index wins */
    if (vis[i]) {
      const float* c = photons[i].color;
      fb[(size_t)pxs[i] + (size_t)W * pys[i]] = (m8(c[0]) << 0) + (m8(c[1]) << 8) + (m8(c[2]) << 16) + (0xffu << 24);
    }
}

}  // extern "C"
