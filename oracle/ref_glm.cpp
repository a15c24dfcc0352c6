/* ref_glm.cpp — reference-derived checker (TEST INFRASTRUCTURE; built only where /root/reference exists, into
 * oracle/_ref/, never committed): the projection of photon-viewer/src/hostCode.cu:53-75 executed with the
 * reference's own glm (externals/glm, compiled from where it lies). Pins orc_view_project / pm_view_matrix. */
#include <cstdint>
#include "glm/glm.hpp"
#include "glm/ext/matrix_transform.hpp"
#include "glm/ext/matrix_clip_space.hpp"

extern "C" void ref_glm_matrix(const float* from, const float* at, const float* up, float fovy, int W, int H, float* out16) {
  auto viewMatrix = glm::lookAt(glm::vec3(from[0], from[1], from[2]), glm::vec3(at[0], at[1], at[2]), glm::vec3(up[0], up[1], up[2]));
  auto perspectiveMatrix = glm::perspective(fovy, W / static_cast<float>(H), 0.1f, 1000.f);
  auto projectionMatrix = perspectiveMatrix * viewMatrix;
  for (int c = 0; c < 4; c++)
    for (int r = 0; r < 4; r++) out16[4 * c + r] = projectionMatrix[c][r];
}

extern "C" void ref_glm_project(const float* from, const float* at, const float* up, float fovy, int W, int H,
                                const float* pos, int64_t n, int32_t* pix) {
  auto viewMatrix = glm::lookAt(glm::vec3(from[0], from[1], from[2]), glm::vec3(at[0], at[1], at[2]), glm::vec3(up[0], up[1], up[2]));
  auto perspectiveMatrix = glm::perspective(fovy, W / static_cast<float>(H), 0.1f, 1000.f);
  auto projectionMatrix = perspectiveMatrix * viewMatrix;
  for (int64_t i = 0; i < n; i++) {
    auto screenPos = projectionMatrix * glm::vec4(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2], 1.f);
    if (screenPos.z < 0) {
      pix[2 * i] = -1; pix[2 * i + 1] = -1;
    } else {
      pix[2 * i] = static_cast<int>((screenPos.x / screenPos.w + 1.f) * 0.5f * W);
      pix[2 * i + 1] = H - static_cast<int>((screenPos.y / screenPos.w + 1.f) * 0.5f * H);
    }
  }
}
