// pbr/CameraData.hpp -- camera block consumed by the passes.
// Mirrors src/pbr_engine/engine/pbr/CameraData.hpp:14-32: {mat4 view, mat4 proj, vec3 position},
// 16-byte aligned (std140 UBO of assets/shaders/Camera.lib.glsl:4-8); makeCameraData builds a
// right-handed view with up = -Y and perspectiveRH_NO(fov, aspect, 0.01, 1024).
// Only `position` reaches the lighting pass (lighting.glsl:25).
#pragma once
#include "pbr/Math.hpp"
#include "pbr/Uniform.hpp"

namespace pbr {
struct alignas(16) CameraData {
  glm::mat4x4 view {};
  glm::mat4x4 proj {};
  glm::vec3 position {};
};
static_assert(sizeof(CameraData) == 144, "CameraData must match the std140 block (144 bytes)");
static_assert(UniformValue<CameraData>, "pbr::CameraData does not satisfy pbr::UniformValue");

[[nodiscard]] inline auto makeCameraData(glm::vec3 position, glm::vec3 target, float fov, float aspect) noexcept
    -> CameraData {
  constexpr float zNear = 0.01f, zFar = 1024.0f;
  CameraData data;
  data.view = glm::lookAtRH(position, target, glm::vec3 {0.0f, -1.0f, 0.0f});
  data.proj = glm::perspectiveRH_NO(fov, aspect, zNear, zFar);
  data.position = position;
  return data;
}
}  // namespace pbr
