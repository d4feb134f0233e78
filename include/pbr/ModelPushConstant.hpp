// pbr/ModelPushConstant.hpp -- mirrors src/pbr_engine/engine/pbr/ModelPushConstant.hpp:13-46:
// {mat4 model, mat3x4 normalModel} with normalModel = mat3(transpose(inverse(model))) padded to three
// vec4 columns, and the TRS overload model = T * R * S.
#pragma once
#include "pbr/Math.hpp"

namespace pbr {
struct ModelPushConstant {
  glm::mat4x4 model;
  glm::mat3x4 normalModel;
};
static_assert(sizeof(ModelPushConstant) == 112, "ModelPushConstant is 64 + 48 bytes");

[[nodiscard]] inline auto makeModelPushConstant(glm::mat4x4 model) -> ModelPushConstant {
  return {model, glm::mat3x4 {glm::toMat3(glm::transpose(glm::inverse(model)))}};
}
[[nodiscard]] inline auto makeModelPushConstant(glm::vec3 position, glm::quat rotation, glm::vec3 scale)
    -> ModelPushConstant {
  auto const t = glm::translate(glm::mat4 {1.0f}, position);
  auto const r = glm::toMat4(rotation);
  auto const s = glm::scale(glm::mat4 {1.0f}, scale);
  return makeModelPushConstant(t * r * s);
}
}  // namespace pbr
