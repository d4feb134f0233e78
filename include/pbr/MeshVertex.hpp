// pbr/MeshVertex.hpp -- mirrors src/pbr_engine/engine/pbr/MeshVertex.hpp:14-27:
// {vec3 position, vec3 normal, vec4 tangent, vec2 texCoords} = 48 bytes, attribute formats
// R32G32B32, R32G32B32, R32G32B32A32, R32G32.
#pragma once
#include "pbr/Math.hpp"

namespace pbr {
struct MeshVertex {
  glm::vec3 position;
  glm::vec3 normal;
  glm::vec4 tangent;
  glm::vec2 texCoords;
};
static_assert(sizeof(MeshVertex) == 48, "MeshVertex must be 48 bytes (MeshVertex.hpp:14-19)");
}  // namespace pbr
