// pbr/Material.hpp -- mirrors src/pbr_engine/engine/pbr/Material.hpp:14-47.
// The reference's MaterialData is a single vec4 colour (Material.hpp:14-16) multiplied with the
// base-colour texel in the geometry pass (geometry_pass/fragment.glsl:29).  The metallic-roughness
// factors below are the Tier-B extension (no reference counterpart) used when an extended G-buffer
// is produced; they default to the glTF defaults.
#pragma once
#include "pbr/Image2D.hpp"
#include "pbr/Math.hpp"
#include "pbr/Uniform.hpp"

#include <memory>

namespace pbr {
struct MaterialData {
  glm::vec4 color {};
};
struct MetallicRoughnessData {  // Tier B
  float metallic {1.0f};
  float roughness {1.0f};
  float occlusion {1.0f};
  glm::vec3 emissive {};
};

class Material {
  Uniform<MaterialData> _materialData;
  std::shared_ptr<Image2D> _colorTexture;
  std::shared_ptr<Image2D> _normalTexture;
  MetallicRoughnessData _pbr {};

public:
  Material(Uniform<MaterialData> matData, std::shared_ptr<Image2D> colorTexture, std::shared_ptr<Image2D> normalTexture)
      : _materialData(std::move(matData)), _colorTexture(std::move(colorTexture)), _normalTexture(std::move(normalTexture)) {}

  [[nodiscard]] auto getMaterialData() const -> MaterialData { return _materialData.get(); }
  [[nodiscard]] auto getColorTexture() const noexcept -> std::shared_ptr<Image2D> const& { return _colorTexture; }
  [[nodiscard]] auto getNormalTexture() const noexcept -> std::shared_ptr<Image2D> const& { return _normalTexture; }
  [[nodiscard]] auto getMetallicRoughness() const noexcept -> MetallicRoughnessData { return _pbr; }
  auto setMetallicRoughness(MetallicRoughnessData d) noexcept -> void { _pbr = d; }
  [[nodiscard]] constexpr auto getDescriptorSet() const noexcept -> vk::DescriptorSet { return this; }
};
}  // namespace pbr
