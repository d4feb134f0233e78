// pbr/CameraUniform.hpp -- mirrors src/pbr_engine/engine/pbr/CameraUniform.hpp:13-31
// (a Uniform<CameraData>; the descriptor set has no CUDA counterpart).
#pragma once
#include "pbr/CameraData.hpp"

namespace pbr {
class CameraUniform : public Uniform<CameraData> {
public:
  explicit CameraUniform(IAllocator& allocator, CameraData init = {}) : Uniform(allocator, init) {}
  [[nodiscard]] constexpr auto getDescriptorSet() const noexcept -> vk::DescriptorSet { return this; }
};
}  // namespace pbr
