// pbr/GBuffer.hpp -- the deferred G-buffer.
// Mirrors src/pbr_engine/engine/pbr/GBuffer.hpp:11-42: four same-extent images with the formats
// below and getPositions/getNormals/getAlbedo/getDepth/getExtent.  CUDA additions (Tier B, no
// reference counterpart): optional ORM (R8G8B8A8_UNORM) and emissive (RGBA16F) planes, and raw() for
// the C ABI.  Cleared like the reference's render pass (PbrRenderSystem.cpp:370-415): colour 0, depth 1.
#pragma once
#include "pbr/Image2D.hpp"

namespace pbr {
class GBuffer {
public:
  static constexpr auto POSITIONS_FORMAT = vk::Format::eR16G16B16A16Sfloat;
  static constexpr auto NORMALS_FORMAT = vk::Format::eR16G16B16A16Sfloat;
  static constexpr auto ALBEDO_FORMAT = vk::Format::eR16G16B16A16Unorm;
  static constexpr auto DEPTH_FORMAT = vk::Format::eD32Sfloat;
  static constexpr auto ORM_FORMAT = vk::Format::eR8G8B8A8Unorm;         // Tier B
  static constexpr auto EMISSIVE_FORMAT = vk::Format::eR16G16B16A16Sfloat;  // Tier B

private:
  Image2D _positions, _normals, _albedo, _depth, _orm, _emissive;
  vk::Extent2D _extent {};
  bool _extended {};

public:
  GBuffer(IAllocator& allocator, vk::Extent2D extent, bool extended = false)
      : _positions(allocator, POSITIONS_FORMAT, extent)
      , _normals(allocator, NORMALS_FORMAT, extent)
      , _albedo(allocator, ALBEDO_FORMAT, extent)
      , _depth(allocator, DEPTH_FORMAT, extent)
      , _extent(extent)
      , _extended(extended) {
    if (extended) {
      _orm = Image2D {allocator, ORM_FORMAT, extent};
      _emissive = Image2D {allocator, EMISSIVE_FORMAT, extent};
    }
    auto const g = raw();
    core::check(pbrcu_gbuffer_clear(&g, nullptr), "pbr::GBuffer");
  }

  [[nodiscard]] auto getPositions() const noexcept -> Image2D const& { return _positions; }
  [[nodiscard]] auto getNormals() const noexcept -> Image2D const& { return _normals; }
  [[nodiscard]] auto getAlbedo() const noexcept -> Image2D const& { return _albedo; }
  [[nodiscard]] auto getDepth() const noexcept -> Image2D const& { return _depth; }
  [[nodiscard]] auto getOrm() const noexcept -> Image2D const& { return _orm; }
  [[nodiscard]] auto getEmissive() const noexcept -> Image2D const& { return _emissive; }
  [[nodiscard]] auto isExtended() const noexcept -> bool { return _extended; }
  [[nodiscard]] auto getExtent() const noexcept -> vk::Extent2D { return _extent; }
  [[nodiscard]] auto getDescriptorSet() const noexcept -> vk::DescriptorSet { return this; }

  // the C-ABI view of the planes
  [[nodiscard]] auto raw() const noexcept -> pbrcu_gbuffer {
    pbrcu_gbuffer g {};
    g.positions = _positions.getImage().data;
    g.normals = _normals.getImage().data;
    g.albedo = _albedo.getImage().data;
    g.depth = static_cast<float*>(_depth.getImage().data);
    g.orm = _orm.getImage().data;
    g.emissive = _emissive.getImage().data;
    g.width = _extent.width;
    g.height = _extent.height;
    g.pitch8 = _positions.getImage().pitch;
    g.pitch4 = _depth.getImage().pitch;
    return g;
  }
};
}  // namespace pbr
