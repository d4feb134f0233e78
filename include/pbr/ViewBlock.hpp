// pbr/ViewBlock.hpp -- the G-buffers and the shaded frames of a block of views of ONE scene (CUDA addition; no reference counterpart:
// the reference renders one view per frame into one G-buffer, src/pbr_engine/engine/pbr/PbrRenderSystem.cpp:357-365, and its only client
// loops frames, src/gltf_viewer/App.cpp:444-483).  BASELINE configs[4] asks for 256 camera views of one scene; PbrRenderSystem::renderViews
// fills a ViewBlock with one batched geometry pass and one batched lighting pass.
//
// Layout: every plane of view v starts viewStride8() (8-byte texels: positions, normals, albedo, emissive) or viewStride4() (depth, ORM)
// bytes after view v-1's, rows tightly packed -- what the batched entry points of pbrcu.h take.  The shaded frames are RGBA16F,
// frameStride() bytes apart.
#pragma once
#include "pbr/memory/IAllocator.hpp"

#include <cstdint>
#include <stdexcept>

namespace pbr {
class ViewBlock {
  Allocation _positions, _normals, _albedo, _depth, _orm, _emissive, _frames;
  vk::Extent2D _extent {};
  std::uint32_t _views {};
  bool _extended {};

public:
  ViewBlock() = default;
  ViewBlock(IAllocator& allocator, vk::Extent2D extent, std::uint32_t views, bool extended) : _extent(extent), _views(views), _extended(extended) {
    if (views == 0 || extent.width == 0 || extent.height == 0) throw std::runtime_error("pbr::ViewBlock: empty block");
    auto const n8 = viewStride8() * views, n4 = viewStride4() * views;
    _positions = allocator.allocateBuffer(n8, {});
    _normals = allocator.allocateBuffer(n8, {});
    _albedo = allocator.allocateBuffer(n8, {});
    _depth = allocator.allocateBuffer(n4, {});
    if (extended) {
      _orm = allocator.allocateBuffer(n4, {});
      _emissive = allocator.allocateBuffer(n8, {});
    }
    _frames = allocator.allocateBuffer(n8, {});
  }

  [[nodiscard]] auto getExtent() const noexcept -> vk::Extent2D { return _extent; }
  [[nodiscard]] auto getViewCount() const noexcept -> std::uint32_t { return _views; }
  [[nodiscard]] auto isExtended() const noexcept -> bool { return _extended; }
  [[nodiscard]] auto pitch8() const noexcept -> std::size_t { return std::size_t(_extent.width) * 8; }
  [[nodiscard]] auto pitch4() const noexcept -> std::size_t { return std::size_t(_extent.width) * 4; }
  [[nodiscard]] auto viewStride8() const noexcept -> std::size_t { return pitch8() * _extent.height; }
  [[nodiscard]] auto viewStride4() const noexcept -> std::size_t { return pitch4() * _extent.height; }
  [[nodiscard]] auto frameStride() const noexcept -> std::size_t { return viewStride8(); }
  [[nodiscard]] auto getFrames() const noexcept -> void* { return _frames.get(); }
  [[nodiscard]] auto getFrame(std::uint32_t view) const noexcept -> vk::Image {
    return vk::Image {static_cast<char*>(_frames.get()) + frameStride() * view, pitch8()};
  }

  // the C-ABI view of view `view`'s planes (view 0: what the batched entry points take as g0)
  [[nodiscard]] auto raw(std::uint32_t view = 0) const noexcept -> pbrcu_gbuffer {
    auto const at = [view](Allocation const& a, std::size_t stride) -> void* { return a.get() ? static_cast<char*>(a.get()) + stride * view : nullptr; };
    pbrcu_gbuffer g {};
    g.positions = at(_positions, viewStride8());
    g.normals = at(_normals, viewStride8());
    g.albedo = at(_albedo, viewStride8());
    g.depth = static_cast<float*>(at(_depth, viewStride4()));
    g.orm = at(_orm, viewStride4());
    g.emissive = at(_emissive, viewStride8());
    g.width = _extent.width;
    g.height = _extent.height;
    g.pitch8 = pitch8();
    g.pitch4 = pitch4();
    return g;
  }
};
}  // namespace pbr
