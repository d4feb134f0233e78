// pbr/Image2D.hpp -- a 2-D device image (one mip, one layer) with its format.
// Mirrors src/pbr_engine/engine/pbr/Image.hpp + Image2D.hpp:12-43 (getImage / getImageView); the
// image memory is a pitched linear device buffer instead of an optimal-tiled VkImage.
#pragma once
#include "pbr/memory/IAllocator.hpp"

namespace pbr {
class Image2D {
  Allocation _allocation;
  vk::Image _image {};
  vk::Format _format {vk::Format::eUndefined};
  vk::Extent2D _extent {};

public:
  Image2D() = default;
  Image2D(IAllocator& allocator, vk::Format format, vk::Extent2D extent) : _format(format), _extent(extent) {
    auto [allocation, pitch] = allocator.allocateImage(extent, format, {});
    _allocation = std::move(allocation);
    _image = vk::Image {_allocation.get(), pitch};
  }
  [[nodiscard]] auto getImage() const noexcept -> vk::Image { return _image; }
  [[nodiscard]] auto getImageView() const noexcept -> vk::ImageView { return {_image, _format}; }
  [[nodiscard]] auto getFormat() const noexcept -> vk::Format { return _format; }
  [[nodiscard]] auto getExtent() const noexcept -> vk::Extent2D { return _extent; }
};
}  // namespace pbr
