// pbr/HdrImage.hpp -- the RGBA16F lighting target plus the LDR image the tonemapper writes.
// Mirrors src/pbr_engine/engine/pbr/HdrImage.hpp:11-32 (updateOutputTexture, getImage, getExtent,
// getDescriptorSet, getOutputImage).
#pragma once
#include "pbr/Image2D.hpp"

namespace pbr {
class HdrImage {
  Image2D _image;
  vk::Extent2D _extent {};
  vk::Image _outputImage {};

public:
  static constexpr auto FORMAT = vk::Format::eR16G16B16A16Sfloat;  // PbrRenderSystem.hpp:21

  HdrImage(IAllocator& allocator, vk::Extent2D extent) : _image(allocator, FORMAT, extent), _extent(extent) {}

  // the swapchain image of the frame in the reference (App.cpp:391); here any RGBA8 device image
  auto updateOutputTexture(vk::Image image, vk::ImageView) -> void { _outputImage = image; }

  [[nodiscard]] auto getImage() const noexcept -> Image2D const& { return _image; }
  [[nodiscard]] auto getExtent() const noexcept -> vk::Extent2D { return _extent; }
  [[nodiscard]] auto getDescriptorSet() const noexcept -> vk::DescriptorSet { return this; }
  [[nodiscard]] auto getOutputImage() const noexcept -> vk::Image { return _outputImage; }
};
}  // namespace pbr
