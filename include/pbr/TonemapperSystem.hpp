// pbr/TonemapperSystem.hpp -- same public surface as src/pbr_engine/engine/pbr/TonemapperSystem.hpp:12-27:
//     TonemapperSystem(core::SharedGpuHandle gpu, vk::PipelineShaderStageCreateInfo shader);
//     auto allocateHdrImage(IAllocator&, vk::Extent2D) -> HdrImage;
//     auto run(vk::CommandBuffer, HdrImage const&) -> void;
// run() enqueues the ACES+gamma pass (assets/shaders/tonemappers/aces+gamma.glsl:30-40) from the HDR
// image into the output image set by HdrImage::updateOutputTexture (TonemapperSystem.cpp:97-134).
#pragma once
#include "pbr/HdrImage.hpp"

#include <stdexcept>

namespace pbr {
class TonemapperSystem {
  core::SharedGpuHandle _gpu;

public:
  TonemapperSystem(core::SharedGpuHandle gpu, vk::PipelineShaderStageCreateInfo = {}) : _gpu(std::move(gpu)) {}

  [[nodiscard]] auto allocateHdrImage(IAllocator& allocator, vk::Extent2D extent) -> HdrImage {
    return HdrImage {allocator, extent};
  }

  auto run(vk::CommandBuffer cmdBuffer, HdrImage const& hdrImage) -> void {
    auto const out = hdrImage.getOutputImage();
    if (!out) throw std::runtime_error("pbr::TonemapperSystem::run: no output image (call HdrImage::updateOutputTexture)");
    auto const hdr = hdrImage.getImage().getImage();
    auto const ext = hdrImage.getExtent();
    core::check(pbrcu_tonemap_aces_gamma(_gpu->getContext(), hdr.data, hdr.pitch, ext.width, ext.height, pbrcu_rect {0, 0, 0, 0},
                                         out.data, out.pitch, cmdBuffer.stream()),
                "TonemapperSystem::run");
  }
};
}  // namespace pbr
