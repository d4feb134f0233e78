// pbr/Vulkan.hpp -- stand-in for the reference's Vulkan-Hpp umbrella header
// (src/pbr_engine/vulkan_include/pbr/Vulkan.hpp:1-8).
//
// The kept pbr:: signatures mention a few vk:: types (render(vk::CommandBuffer, ..., vk::Extent2D),
// vk::Format constants on GBuffer, vk::Image / vk::ImageView handed to HdrImage::updateOutputTexture).
// There is no graphics stack on the B200 boxes, so this header defines exactly those names as thin
// CUDA-side equivalents, keeping call sites textually identical to the reference's:
//
//   vk::CommandBuffer  -> a CUDA stream (work recorded into it is enqueued, i.e. asynchronous, like a
//                         recorded-then-submitted command buffer)
//   vk::Image          -> device pointer + row pitch of a headless image
//   vk::ImageView      -> the same image seen with a format
//   vk::Extent2D / vk::Rect2D / vk::Format : plain data
#pragma once
#include <cstddef>
#include <cstdint>

namespace vk {
struct Extent2D {
  std::uint32_t width {};
  std::uint32_t height {};
  constexpr bool operator==(Extent2D const&) const = default;
};
struct Offset2D {
  std::int32_t x {};
  std::int32_t y {};
};
struct Rect2D {
  Offset2D offset {};
  Extent2D extent {};
};
enum class Format : std::uint32_t {  // numeric values are Vulkan's VkFormat
  eUndefined = 0,
  eR8G8B8A8Unorm = 37,
  eR16G16B16A16Unorm = 91,
  eR16G16B16A16Sfloat = 97,
  eR32G32B32A32Sfloat = 109,
  eD32Sfloat = 126,
};
constexpr std::uint32_t texelBytes(Format f) noexcept {
  switch (f) {
    case Format::eR8G8B8A8Unorm: return 4;
    case Format::eD32Sfloat: return 4;
    case Format::eR16G16B16A16Unorm: return 8;
    case Format::eR16G16B16A16Sfloat: return 8;
    case Format::eR32G32B32A32Sfloat: return 16;
    default: return 0;
  }
}
// A recorded command buffer becomes "work enqueued on this stream" (cudaStream_t as void*).
class CommandBuffer {
  void* _stream {};

public:
  constexpr CommandBuffer() = default;
  constexpr explicit CommandBuffer(void* cudaStream) : _stream(cudaStream) {}
  [[nodiscard]] constexpr void* stream() const noexcept { return _stream; }
};
// Headless image handle: device memory, row pitch in bytes.
struct Image {
  void* data {};
  std::size_t pitch {};
  constexpr explicit operator bool() const noexcept { return data != nullptr; }
  constexpr bool operator==(Image const&) const = default;
};
struct ImageView {
  Image image {};
  Format format {Format::eUndefined};
  constexpr explicit operator bool() const noexcept { return bool(image); }
};
// The reference hands SPIR-V modules to the render systems through this struct
// (PbrRenderSystem.hpp:13-18).  The CUDA kernels are linked into libpbrcu.so, so it carries nothing;
// it is kept so that construction code reads the same.
struct PipelineShaderStageCreateInfo {
  char const* pName {"main"};
};
using DescriptorSet = void const*;  // opaque; nothing to bind on the CUDA side
}  // namespace vk
