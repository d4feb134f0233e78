// pbr/TransferStager.hpp -- batched host-to-device uploads of mesh buffers and textures.
// Mirrors src/pbr_engine/engine/pbr/TransferStager.hpp:18-52 and TransferStager.cpp:51-172: addTransfer()
// creates the destination (a device buffer, or a 2-D image) and remembers the bytes; submit() packs every payload
// into ONE staging buffer and records all copies; wait() blocks until they are done and drops the staging buffer.
// The Vulkan staging buffer + vkCmdCopyBuffer/vkCmdCopyBufferToImage become one pinned host buffer +
// cudaMemcpyAsync on the submit stream (pbrcu_memcpy_h2d); image layout transitions have no CUDA counterpart.
#pragma once
#include "pbr/AsyncSubmitter.hpp"
#include "pbr/Image2D.hpp"
#include "pbr/Mesh.hpp"

#include <cstddef>
#include <cstring>
#include <memory>
#include <optional>
#include <stdexcept>
#include <vector>

namespace vk {
struct Extent3D {
  std::uint32_t width {}, height {}, depth {1};
};
struct ImageCreateInfo {  // the fields the reference's loaders fill (image/LoadImage.cpp:27-47)
  Format format {Format::eR8G8B8A8Unorm};
  Extent3D extent {};
  std::uint32_t mipLevels {1};
  std::uint32_t arrayLayers {1};
};
enum class BufferUsageFlagBits : std::uint32_t { eVertexBuffer = 0x80, eIndexBuffer = 0x40 };
using BufferUsageFlags = BufferUsageFlagBits;
}  // namespace vk

namespace pbr {
class TransferStager {
  struct BufferTransfer {
    std::vector<std::byte> data;
    void* dst;
  };
  struct ImageTransfer {
    std::vector<std::byte> data;
    vk::Image image;
    std::size_t rowBytes;
    std::uint32_t rows;
  };
  core::SharedGpuHandle _gpu;
  std::shared_ptr<IAllocator> _allocator;
  std::vector<BufferTransfer> _bufferTransfers {};
  std::vector<ImageTransfer> _imageTransfers {};
  std::optional<Allocation> _stagingBuffer = std::nullopt;
  AsyncSubmitter _submitter;

public:
  TransferStager(core::SharedGpuHandle gpu, std::shared_ptr<IAllocator> allocator)
      : _gpu(gpu), _allocator(std::move(allocator)), _submitter(std::move(gpu)) {}

  [[nodiscard]] auto addTransfer(std::vector<std::byte> data, vk::BufferUsageFlags) -> Buffer {
    auto buffer = _allocator->allocateBuffer(data.size(), {});
    _bufferTransfers.push_back({std::move(data), buffer.get()});
    return buffer;
  }
  [[nodiscard]] auto addTransfer(std::vector<std::byte> data, vk::ImageCreateInfo imageInfo) -> Image2D {
    vk::Extent2D const extent {imageInfo.extent.width, imageInfo.extent.height};
    std::size_t const rowBytes = std::size_t(extent.width) * vk::texelBytes(imageInfo.format);
    if (imageInfo.extent.depth != 1 || imageInfo.mipLevels != 1 || imageInfo.arrayLayers != 1 || data.size() != rowBytes * extent.height)
      throw std::runtime_error("pbr::TransferStager::addTransfer: only tightly packed single-level 2-D images are supported");
    Image2D image {*_allocator, imageInfo.format, extent};
    _imageTransfers.push_back({std::move(data), image.getImage(), rowBytes, extent.height});
    return image;
  }

  // Records every pending copy on the command buffer's stream through one staging buffer (TransferStager.cpp:51-172).
  auto submit(vk::CommandBuffer cmdBuffer) -> void {
    if (_submitter.isSubmitted()) (void)wait();
    std::size_t total = 0;
    for (auto const& t : _bufferTransfers) total += t.data.size();
    for (auto const& t : _imageTransfers) total += t.data.size();
    if (total == 0) return;
    _stagingBuffer = _allocator->allocateBuffer(total, {.preference = AllocationPreference::Host, .ableToBeMapped = true});
    auto* staging = static_cast<std::byte*>(_stagingBuffer->get());
    std::size_t offset = 0;
    for (auto const& t : _bufferTransfers) {
      std::memcpy(staging + offset, t.data.data(), t.data.size());
      core::check(pbrcu_memcpy_h2d(t.dst, staging + offset, t.data.size(), cmdBuffer.stream()), "TransferStager: buffer copy");
      offset += t.data.size();
    }
    for (auto const& t : _imageTransfers) {
      std::memcpy(staging + offset, t.data.data(), t.data.size());
      // one pitched copy per image (the staging rows are tightly packed, the image rows pitch apart)
      core::check(pbrcu_memcpy_h2d_2d(t.image.data, t.image.pitch, staging + offset, t.rowBytes, t.rowBytes, t.rows, cmdBuffer.stream()),
                  "TransferStager: image copy");
      offset += t.data.size();
    }
    _bufferTransfers.clear();
    _imageTransfers.clear();
    _submitter.submit({.cmdBuffer = cmdBuffer});
  }
  auto wait() -> void {
    if (_submitter.isSubmitted()) (void)_submitter.wait();
    _stagingBuffer.reset();
  }
};
}  // namespace pbr
