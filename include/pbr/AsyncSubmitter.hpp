// pbr/AsyncSubmitter.hpp -- mirrors src/pbr_engine/engine/pbr/AsyncSubmitter.hpp:17-66: one in-flight
// submission; submit() takes ownership of the info, wait() blocks until the device is done and gives
// it back; the destructor blocks.  A Vulkan queue submit + fence becomes "the stream the passes were
// enqueued on" + stream synchronisation.
#pragma once
#include "pbr/core/GpuHandle.hpp"

#include <optional>
#include <utility>

namespace pbr {
struct AsyncSubmitInfo {
  vk::CommandBuffer cmdBuffer {};
};

class AsyncSubmitter {
  core::SharedGpuHandle _gpu;
  std::optional<AsyncSubmitInfo> _submittedInfo = std::nullopt;

public:
  explicit AsyncSubmitter(core::SharedGpuHandle gpu) : _gpu(std::move(gpu)) {}
  AsyncSubmitter(AsyncSubmitter const&) = delete;
  auto operator=(AsyncSubmitter const&) -> AsyncSubmitter& = delete;
  AsyncSubmitter(AsyncSubmitter&&) = default;
  auto operator=(AsyncSubmitter&&) -> AsyncSubmitter& = default;
  ~AsyncSubmitter() noexcept {
    if (_submittedInfo) pbrcu_stream_synchronize(_submittedInfo->cmdBuffer.stream());
  }

  auto submit(AsyncSubmitInfo info) -> void { _submittedInfo = std::move(info); }  // work is already enqueued
  [[nodiscard]] auto isSubmitted() const noexcept -> bool { return _submittedInfo.has_value(); }
  auto wait() -> AsyncSubmitInfo {
    core::check(pbrcu_stream_synchronize(_submittedInfo->cmdBuffer.stream()), "AsyncSubmitter::wait");
    return std::exchange(_submittedInfo, std::nullopt).value();
  }
};
}  // namespace pbr
