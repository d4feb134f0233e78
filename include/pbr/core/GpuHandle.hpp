// pbr/core/GpuHandle.hpp -- device handle of the CUDA build.
// Mirrors src/pbr_engine/core/pbr/core/GpuHandle.hpp:12-46 (class GpuHandle, SharedGpuHandle,
// makeGpuHandle) and GpuHandleCreateInfo.hpp:14-28.  The reference picks the first Vulkan device with a
// graphics+transfer+present queue and throws std::runtime_error if none fits (GpuHandle.cpp:42-71);
// here the handle owns a pbrcu_ctx on one CUDA device and throws the same exception type.
#pragma once
#include "pbr/Vulkan.hpp"
#include "pbrcu.h"

#include <memory>
#include <span>
#include <stdexcept>
#include <string>

namespace pbr::core {
struct GpuHandleCreateInfo {
  std::span<char const* const> extensions {};  // kept for source compatibility; unused
  bool enableValidation {};                    // kept; CUDA errors are always checked
  int deviceOrdinal {0};                       // CUDA addition: which GPU (one process per GPU)
};

// Every failing C-ABI call becomes the exception the reference would throw (std::runtime_error family).
inline void check(int status, char const* what) {
  if (status != PBRCU_OK) throw std::runtime_error(std::string(what) + ": " + pbrcu_last_error());
}

class GpuHandle {
  pbrcu_ctx* _ctx {};
  int _device {};

public:
  explicit GpuHandle(GpuHandleCreateInfo const& info) : _device(info.deviceOrdinal) {
    check(pbrcu_ctx_create(info.deviceOrdinal, &_ctx), "pbr::core::GpuHandle");
  }
  GpuHandle(GpuHandle const&) = delete;
  auto operator=(GpuHandle const&) -> GpuHandle& = delete;
  ~GpuHandle() { pbrcu_ctx_destroy(_ctx); }

  [[nodiscard]] auto getContext() const noexcept -> pbrcu_ctx* { return _ctx; }
  [[nodiscard]] auto getDeviceOrdinal() const noexcept -> int { return _device; }
  [[nodiscard]] auto getSmCount() const noexcept -> int { return pbrcu_ctx_sm_count(_ctx); }
};

using SharedGpuHandle = std::shared_ptr<GpuHandle const>;

[[nodiscard]] inline auto makeGpuHandle(GpuHandleCreateInfo const& info) -> SharedGpuHandle {
  return std::make_shared<GpuHandle>(info);
}
}  // namespace pbr::core
