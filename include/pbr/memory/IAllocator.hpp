// pbr/memory/IAllocator.hpp -- allocation seam of the render systems.
// Mirrors src/pbr_engine/engine/pbr/memory/IAllocator.hpp:10-20 (abstract allocateBuffer /
// allocateImage) and MemoryAllocator (VMA) -- here device memory comes from pbrcu_malloc and
// host-visible ("mappable") buffers from pinned host memory.  Allocation is move-only RAII like the
// reference's (memory/Allocation.hpp:30-33).
#pragma once
#include "pbr/core/GpuHandle.hpp"

#include <cstddef>
#include <utility>

namespace pbr {
enum class AllocationPreference { Device, Host };
struct AllocationInfo {
  AllocationPreference preference {AllocationPreference::Device};
  bool ableToBeMapped {};
  bool persistentlyMapped {};
};

class Allocation {
  core::SharedGpuHandle _gpu {};
  void* _ptr {};
  std::size_t _size {};
  bool _host {};
  void (*_release)(void*) noexcept {};  // custom release (fake allocators in host-only tests)

public:
  Allocation() = default;
  Allocation(core::SharedGpuHandle gpu, void* ptr, std::size_t size, bool host)
      : _gpu(std::move(gpu)), _ptr(ptr), _size(size), _host(host) {}
  Allocation(void* ptr, std::size_t size, void (*release)(void*) noexcept)
      : _ptr(ptr), _size(size), _host(true), _release(release) {}
  Allocation(Allocation const&) = delete;
  auto operator=(Allocation const&) -> Allocation& = delete;
  Allocation(Allocation&& o) noexcept { swap(o); }
  auto operator=(Allocation&& o) noexcept -> Allocation& {
    if (this != &o) { reset(); swap(o); }
    return *this;
  }
  ~Allocation() { reset(); }

  [[nodiscard]] auto get() const noexcept -> void* { return _ptr; }
  [[nodiscard]] auto size() const noexcept -> std::size_t { return _size; }
  [[nodiscard]] auto isHost() const noexcept -> bool { return _host; }

private:
  void swap(Allocation& o) noexcept {
    std::swap(_gpu, o._gpu); std::swap(_ptr, o._ptr); std::swap(_size, o._size); std::swap(_host, o._host);
    std::swap(_release, o._release);
  }
  void reset() noexcept {
    if (_ptr) {
      if (_release) _release(_ptr);
      else if (_host) pbrcu_host_free(_ptr);
      else pbrcu_free(_gpu ? _gpu->getContext() : nullptr, _ptr);
    }
    _ptr = nullptr; _size = 0; _release = nullptr; _gpu.reset();
  }
};

class IAllocator {
public:
  virtual ~IAllocator() = default;
  // size in bytes; the reference passes vk::BufferCreateInfo{.size,...} (IAllocator.hpp:14-16)
  [[nodiscard]] virtual auto allocateBuffer(std::size_t size, AllocationInfo info) -> Allocation = 0;
  // a pitched 2-D image of `format` texels; returns the allocation and its row pitch in bytes
  [[nodiscard]] virtual auto allocateImage(vk::Extent2D extent, vk::Format format, AllocationInfo info)
      -> std::pair<Allocation, std::size_t> = 0;
};

class MemoryAllocator final : public IAllocator {
  core::SharedGpuHandle _gpu;

public:
  explicit MemoryAllocator(core::SharedGpuHandle gpu) : _gpu(std::move(gpu)) {}

  [[nodiscard]] auto allocateBuffer(std::size_t size, AllocationInfo info) -> Allocation override {
    void* p = nullptr;
    if (info.preference == AllocationPreference::Host || info.ableToBeMapped) {
      core::check(pbrcu_host_alloc(size, &p), "pbr::MemoryAllocator::allocateBuffer(host)");
      return Allocation {_gpu, p, size, true};
    }
    core::check(pbrcu_malloc(_gpu->getContext(), size, &p), "pbr::MemoryAllocator::allocateBuffer");
    return Allocation {_gpu, p, size, false};
  }
  [[nodiscard]] auto allocateImage(vk::Extent2D extent, vk::Format format, AllocationInfo)
      -> std::pair<Allocation, std::size_t> override {
    void* p = nullptr;
    std::size_t pitch = 0;
    core::check(pbrcu_image_alloc(_gpu->getContext(), extent.width, extent.height, vk::texelBytes(format), &p, &pitch),
                "pbr::MemoryAllocator::allocateImage");
    return {Allocation {_gpu, p, pitch * extent.height, false}, pitch};
  }
};
}  // namespace pbr
