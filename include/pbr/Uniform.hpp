// pbr/Uniform.hpp -- a host-visible value read by device passes.
// Mirrors src/pbr_engine/engine/pbr/Uniform.hpp:14-63: Uniform<T>(allocator, init), get(), set(T).
// The reference keeps a persistently mapped host-coherent UBO and memcpy's into it; here the value
// lives in pinned host memory and is passed to kernels as an argument at record time.
#pragma once
#include "pbr/memory/IAllocator.hpp"

#include <concepts>
#include <cstring>
#include <memory>
#include <type_traits>

namespace pbr {
template <typename T>
concept UniformValue = std::semiregular<T> && std::is_trivially_copyable_v<T>;

template <UniformValue T> class Uniform {
  Allocation _buffer;

public:
  explicit Uniform(IAllocator& allocator, T init = {})
      : _buffer(allocator.allocateBuffer(sizeof(T), {.preference = AllocationPreference::Host,
                                                     .ableToBeMapped = true,
                                                     .persistentlyMapped = true})) {
    set(init);
  }
  [[nodiscard]] auto get() const -> T {
    T value {};
    std::memcpy(std::addressof(value), _buffer.get(), sizeof(T));
    return value;
  }
  auto set(T value) -> void { std::memcpy(_buffer.get(), std::addressof(value), sizeof(T)); }
  [[nodiscard]] auto getUniformBuffer() const noexcept -> Allocation const& { return _buffer; }
};
}  // namespace pbr
