// pbr/Mesh.hpp -- mirrors src/pbr_engine/engine/pbr/Mesh.hpp:12-46: PrimitiveSpan
// {material, firstVertex, vertexCount, firstIndex, indexCount} and Mesh {vertex buffer, index buffer
// (uint16), primitives}.  Buffers are device allocations.
#pragma once
#include "pbr/Material.hpp"
#include "pbr/memory/IAllocator.hpp"

#include <cstdint>
#include <memory>
#include <span>
#include <vector>

namespace pbr {
struct PrimitiveSpan {
  std::shared_ptr<Material> material;
  std::uint32_t firstVertex;
  std::uint32_t vertexCount;
  std::uint32_t firstIndex;
  std::uint32_t indexCount;
};
using Buffer = Allocation;

class Mesh {
  Buffer _vertexBuffer;
  Buffer _indexBuffer;
  std::vector<PrimitiveSpan> _primitives;

public:
  Mesh(Buffer vertexBuffer, Buffer indexBuffer, std::vector<PrimitiveSpan> primitives) noexcept
      : _vertexBuffer(std::move(vertexBuffer)), _indexBuffer(std::move(indexBuffer)), _primitives(std::move(primitives)) {}
  [[nodiscard]] auto getVertexBuffer() const noexcept -> Buffer const& { return _vertexBuffer; }
  [[nodiscard]] auto getIndexBuffer() const noexcept -> Buffer const& { return _indexBuffer; }
  [[nodiscard]] auto getPrimitives() const noexcept -> std::span<PrimitiveSpan const> { return _primitives; }
};
}  // namespace pbr
