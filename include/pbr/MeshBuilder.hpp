// pbr/MeshBuilder.hpp -- mirrors src/pbr_engine/engine/pbr/MeshBuilder.hpp:13-51 and
// MeshBuilder.cpp:11-55: primitives are concatenated in insertion order; each PrimitiveSpan records the
// running firstVertex / firstIndex offsets and its own counts; indices are NOT rebased (the draw uses
// vertexOffset = firstVertex, PbrRenderSystem.cpp:458-459).
#pragma once
#include "pbr/Mesh.hpp"
#include "pbr/MeshVertex.hpp"

#include <cstdint>
#include <utility>
#include <vector>

namespace pbr {
class MeshBuilder {
public:
  struct Primitive {
    std::shared_ptr<Material> material {};
    std::vector<MeshVertex> vertices {};
    std::vector<std::uint16_t> indices {};
  };
  struct BuiltMesh {
    std::vector<MeshVertex> vertices;
    std::vector<std::uint16_t> indices;
    std::vector<PrimitiveSpan> primitives;
  };

private:
  std::vector<Primitive> _primitives {};

public:
  MeshBuilder() = default;

  auto addPrimitive(Primitive primitive) -> MeshBuilder& {
    _primitives.push_back(std::move(primitive));
    return *this;
  }
  auto addPrimitive(std::vector<MeshVertex> vertices, std::vector<std::uint16_t> indices) -> MeshBuilder& {
    Primitive p;
    p.vertices = std::move(vertices);
    p.indices = std::move(indices);
    return addPrimitive(std::move(p));
  }

  [[nodiscard]] auto build() const -> BuiltMesh {
    BuiltMesh out;
    std::size_t nv = 0, ni = 0;
    for (auto const& p : _primitives) { nv += p.vertices.size(); ni += p.indices.size(); }
    out.vertices.reserve(nv);
    out.indices.reserve(ni);
    out.primitives.reserve(_primitives.size());
    for (auto const& p : _primitives) {
      PrimitiveSpan span {};
      span.material = p.material;
      span.firstVertex = static_cast<std::uint32_t>(out.vertices.size());
      span.vertexCount = static_cast<std::uint32_t>(p.vertices.size());
      span.firstIndex = static_cast<std::uint32_t>(out.indices.size());
      span.indexCount = static_cast<std::uint32_t>(p.indices.size());
      out.vertices.insert(out.vertices.end(), p.vertices.begin(), p.vertices.end());
      out.indices.insert(out.indices.end(), p.indices.begin(), p.indices.end());
      out.primitives.push_back(std::move(span));
    }
    return out;
  }
};
}  // namespace pbr
