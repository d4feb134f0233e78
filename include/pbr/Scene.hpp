// pbr/Scene.hpp -- scene tree kept from the reference (src/pbr_engine/engine/pbr/Scene.hpp:19-116,
// Scene.cpp:13-109): Transform {position, rotation, scale}, Node (name, children, transform, optional
// mesh and camera) with pmr allocators, Scene (top-level nodes).  Behaviour that callers rely on and
// tests pin:
//   * iterateChildren()/iterateAllNodes() walk POST-ORDER: all descendants of a node are yielded
//     before the node itself (Scene.cpp:56-61, 77-82);
//   * findCamera() returns the first camera in that order, or nullopt (Scene.cpp:97-104);
//   * addChild()/addNode() return a reference to the stored node and invalidate earlier references.
// CUDA additions (Tier B, no reference counterpart): point lights and an environment on the Scene.
#pragma once
#include "pbr/CameraUniform.hpp"
#include "pbr/Generator.hpp"
#include "pbr/Mesh.hpp"

#include <memory>
#include <memory_resource>
#include <optional>
#include <span>
#include <string>
#include <vector>

namespace pbr {
struct Transform {
  glm::vec3 position {};
  glm::quat rotation = glm::identity<glm::quat>();
  glm::vec3 scale {1.0f};
};

class Node {
  std::pmr::string _name;
  std::pmr::vector<Node> _children;
  Transform _transform {};
  std::shared_ptr<Mesh> _mesh = nullptr;
  std::shared_ptr<CameraUniform> _camera = nullptr;

public:
  using allocator_type = std::pmr::polymorphic_allocator<>;

  explicit Node(std::pmr::string name, allocator_type alloc = {}) : _name(std::move(name), alloc), _children(alloc) {}
  explicit Node(std::pmr::string name, Transform transform, allocator_type alloc = {}) : Node(std::move(name), alloc) {
    _transform = transform;
  }
  Node(Node const&) = delete;
  auto operator=(Node const&) -> Node& = delete;
  Node(Node&& node, allocator_type alloc = {}) noexcept
      : _name(std::move(node._name), alloc)
      , _children(std::move(node._children), alloc)
      , _transform(node._transform)
      , _mesh(std::move(node._mesh))
      , _camera(std::move(node._camera)) {}
  auto operator=(Node&&) -> Node& = default;
  ~Node() = default;

  [[nodiscard]] auto getName() const noexcept -> std::pmr::string const& { return _name; }
  [[nodiscard]] auto getChildren() const noexcept -> std::span<Node const> { return _children; }
  [[nodiscard]] auto getChildren() noexcept -> std::span<Node> { return _children; }
  [[nodiscard]] auto getTransform() const noexcept -> Transform { return _transform; }
  auto setTransform(Transform t) noexcept -> void { _transform = t; }
  [[nodiscard]] auto getMesh() const noexcept -> std::shared_ptr<Mesh> const& { return _mesh; }
  auto setMesh(std::shared_ptr<Mesh> m) noexcept -> void { _mesh = std::move(m); }
  [[nodiscard]] auto getCamera() const noexcept -> std::shared_ptr<CameraUniform> const& { return _camera; }
  auto setCamera(std::shared_ptr<CameraUniform> c) noexcept -> void { _camera = std::move(c); }

  [[nodiscard]] auto iterateChildren() const -> generator<Node const&> {
    for (auto const& child : _children) {
      for (auto const& below : child.iterateChildren()) co_yield below;
      co_yield child;
    }
  }
  [[nodiscard]] auto iterateChildren() -> generator<Node&> {
    for (auto& child : _children) {
      for (auto& below : child.iterateChildren()) co_yield below;
      co_yield child;
    }
  }
  auto addChild(Node node) -> Node& {
    _children.emplace_back(std::move(node));
    return _children.back();
  }
};

// ---- Tier-B scene data (no reference counterpart: Scene.hpp:24-29 holds only mesh + camera) ----
struct PointLight {  // KHR_lights_punctual point light; layout == pbrcu_light
  glm::vec3 position {};
  float range {0.0f};  // <= 0: unlimited
  glm::vec3 color {1.0f};
  float intensity {1.0f};
};
static_assert(sizeof(PointLight) == sizeof(pbrcu_light), "PointLight must match pbrcu_light");

class Scene {
  std::pmr::vector<Node> _nodes;
  std::vector<PointLight> _lights;

public:
  using allocator_type = std::pmr::polymorphic_allocator<>;

  explicit Scene(allocator_type alloc = {}) : _nodes(alloc) {}

  [[nodiscard]] auto iterateAllNodes() const -> generator<Node const&> {
    for (auto const& node : _nodes) {
      for (auto const& below : node.iterateChildren()) co_yield below;
      co_yield node;
    }
  }
  [[nodiscard]] auto iterateAllNodes() -> generator<Node&> {
    for (auto& node : _nodes) {
      for (auto& below : node.iterateChildren()) co_yield below;
      co_yield node;
    }
  }
  [[nodiscard]] auto getTopLevelNodes() const noexcept -> std::span<Node const> { return _nodes; }
  [[nodiscard]] auto getTopLevelNodes() noexcept -> std::span<Node> { return _nodes; }

  [[nodiscard]] auto findCamera() const -> std::optional<std::shared_ptr<CameraUniform>> {
    for (auto const& node : iterateAllNodes())
      if (auto const& camera = node.getCamera(); camera) return camera;
    return std::nullopt;
  }
  auto addNode(Node node) -> Node& {
    _nodes.emplace_back(std::move(node));
    return _nodes.back();
  }

  auto addLight(PointLight light) -> void { _lights.push_back(light); }
  [[nodiscard]] auto getLights() const noexcept -> std::span<PointLight const> { return _lights; }
};
}  // namespace pbr
