// pbr/gltf/Asset.hpp -- a loaded glTF asset and its conversion into the engine's Scene / Node / Mesh / Material.
// Mirrors src/pbr_engine/gltf/pbr/gltf/Asset.hpp:25-78 and Asset.cpp:99-273 method by method:
//   loadImage2D  (Asset.cpp:118-129)  embedded bufferView or uri -> image::loadImage2D, cached BY NAME
//   loadMaterial (Asset.cpp:131-156)  baseColorFactor -> MaterialData.color; base-colour and normal textures are
//                                     REQUIRED (`.value()` there: std::bad_optional_access; here std::runtime_error)
//   loadPrimitive(Asset.cpp:158-204)  POSITION / NORMAL / TANGENT / TEXCOORD_0 -> MeshVertex, indices -> uint16
//   loadMesh     (Asset.cpp:206-232)  MeshBuilder over the primitives, vertex + index buffer through the stager
//   loadNode     (Asset.cpp:234-256)  TRS (quaternion re-ordered w, x, y, z), children first, then the mesh
//   loadScene    (Asset.cpp:258-273)  a "DefaultCamera" node with a CameraUniform, then the scene's root nodes
// Like the reference, the caches are keyed by the glTF object's NAME (Asset.cpp:104,121,134,209): two images,
// materials or meshes with the same name -- e.g. all unnamed ones -- resolve to whichever was loaded first.
// Vulkan-only pieces have no counterpart: samplers (the reference creates every sampler with default parameters =
// nearest + repeat, Asset.cpp:108-109, which is what K10 implements) and the descriptor-set allocators.
// Tier B: metallicFactor / roughnessFactor / emissiveFactor are carried into Material::setMetallicRoughness.
#pragma once
#include "pbr/MeshBuilder.hpp"
#include "pbr/Scene.hpp"
#include "pbr/TransferStager.hpp"
#include "pbr/gltf/Document.hpp"
#include "pbr/image/LoadImage.hpp"

#include <memory>
#include <memory_resource>
#include <string>
#include <unordered_map>

namespace pbr::gltf {
struct AssetDependencies {
  core::SharedGpuHandle gpu;
  std::shared_ptr<IAllocator> allocator;
};

class Asset {
  AssetDependencies _dependencies;
  Document _asset;

  template <typename T> using Cache = std::pmr::unordered_map<std::pmr::string, std::shared_ptr<T>>;
  Cache<Image2D> _imageCache;
  Cache<Material> _materialCache;
  Cache<Mesh> _meshCache;

  template <typename T> static auto need(std::optional<T> const& v, char const* what) -> T const& {
    if (!v) throw std::runtime_error(std::string("pbr::gltf::Asset: ") + what);
    return *v;
  }

public:
  static constexpr std::size_t MAX_NODE_DEPTH = 256;

  Asset(Document asset, AssetDependencies dependencies) noexcept : _dependencies(std::move(dependencies)), _asset(std::move(asset)) {}

  [[nodiscard]] auto getDocument() const noexcept -> Document const& { return _asset; }

  [[nodiscard]] auto loadImage2D(TransferStager& stager, std::size_t index) -> std::shared_ptr<Image2D> {
    auto const& image = _asset.images.at(index);
    if (auto iter = _imageCache.find(std::pmr::string(image.name)); iter != _imageCache.end()) return iter->second;
    std::shared_ptr<Image2D> image2D;
    if (image.bufferView) {
      image2D = std::make_shared<Image2D>(image::loadImage2D(*_dependencies.gpu, stager, _asset.viewBytes(*image.bufferView)));
    } else {
      auto const bytes = _asset.uriBytes(image.uri);
      image2D = std::make_shared<Image2D>(image::loadImage2D(*_dependencies.gpu, stager, std::span<std::uint8_t const>(bytes)));
    }
    _imageCache[std::pmr::string(image.name)] = image2D;
    return image2D;
  }

  [[nodiscard]] auto loadMaterial(TransferStager& stager, std::size_t index) -> std::shared_ptr<Material> {
    auto const& matInfo = _asset.materials.at(index);
    if (auto iter = _materialCache.find(std::pmr::string(matInfo.name)); iter != _materialCache.end()) return iter->second;
    auto const& f = matInfo.baseColorFactor;
    MaterialData const matData {.color {f[0], f[1], f[2], f[3]}};
    auto const& colorTexture = _asset.textures.at(need(matInfo.baseColorTexture, "material without a base colour texture").textureIndex);
    auto const& normalTexture = _asset.textures.at(need(matInfo.normalTexture, "material without a normal texture").textureIndex);
    (void)need(colorTexture.sampler, "texture without a sampler");  // Asset.cpp:150,152: samplerIndex.value()
    (void)need(normalTexture.sampler, "texture without a sampler");
    auto material = std::make_shared<Material>(Uniform<MaterialData>(*_dependencies.allocator, matData),
                                               loadImage2D(stager, need(colorTexture.source, "texture without an image")),
                                               loadImage2D(stager, need(normalTexture.source, "texture without an image")));
    material->setMetallicRoughness({.metallic = matInfo.metallicFactor,
                                    .roughness = matInfo.roughnessFactor,
                                    .occlusion = 1.0f,
                                    .emissive {matInfo.emissiveFactor[0], matInfo.emissiveFactor[1], matInfo.emissiveFactor[2]}});
    _materialCache[std::pmr::string(matInfo.name)] = material;
    return material;
  }

  [[nodiscard]] auto loadPrimitive(TransferStager& stager, PrimitiveInfo const& primitive) -> MeshBuilder::Primitive {
    auto const getAccessor = [&](char const* attr) {
      if (auto const acc = primitive.findAttribute(attr)) return *acc;
      throw std::runtime_error(std::string("Primitive does not have ") + attr + " attribute");
    };
    auto const positionAcc = getAccessor("POSITION"), normalAcc = getAccessor("NORMAL"), tangentAcc = getAccessor("TANGENT"),
               texCoordsAcc = getAccessor("TEXCOORD_0");
    std::vector<MeshVertex> vertices(_asset.accessors.at(positionAcc).count);
    _asset.iterateFloats<3>(positionAcc, [&](std::array<float, 3> v, std::size_t i) { vertices.at(i).position = {v[0], v[1], v[2]}; });
    _asset.iterateFloats<3>(normalAcc, [&](std::array<float, 3> v, std::size_t i) { vertices.at(i).normal = {v[0], v[1], v[2]}; });
    _asset.iterateFloats<4>(tangentAcc, [&](std::array<float, 4> v, std::size_t i) { vertices.at(i).tangent = {v[0], v[1], v[2], v[3]}; });
    _asset.iterateFloats<2>(texCoordsAcc, [&](std::array<float, 2> v, std::size_t i) { vertices.at(i).texCoords = {v[0], v[1]}; });

    std::vector<std::uint16_t> indices;
    if (primitive.indices) {
      indices.reserve(_asset.accessors.at(*primitive.indices).count);
      _asset.iterateIndices(*primitive.indices, [&](std::uint16_t idx) { indices.push_back(idx); });
    } else {  // GenerateMeshIndices (gltf/Loader.cpp:22)
      indices.resize(vertices.size());
      for (std::size_t i = 0; i < indices.size(); ++i) indices[i] = static_cast<std::uint16_t>(i);
    }
    // K10 addresses its vertex cache at vertex_offset + index: an index past the primitive's vertices (a malformed or
    // hostile file) would read another draw's vertices or run off the buffer
    Document::requireIndicesBelow(indices, vertices.size());
    MeshBuilder::Primitive out;
    out.material = loadMaterial(stager, need(primitive.material, "primitive without a material"));
    out.vertices = std::move(vertices);
    out.indices = std::move(indices);
    return out;
  }

  [[nodiscard]] auto loadMesh(TransferStager& stager, std::size_t index) -> std::shared_ptr<Mesh> {
    auto const& meshInfo = _asset.meshes.at(index);
    if (auto iter = _meshCache.find(std::pmr::string(meshInfo.name)); iter != _meshCache.end()) return iter->second;
    MeshBuilder meshBuilder;
    for (auto const& primitive : meshInfo.primitives) meshBuilder.addPrimitive(loadPrimitive(stager, primitive));
    auto builtMesh = meshBuilder.build();
    std::vector<std::byte> vertices(builtMesh.vertices.size() * sizeof(MeshVertex));
    std::memcpy(vertices.data(), builtMesh.vertices.data(), vertices.size());
    std::vector<std::byte> indices(builtMesh.indices.size() * sizeof(std::uint16_t));
    std::memcpy(indices.data(), builtMesh.indices.data(), indices.size());
    auto mesh = std::make_shared<Mesh>(stager.addTransfer(std::move(vertices), vk::BufferUsageFlagBits::eVertexBuffer),
                                       stager.addTransfer(std::move(indices), vk::BufferUsageFlagBits::eIndexBuffer), std::move(builtMesh.primitives));
    _meshCache[std::pmr::string(meshInfo.name)] = mesh;
    return mesh;
  }

  // (depth: a node graph with a cycle, or an absurdly deep one, is rejected instead of overflowing the stack)
  [[nodiscard]] auto loadNode(TransferStager& stager, std::size_t index, std::pmr::polymorphic_allocator<> alloc = {}, std::size_t depth = 0) -> Node {
    if (depth > MAX_NODE_DEPTH) throw std::runtime_error("glTF node hierarchy deeper than " + std::to_string(MAX_NODE_DEPTH) + " levels (cycle?)");
    auto const& gltfNode = _asset.nodes.at(index);
    auto const& trs = gltfNode.transform;
    Transform const transform {
        .position {trs.translation[0], trs.translation[1], trs.translation[2]},
        .rotation {trs.rotation[3], trs.rotation[0], trs.rotation[1], trs.rotation[2]},  // glm::quat(w, x, y, z)
        .scale {trs.scale[0], trs.scale[1], trs.scale[2]},
    };
    Node node(std::pmr::string(gltfNode.name, alloc), transform, alloc);
    for (auto const childIdx : gltfNode.children) node.addChild(loadNode(stager, childIdx, alloc, depth + 1));
    if (gltfNode.mesh) node.setMesh(loadMesh(stager, *gltfNode.mesh));
    return node;
  }

  [[nodiscard]] auto loadScene(TransferStager& stager, std::size_t index, std::pmr::polymorphic_allocator<> alloc = {}) -> Scene {
    Scene scene(alloc);
    scene.addNode(Node(std::pmr::string("DefaultCamera", alloc), alloc)).setCamera(std::make_shared<CameraUniform>(*_dependencies.allocator));
    auto const& gltfScene = _asset.scenes.at(index);
    for (auto const nodeIdx : gltfScene.nodes) scene.addNode(loadNode(stager, nodeIdx, alloc));
    return scene;
  }
};
}  // namespace pbr::gltf
