// pbr/gltf/Loader.hpp -- mirrors src/pbr_engine/gltf/pbr/gltf/Loader.hpp:13-22 and Loader.cpp:10-32:
// loadAsset(path, dependencies) reads a .glb / .gltf (external buffers loaded, node matrices decomposed, missing
// indices generated -- the three fastgltf options the reference sets) and returns the Asset; failures throw
// std::runtime_error("Cant load gltf asset from <path> error: <why>") like the reference.
#pragma once
#include "pbr/gltf/Asset.hpp"

namespace pbr::gltf {
class Loader {
public:
  [[nodiscard]] auto loadAsset(std::filesystem::path const& path, AssetDependencies dependencies) -> Asset {
    try {
      return {Document::fromPath(path), std::move(dependencies)};
    } catch (std::exception const& e) {
      throw std::runtime_error("Cant load gltf asset from " + path.string() + " error: " + e.what());
    }
  }
};
}  // namespace pbr::gltf
