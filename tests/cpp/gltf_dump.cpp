// gltf_dump ASSET -- parses a .glb / .gltf with pbr::gltf::Document (CPU only) and prints what the reference's
// gltf::Asset reads from it, in loadScene / loadNode / loadMesh / loadPrimitive / loadMaterial order, as text:
//   node <name> T x y z R x y z w S x y z mesh <index|-> children <n>
//   prim vertices <n> indices <n> vhash <fnv1a64 of the MeshVertex bytes> ihash <fnv1a64 of the uint16 indices> material <index>
//   material <name> color r g b a metallic m roughness r emissive r g b ctex <w>x<h> <fnv of RGBA8> ntex <w>x<h> <fnv>
// tests/test_gltf_cpu.py computes the same lines from the data it wrote.
#include "pbr/MeshVertex.hpp"
#include "pbr/gltf/Document.hpp"
#include "pbr/image/Decode.hpp"

#include <cinttypes>
#include <cstdio>

namespace {
auto fnv(void const* data, std::size_t n) -> std::uint64_t {
  auto const* p = static_cast<std::uint8_t const*>(data);
  std::uint64_t h = 1469598103934665603ull;
  for (std::size_t i = 0; i < n; ++i) { h ^= p[i]; h *= 1099511628211ull; }
  return h;
}
void image(pbr::gltf::Document const& doc, std::size_t textureIndex, char const* tag) {
  auto const& tex = doc.textures.at(textureIndex);
  auto const& img = doc.images.at(tex.source.value());
  std::vector<std::uint8_t> owned;
  std::span<std::uint8_t const> bytes;
  if (img.bufferView) bytes = doc.viewBytes(*img.bufferView);
  else { owned = doc.uriBytes(img.uri); bytes = owned; }
  auto const dec = pbr::image::decode(bytes);
  std::printf(" %s %ux%u %016" PRIx64, tag, dec.width, dec.height, fnv(dec.rgba.data(), dec.rgba.size()));
}
void node(pbr::gltf::Document const& doc, std::size_t index, int depth) {
  auto const& n = doc.nodes.at(index);
  auto const& t = n.transform;
  std::printf("%*snode %s T %.9g %.9g %.9g R %.9g %.9g %.9g %.9g S %.9g %.9g %.9g mesh %s children %zu\n", depth * 2, "", n.name.c_str(), t.translation[0],
              t.translation[1], t.translation[2], t.rotation[0], t.rotation[1], t.rotation[2], t.rotation[3], t.scale[0], t.scale[1], t.scale[2],
              n.mesh ? std::to_string(*n.mesh).c_str() : "-", n.children.size());
  for (auto const c : n.children) node(doc, c, depth + 1);
  if (!n.mesh) return;
  for (auto const& prim : doc.meshes.at(*n.mesh).primitives) {
    auto const need = [&](char const* name) {
      auto const a = prim.findAttribute(name);
      if (!a) throw std::runtime_error(std::string("Primitive does not have ") + name + " attribute");
      return *a;
    };
    std::vector<pbr::MeshVertex> v(doc.accessors.at(need("POSITION")).count);
    doc.iterateFloats<3>(need("POSITION"), [&](std::array<float, 3> a, std::size_t i) { v.at(i).position = {a[0], a[1], a[2]}; });
    doc.iterateFloats<3>(need("NORMAL"), [&](std::array<float, 3> a, std::size_t i) { v.at(i).normal = {a[0], a[1], a[2]}; });
    doc.iterateFloats<4>(need("TANGENT"), [&](std::array<float, 4> a, std::size_t i) { v.at(i).tangent = {a[0], a[1], a[2], a[3]}; });
    doc.iterateFloats<2>(need("TEXCOORD_0"), [&](std::array<float, 2> a, std::size_t i) { v.at(i).texCoords = {a[0], a[1]}; });
    std::vector<std::uint16_t> idx;
    if (prim.indices) doc.iterateIndices(*prim.indices, [&](std::uint16_t i) { idx.push_back(i); });
    else for (std::size_t i = 0; i < v.size(); ++i) idx.push_back(std::uint16_t(i));
    pbr::gltf::Document::requireIndicesBelow(idx, v.size());  // as gltf::Asset::loadPrimitive does
    std::printf("%*sprim vertices %zu indices %zu vhash %016" PRIx64 " ihash %016" PRIx64 " material %zu\n", depth * 2 + 2, "", v.size(), idx.size(),
                fnv(v.data(), v.size() * sizeof(pbr::MeshVertex)), fnv(idx.data(), idx.size() * 2), prim.material.value());
    auto const& m = doc.materials.at(prim.material.value());
    std::printf("%*smaterial %s color %.9g %.9g %.9g %.9g metallic %.9g roughness %.9g emissive %.9g %.9g %.9g", depth * 2 + 2, "", m.name.c_str(), m.baseColorFactor[0],
                m.baseColorFactor[1], m.baseColorFactor[2], m.baseColorFactor[3], m.metallicFactor, m.roughnessFactor, m.emissiveFactor[0], m.emissiveFactor[1], m.emissiveFactor[2]);
    image(doc, m.baseColorTexture.value().textureIndex, "ctex");
    image(doc, m.normalTexture.value().textureIndex, "ntex");
    std::printf("\n");
  }
}
}  // namespace

int main(int argc, char** argv) {
  if (argc != 2) { std::fprintf(stderr, "usage: gltf_dump ASSET\n"); return 1; }
  try {
    auto const doc = pbr::gltf::Document::fromPath(argv[1]);
    auto const& scene = doc.scenes.at(doc.defaultScene.value_or(0));
    std::printf("scene %s nodes %zu\n", scene.name.c_str(), scene.nodes.size());
    for (auto const n : scene.nodes) node(doc, n, 0);
  } catch (std::exception const& e) {
    std::fprintf(stderr, "gltf_dump: %s\n", e.what());
    return 2;
  }
  return 0;
}
