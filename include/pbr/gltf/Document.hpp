// pbr/gltf/Document.hpp -- the parsed glTF 2.0 asset on the host (CPU-only): what `fastgltf::Asset` is to the
// reference's gltf::Asset (src/pbr_engine/gltf/pbr/gltf/Asset.hpp:38-39).  fastgltf 0.8.0 (cmake/Dependencies.cmake:8)
// is a third-party dependency that is absent from the reference tree, so this restates the glTF 2.0 specification
// for exactly what the reference's loader reads, with the three parser options it sets (gltf/Loader.cpp:19-22):
//   LoadExternalBuffers   -- `uri` buffers are read from disk (relative to the asset) or decoded from data: URIs;
//   DecomposeNodeMatrices -- a node `matrix` becomes translation / rotation / scale;
//   GenerateMeshIndices   -- a primitive without `indices` gets 0, 1, 2, ...
// Accessor reads follow fastgltf::iterateAccessor: any component type converts to the requested one
// (`normalized` integers by the specification's dequantisation formulas), a missing bufferView reads as zeros;
// sparse accessors are rejected.  Parity of this layer is UNPINNED against fastgltf itself (no build of it here);
// it is anchored on the reference's call sites (gltf/Asset.cpp:167-212) and on assets written by tests/gltfwriter.py.
#pragma once
#include "pbr/gltf/Json.hpp"

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <iterator>
#include <optional>
#include <span>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace pbr::gltf {

enum class ComponentType : std::uint32_t { Byte = 5120, UnsignedByte = 5121, Short = 5122, UnsignedShort = 5123, UnsignedInt = 5125, Float = 5126 };

struct BufferView {
  std::size_t buffer {}, byteOffset {}, byteLength {};
  std::optional<std::size_t> byteStride;
};
struct Accessor {
  std::optional<std::size_t> bufferView;
  std::size_t byteOffset {}, count {};
  ComponentType componentType {ComponentType::Float};
  int components {1};  // SCALAR 1, VEC2 2, VEC3 3, VEC4 4 (matrices are not read by this path)
  bool normalized {false};
};
struct ImageSource {
  std::string name;
  std::optional<std::size_t> bufferView;  // embedded (GLB) image
  std::string uri;                        // else: external file (relative to the asset) or data: URI
};
struct Texture {
  std::optional<std::size_t> sampler, source;
};
struct TextureRef {
  std::size_t textureIndex {};
};
struct MaterialInfo {
  std::string name;
  std::array<float, 4> baseColorFactor {1.0f, 1.0f, 1.0f, 1.0f};
  std::optional<TextureRef> baseColorTexture, normalTexture;
  // Tier B (not read by the reference): the metallic-roughness constants K3 shades with
  float metallicFactor {1.0f}, roughnessFactor {1.0f};
  std::array<float, 3> emissiveFactor {0.0f, 0.0f, 0.0f};
};
struct PrimitiveInfo {
  std::vector<std::pair<std::string, std::size_t>> attributes;  // semantic -> accessor
  std::optional<std::size_t> indices, material;
  [[nodiscard]] auto findAttribute(std::string_view name) const -> std::optional<std::size_t> {
    for (auto const& a : attributes)
      if (a.first == name) return a.second;
    return std::nullopt;
  }
};
struct MeshInfo {
  std::string name;
  std::vector<PrimitiveInfo> primitives;
};
struct TRS {
  std::array<float, 3> translation {0.0f, 0.0f, 0.0f};
  std::array<float, 4> rotation {0.0f, 0.0f, 0.0f, 1.0f};  // x, y, z, w (glTF order)
  std::array<float, 3> scale {1.0f, 1.0f, 1.0f};
};
struct NodeInfo {
  std::string name;
  TRS transform;
  std::vector<std::size_t> children;
  std::optional<std::size_t> mesh;
};
struct SceneInfo {
  std::string name;
  std::vector<std::size_t> nodes;
};
struct SamplerInfo {
  std::string name;
};

// Column-major 4x4 -> T, R, S.  Follows the decomposition fastgltf 0.8.0 publishes for DecomposeNodeMatrices
// (fastgltf/math.hpp decomposeTransformMatrix): translation = column 3; scale = the Euclidean lengths of columns
// 0..2; rotation from the scale-free matrix by the "Christian" matrix-to-quaternion method (euclideanspace.com):
// each component sqrt(max(0, 1 +- m00 +- m11 +- m22)) / 2, signs of x, y, z copied from the antisymmetric part.
[[nodiscard]] inline auto decomposeTransformMatrix(std::array<float, 16> m) -> TRS {
  TRS out;
  auto col = [&](int c, int r) -> float& { return m[std::size_t(4 * c + r)]; };
  out.translation = {col(3, 0), col(3, 1), col(3, 2)};
  for (int c = 0; c < 3; ++c) {
    float const len = std::sqrt(col(c, 0) * col(c, 0) + col(c, 1) * col(c, 1) + col(c, 2) * col(c, 2));
    out.scale[std::size_t(c)] = len;
    for (int r = 0; r < 3; ++r) col(c, r) /= len;
  }
  float const m00 = col(0, 0), m11 = col(1, 1), m22 = col(2, 2);
  float x = std::sqrt(std::max(0.0f, 1.0f + m00 - m11 - m22)) / 2.0f;
  float y = std::sqrt(std::max(0.0f, 1.0f - m00 + m11 - m22)) / 2.0f;
  float z = std::sqrt(std::max(0.0f, 1.0f - m00 - m11 + m22)) / 2.0f;
  float const w = std::sqrt(std::max(0.0f, 1.0f + m00 + m11 + m22)) / 2.0f;
  x = std::copysign(x, col(1, 2) - col(2, 1));
  y = std::copysign(y, col(2, 0) - col(0, 2));
  z = std::copysign(z, col(0, 1) - col(1, 0));
  out.rotation = {x, y, z, w};
  return out;
}

class Document {
public:
  std::vector<std::vector<std::uint8_t>> buffers;
  std::vector<BufferView> bufferViews;
  std::vector<Accessor> accessors;
  std::vector<ImageSource> images;
  std::vector<SamplerInfo> samplers;
  std::vector<Texture> textures;
  std::vector<MaterialInfo> materials;
  std::vector<MeshInfo> meshes;
  std::vector<NodeInfo> nodes;
  std::vector<SceneInfo> scenes;
  std::optional<std::size_t> defaultScene;
  std::filesystem::path directory;

private:
  [[noreturn]] static auto fail(std::string const& what) -> void { throw std::runtime_error("pbr::gltf: " + what); }

  static auto readFile(std::filesystem::path const& path) -> std::vector<std::uint8_t> {
    std::ifstream in(path, std::ios::binary);
    if (!in) fail("cannot open " + path.string());
    return {std::istreambuf_iterator<char>(in), std::istreambuf_iterator<char>()};
  }
  static auto base64(std::string_view text) -> std::vector<std::uint8_t> {
    std::vector<std::uint8_t> out;
    std::uint32_t acc = 0;
    int bits = 0;
    for (char const c : text) {
      int v = -1;
      if (c >= 'A' && c <= 'Z') v = c - 'A';
      else if (c >= 'a' && c <= 'z') v = c - 'a' + 26;
      else if (c >= '0' && c <= '9') v = c - '0' + 52;
      else if (c == '+' || c == '-') v = 62;
      else if (c == '/' || c == '_') v = 63;
      else if (c == '=') break;
      else continue;
      acc = (acc << 6) | std::uint32_t(v);
      bits += 6;
      if (bits >= 8) {
        bits -= 8;
        out.push_back(std::uint8_t((acc >> bits) & 0xFFu));
      }
    }
    return out;
  }
  static auto percentDecode(std::string_view uri) -> std::string {
    std::string out;
    for (std::size_t i = 0; i < uri.size(); ++i) {
      if (uri[i] == '%' && i + 2 < uri.size() + 0 && std::isxdigit(static_cast<unsigned char>(uri[i + 1])) && std::isxdigit(static_cast<unsigned char>(uri[i + 2]))) {
        out += char(std::stoi(std::string(uri.substr(i + 1, 2)), nullptr, 16));
        i += 2;
      } else {
        out += uri[i];
      }
    }
    return out;
  }
  template <std::size_t N> static auto floats(json::Value const& v, std::array<float, N>& out) -> void {
    if (v.size() != N) fail("array property has the wrong length");
    for (std::size_t i = 0; i < N; ++i) out[i] = static_cast<float>(v.at(i).number());
  }
  static auto optionalIndex(json::Value const& obj, std::string_view key) -> std::optional<std::size_t> {
    auto const* v = obj.find(key);
    return v ? std::optional<std::size_t>(v->index()) : std::nullopt;
  }

  auto build(json::Value const& root, std::vector<std::uint8_t> glbBinary, bool haveGlbBinary) -> void {
    auto const* assetInfo = root.find("asset");
    if (!assetInfo || assetInfo->stringOr("version", "").substr(0, 2) != "2.") fail("not a glTF 2.x asset");
    auto const each = [&](char const* key, auto&& fn) {
      if (auto const* arr = root.find(key)) {
        if (!arr->isArray()) fail(std::string(key) + " is not an array");
        for (auto const& item : arr->items()) fn(item);
      }
    };
    each("buffers", [&](json::Value const& b) {
      std::size_t const length = b.indexOr("byteLength", 0);
      std::vector<std::uint8_t> data;
      if (auto const* uri = b.find("uri")) {
        std::string const& u = uri->string();
        if (u.rfind("data:", 0) == 0) {
          auto const comma = u.find(',');
          if (comma == std::string::npos || u.substr(0, comma).find(";base64") == std::string::npos) fail("unsupported data: URI");
          data = base64(std::string_view(u).substr(comma + 1));
        } else {
          data = readFile(directory / percentDecode(u));
        }
      } else {
        if (!haveGlbBinary || !buffers.empty()) fail("buffer without uri (only the first buffer of a GLB may omit it)");
        data = std::move(glbBinary);
      }
      if (data.size() < length) fail("buffer is shorter than its byteLength");
      buffers.push_back(std::move(data));
    });
    each("bufferViews", [&](json::Value const& v) {
      BufferView bv;
      bv.buffer = v.indexOr("buffer", 0);
      bv.byteOffset = v.indexOr("byteOffset", 0);
      bv.byteLength = v.indexOr("byteLength", 0);
      bv.byteStride = optionalIndex(v, "byteStride");
      if (bv.buffer >= buffers.size() || bv.byteOffset > buffers[bv.buffer].size() || bv.byteLength > buffers[bv.buffer].size() - bv.byteOffset)
        fail("bufferView outside its buffer");
      bufferViews.push_back(bv);
    });
    each("accessors", [&](json::Value const& v) {
      Accessor a;
      a.bufferView = optionalIndex(v, "bufferView");
      a.byteOffset = v.indexOr("byteOffset", 0);
      a.count = v.indexOr("count", 0);
      auto const ct = v.indexOr("componentType", 0);
      if (ct != 5120 && ct != 5121 && ct != 5122 && ct != 5123 && ct != 5125 && ct != 5126) fail("unknown accessor componentType");
      a.componentType = static_cast<ComponentType>(ct);
      std::string const type = v.stringOr("type", "");
      a.components = type == "SCALAR" ? 1 : type == "VEC2" ? 2 : type == "VEC3" ? 3 : type == "VEC4" ? 4 : type == "MAT2" ? 4 : type == "MAT3" ? 9 : type == "MAT4" ? 16 : 0;
      if (a.components == 0) fail("unknown accessor type");
      if (auto const* n = v.find("normalized")) a.normalized = n->boolean();
      if (v.find("sparse")) fail("sparse accessors are not supported");
      if (a.bufferView && *a.bufferView >= bufferViews.size()) fail("accessor bufferView out of range");
      accessors.push_back(a);
    });
    each("images", [&](json::Value const& v) {
      ImageSource img;
      img.name = v.stringOr("name", "");
      img.bufferView = optionalIndex(v, "bufferView");
      img.uri = v.stringOr("uri", "");
      if (!img.bufferView && img.uri.empty()) fail("image without bufferView or uri");
      images.push_back(std::move(img));
    });
    each("samplers", [&](json::Value const& v) { samplers.push_back({v.stringOr("name", "")}); });
    each("textures", [&](json::Value const& v) { textures.push_back({optionalIndex(v, "sampler"), optionalIndex(v, "source")}); });
    each("materials", [&](json::Value const& v) {
      MaterialInfo m;
      m.name = v.stringOr("name", "");
      if (auto const* pbr = v.find("pbrMetallicRoughness")) {
        if (auto const* f = pbr->find("baseColorFactor")) floats(*f, m.baseColorFactor);
        if (auto const* t = pbr->find("baseColorTexture")) m.baseColorTexture = TextureRef {t->indexOr("index", 0)};
        m.metallicFactor = static_cast<float>(pbr->numberOr("metallicFactor", 1.0));
        m.roughnessFactor = static_cast<float>(pbr->numberOr("roughnessFactor", 1.0));
      }
      if (auto const* t = v.find("normalTexture")) m.normalTexture = TextureRef {t->indexOr("index", 0)};
      if (auto const* e = v.find("emissiveFactor")) floats(*e, m.emissiveFactor);
      materials.push_back(std::move(m));
    });
    each("meshes", [&](json::Value const& v) {
      MeshInfo mesh;
      mesh.name = v.stringOr("name", "");
      if (auto const* prims = v.find("primitives"))
        for (auto const& p : prims->items()) {
          PrimitiveInfo info;
          if (auto const* attrs = p.find("attributes"); attrs && attrs->isObject())
            for (auto const& [semantic, acc] : attrs->members()) info.attributes.emplace_back(semantic, acc.index());
          info.indices = optionalIndex(p, "indices");
          info.material = optionalIndex(p, "material");
          mesh.primitives.push_back(std::move(info));
        }
      meshes.push_back(std::move(mesh));
    });
    each("nodes", [&](json::Value const& v) {
      NodeInfo node;
      node.name = v.stringOr("name", "");
      if (auto const* m = v.find("matrix")) {  // DecomposeNodeMatrices
        std::array<float, 16> mat {};
        floats(*m, mat);
        node.transform = decomposeTransformMatrix(mat);
      } else {
        if (auto const* t = v.find("translation")) floats(*t, node.transform.translation);
        if (auto const* r = v.find("rotation")) floats(*r, node.transform.rotation);
        if (auto const* s = v.find("scale")) floats(*s, node.transform.scale);
      }
      if (auto const* c = v.find("children"))
        for (auto const& child : c->items()) node.children.push_back(child.index());
      node.mesh = optionalIndex(v, "mesh");
      nodes.push_back(std::move(node));
    });
    each("scenes", [&](json::Value const& v) {
      SceneInfo scene;
      scene.name = v.stringOr("name", "");
      if (auto const* n = v.find("nodes"))
        for (auto const& node : n->items()) scene.nodes.push_back(node.index());
      scenes.push_back(std::move(scene));
    });
    defaultScene = optionalIndex(root, "scene");
    for (auto const& n : nodes) {
      for (auto const c : n.children)
        if (c >= nodes.size()) fail("node child out of range");
      if (n.mesh && *n.mesh >= meshes.size()) fail("node mesh out of range");
    }
    for (auto const& s : scenes)
      for (auto const n : s.nodes)
        if (n >= nodes.size()) fail("scene node out of range");
    // glTF 2.0, 3.5.2: the node hierarchy is a set of strict trees.  A cycle (or a child with two parents) would send
    // the recursive loaders into unbounded recursion; reject it here, iteratively.
    std::vector<std::uint8_t> hasParent(nodes.size(), 0);
    for (auto const& n : nodes)
      for (auto const c : n.children) {
        if (hasParent[c]) fail("node has more than one parent");
        hasParent[c] = 1;
      }
    std::vector<std::uint8_t> reached(nodes.size(), 0);
    std::vector<std::size_t> stack;
    for (std::size_t r = 0; r < nodes.size(); ++r)
      if (!hasParent[r]) {
        stack.push_back(r);
        while (!stack.empty()) {
          std::size_t const i = stack.back();
          stack.pop_back();
          reached[i] = 1;
          for (auto const c : nodes[i].children) stack.push_back(c);
        }
      }
    for (std::size_t i = 0; i < nodes.size(); ++i)
      if (!reached[i]) fail("node hierarchy contains a cycle");  // every node of a forest hangs below a parentless root
  }

  // one component of one element, converted to double (exactly representable for every component type)
  [[nodiscard]] static auto component(std::uint8_t const* p, ComponentType t) -> double {
    switch (t) {
      case ComponentType::Byte: { std::int8_t v; std::memcpy(&v, p, 1); return v; }
      case ComponentType::UnsignedByte: return *p;
      case ComponentType::Short: { std::int16_t v; std::memcpy(&v, p, 2); return v; }
      case ComponentType::UnsignedShort: { std::uint16_t v; std::memcpy(&v, p, 2); return v; }
      case ComponentType::UnsignedInt: { std::uint32_t v; std::memcpy(&v, p, 4); return v; }
      default: { float v; std::memcpy(&v, p, 4); return v; }
    }
  }

public:
  [[nodiscard]] static auto componentSize(ComponentType t) -> std::size_t {
    return t == ComponentType::Byte || t == ComponentType::UnsignedByte ? 1 : t == ComponentType::Short || t == ComponentType::UnsignedShort ? 2 : 4;
  }

  // .glb (binary container) or .gltf (JSON + external / data: buffers), chosen by the first four bytes
  [[nodiscard]] static auto fromPath(std::filesystem::path const& path) -> Document {
    return fromBytes(readFile(path), path.parent_path());
  }
  [[nodiscard]] static auto fromBytes(std::span<std::uint8_t const> bytes, std::filesystem::path directory) -> Document {
    Document doc;
    doc.directory = std::move(directory);
    auto const u32 = [&](std::size_t off) { std::uint32_t v; std::memcpy(&v, bytes.data() + off, 4); return v; };  // little-endian hosts only
    if (bytes.size() >= 12 && u32(0) == 0x46546C67u) {  // "glTF"
      if (u32(4) != 2) fail("unsupported GLB container version");
      std::size_t const total = std::min<std::size_t>(u32(8), bytes.size());
      std::string jsonText;
      std::vector<std::uint8_t> bin;
      bool haveJson = false, haveBin = false;
      for (std::size_t pos = 12; pos + 8 <= total;) {
        std::size_t const len = u32(pos);
        std::uint32_t const kind = u32(pos + 4);
        if (len > total - pos - 8) fail("GLB chunk runs past the end of the file");
        if (kind == 0x4E4F534Au && !haveJson) {  // "JSON"
          jsonText.assign(reinterpret_cast<char const*>(bytes.data() + pos + 8), len);
          haveJson = true;
        } else if (kind == 0x004E4942u && !haveBin) {  // "BIN\0"
          bin.assign(bytes.begin() + std::ptrdiff_t(pos + 8), bytes.begin() + std::ptrdiff_t(pos + 8 + len));
          haveBin = true;
        }
        pos += 8 + ((len + 3) & ~std::size_t(3));
      }
      if (!haveJson) fail("GLB without a JSON chunk");
      doc.build(json::parse(jsonText), std::move(bin), haveBin);
    } else {
      doc.build(json::parse(std::string_view(reinterpret_cast<char const*>(bytes.data()), bytes.size())), {}, false);
    }
    return doc;
  }

  // every index of a primitive must address one of its vertices (K10 reads its vertex cache at vertex_offset + index)
  static auto requireIndicesBelow(std::span<std::uint16_t const> indices, std::size_t vertexCount) -> void {
    for (auto const idx : indices)
      if (idx >= vertexCount) fail("primitive index " + std::to_string(idx) + " addresses a vertex past its " + std::to_string(vertexCount) + " vertices");
  }

  // count elements of elem bytes, stride apart, starting byteOffset bytes into a view of byteLength bytes -- without
  // the size_t wrap-around a huge count or stride from an untrusted file would cause in (count - 1) * stride + elem
  [[nodiscard]] static auto accessorFits(std::size_t count, std::size_t byteOffset, std::size_t stride, std::size_t elem, std::size_t byteLength) noexcept
      -> bool {
    if (count == 0) return true;
    if (byteOffset > byteLength || elem > byteLength - byteOffset) return false;
    if (count == 1) return true;
    if (stride == 0) return false;
    return count - 1 <= (byteLength - byteOffset - elem) / stride;
  }

  // fastgltf::iterateAccessorWithIndex<vecN> for float targets: fn(std::array<float, N>, index).  The accessor must
  // have at least N components per element (extra ones are ignored, like reading a VEC4 colour as vec3).
  template <std::size_t N, typename Fn> auto iterateFloats(std::size_t accessorIndex, Fn&& fn) const -> void {
    Accessor const& a = accessors.at(accessorIndex);
    if (std::size_t(a.components) < N) fail("accessor has fewer components than the attribute needs");
    std::size_t const csize = componentSize(a.componentType);
    if (!a.bufferView) {
      for (std::size_t i = 0; i < a.count; ++i) fn(std::array<float, N> {}, i);
      return;
    }
    BufferView const& bv = bufferViews.at(*a.bufferView);
    std::size_t const elem = csize * std::size_t(a.components), stride = bv.byteStride.value_or(elem);
    if (!accessorFits(a.count, a.byteOffset, stride, elem, bv.byteLength)) fail("accessor outside its bufferView");
    std::uint8_t const* base = buffers[bv.buffer].data() + bv.byteOffset + a.byteOffset;
    for (std::size_t i = 0; i < a.count; ++i) {
      std::array<float, N> out {};
      for (std::size_t c = 0; c < N; ++c) {
        double const v = component(base + i * stride + c * csize, a.componentType);
        if (a.normalized && a.componentType != ComponentType::Float) {
          // glTF 2.0, 3.11 (accessor data types): dequantisation of normalised integers
          switch (a.componentType) {
            case ComponentType::Byte: out[c] = std::max(static_cast<float>(v) / 127.0f, -1.0f); break;
            case ComponentType::UnsignedByte: out[c] = static_cast<float>(v) / 255.0f; break;
            case ComponentType::Short: out[c] = std::max(static_cast<float>(v) / 32767.0f, -1.0f); break;
            case ComponentType::UnsignedShort: out[c] = static_cast<float>(v) / 65535.0f; break;
            default: out[c] = static_cast<float>(v) / 4294967295.0f; break;
          }
        } else {
          out[c] = static_cast<float>(v);
        }
      }
      fn(out, i);
    }
  }
  // fastgltf::iterateAccessor<std::uint16_t>: fn(value) per element; wider components are narrowed like a cast
  template <typename Fn> auto iterateIndices(std::size_t accessorIndex, Fn&& fn) const -> void {
    Accessor const& a = accessors.at(accessorIndex);
    std::size_t const csize = componentSize(a.componentType);
    if (!a.bufferView) {
      for (std::size_t i = 0; i < a.count; ++i) fn(std::uint16_t {0});
      return;
    }
    BufferView const& bv = bufferViews.at(*a.bufferView);
    std::size_t const elem = csize * std::size_t(a.components), stride = bv.byteStride.value_or(elem);
    if (!accessorFits(a.count, a.byteOffset, stride, elem, bv.byteLength)) fail("accessor outside its bufferView");
    std::uint8_t const* base = buffers[bv.buffer].data() + bv.byteOffset + a.byteOffset;
    for (std::size_t i = 0; i < a.count; ++i) {
      double const v = component(base + i * stride, a.componentType);
      fn(static_cast<std::uint16_t>(static_cast<std::uint32_t>(static_cast<std::int64_t>(v))));
    }
  }
  // the bytes of an embedded image
  [[nodiscard]] auto viewBytes(std::size_t bufferViewIndex) const -> std::span<std::uint8_t const> {
    BufferView const& bv = bufferViews.at(bufferViewIndex);
    return {buffers[bv.buffer].data() + bv.byteOffset, bv.byteLength};
  }
  // an image `uri`: the decoded payload of a data: URI, or the file it names (relative to the asset)
  [[nodiscard]] auto uriBytes(std::string const& uri) const -> std::vector<std::uint8_t> {
    if (uri.rfind("data:", 0) == 0) {
      auto const comma = uri.find(',');
      if (comma == std::string::npos || uri.substr(0, comma).find(";base64") == std::string::npos) fail("unsupported data: URI");
      return base64(std::string_view(uri).substr(comma + 1));
    }
    return readFile(directory / percentDecode(uri));
  }
};

}  // namespace pbr::gltf
