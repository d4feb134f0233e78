// A mesh scene through the C++ facade the way the reference's client builds and draws one:
//   MeshBuilder::addPrimitive/build -> TransferStager::addTransfer (vertex / index buffers, textures) -> Mesh,
//   Material, Node::setMesh, Scene::addNode (gltf/Asset.cpp:122-273), then per frame App.cpp:387-392:
//   pbrSystem.render(cmd, scene, gBuffer, hdr, extent)   = geometry pass (K10) + lighting pass (K1)
//   tonemapper.run(cmd, hdrImage)
// usage: facade_scene <dir>   (reads <dir>/scene.txt + the binary files it names; see tests/test_gpu_facade.py)
// writes the G-buffer planes, hdr.bin, ldr.bin and matrices.bin (view, proj, then model + normalModel per draw, as
// computed here in fp32 by makeCameraData / makeModelPushConstant) so the test can run the oracle on the same inputs.
#include "pbr/AsyncSubmitter.hpp"
#include "pbr/MeshBuilder.hpp"
#include "pbr/PbrRenderSystem.hpp"
#include "pbr/TonemapperSystem.hpp"
#include "pbr/TransferStager.hpp"

#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

namespace {
template <class T> auto readBin(std::string const& path) -> std::vector<T> {
  std::ifstream f(path, std::ios::binary | std::ios::ate);
  if (!f) throw std::runtime_error("cannot open " + path);
  std::vector<T> data(static_cast<std::size_t>(f.tellg()) / sizeof(T));
  f.seekg(0);
  f.read(reinterpret_cast<char*>(data.data()), static_cast<std::streamsize>(data.size() * sizeof(T)));
  return data;
}
template <class T> auto bytes(std::vector<T> const& v) -> std::vector<std::byte> {
  std::vector<std::byte> b(v.size() * sizeof(T));
  std::memcpy(b.data(), v.data(), b.size());
  return b;
}
void dumpImage(std::string const& path, vk::Image img, std::uint32_t w, std::uint32_t h, std::uint32_t texel, void* stream) {
  std::vector<char> host(std::size_t(w) * h * texel);
  for (std::uint32_t y = 0; y < h; ++y)
    pbr::core::check(pbrcu_memcpy_d2h(host.data() + std::size_t(y) * w * texel, static_cast<char*>(img.data) + y * img.pitch, std::size_t(w) * texel, stream), "download");
  pbr::core::check(pbrcu_stream_synchronize(stream), "sync");
  std::ofstream(path, std::ios::binary).write(host.data(), static_cast<std::streamsize>(host.size()));
}
}  // namespace

int main(int argc, char** argv) try {
  if (argc < 2) { std::fprintf(stderr, "usage: %s <dir>\n", argv[0]); return 2; }
  std::string const dir = argv[1];
  std::ifstream in(dir + "/scene.txt");
  if (!in) throw std::runtime_error("cannot open scene.txt");
  vk::Extent2D extent;
  glm::vec3 eye, target;
  float fov;
  int nMeshes;
  in >> extent.width >> extent.height >> eye.x >> eye.y >> eye.z >> target.x >> target.y >> target.z >> fov >> nMeshes;

  auto gpu = pbr::core::makeGpuHandle({});
  auto allocator = std::make_shared<pbr::MemoryAllocator>(gpu);
  pbr::PbrRenderSystem pbrSystem {gpu, {}};
  pbr::TonemapperSystem tonemapper {gpu, {}};
  auto gBuffer = pbrSystem.allocateGBuffer(*allocator, extent);
  auto hdrImage = tonemapper.allocateHdrImage(*allocator, extent);
  pbr::Image2D outputImage {*allocator, vk::Format::eR8G8B8A8Unorm, extent};
  pbrcu_stream stream {};
  pbr::core::check(pbrcu_stream_create(&stream), "stream");
  vk::CommandBuffer cmd {stream};

  pbr::TransferStager stager {gpu, allocator};
  pbr::Scene scene;
  std::vector<float> matrices;
  for (int m = 0; m < nMeshes; ++m) {
    int nPrims;
    pbr::Transform tr;
    in >> nPrims >> tr.position.x >> tr.position.y >> tr.position.z >> tr.rotation.w >> tr.rotation.x >> tr.rotation.y >> tr.rotation.z >> tr.scale.x >> tr.scale.y >> tr.scale.z;
    pbr::MeshBuilder builder;
    for (int p = 0; p < nPrims; ++p) {
      std::string prefix;
      glm::vec4 color;
      std::uint32_t cw, ch, nw, nh;
      in >> prefix >> color.x >> color.y >> color.z >> color.w >> cw >> ch >> nw >> nh;
      auto colorTex = std::make_shared<pbr::Image2D>(stager.addTransfer(bytes(readBin<std::uint8_t>(dir + "/" + prefix + "_c.bin")), vk::ImageCreateInfo {.extent = {cw, ch, 1}}));
      auto normalTex = std::make_shared<pbr::Image2D>(stager.addTransfer(bytes(readBin<std::uint8_t>(dir + "/" + prefix + "_n.bin")), vk::ImageCreateInfo {.extent = {nw, nh, 1}}));
      pbr::MeshBuilder::Primitive prim;
      prim.material = std::make_shared<pbr::Material>(pbr::Uniform<pbr::MaterialData>(*allocator, pbr::MaterialData {color}), colorTex, normalTex);
      prim.vertices = readBin<pbr::MeshVertex>(dir + "/" + prefix + "_v.bin");
      prim.indices = readBin<std::uint16_t>(dir + "/" + prefix + "_i.bin");
      builder.addPrimitive(std::move(prim));
    }
    auto built = builder.build();
    auto mesh = std::make_shared<pbr::Mesh>(stager.addTransfer(bytes(built.vertices), vk::BufferUsageFlagBits::eVertexBuffer),
                                            stager.addTransfer(bytes(built.indices), vk::BufferUsageFlagBits::eIndexBuffer), std::move(built.primitives));
    scene.addNode(pbr::Node {std::pmr::string {"mesh"}, tr}).setMesh(mesh);
    auto const pc = pbr::makeModelPushConstant(tr.position, tr.rotation, tr.scale);
    for (int p = 0; p < nPrims; ++p) {  // one draw per primitive, same matrices
      matrices.insert(matrices.end(), reinterpret_cast<float const*>(&pc.model), reinterpret_cast<float const*>(&pc.model) + 16);
      matrices.insert(matrices.end(), reinterpret_cast<float const*>(&pc.normalModel), reinterpret_cast<float const*>(&pc.normalModel) + 12);
    }
  }
  stager.submit(cmd);
  stager.wait();
  auto const camData = pbr::makeCameraData(eye, target, fov, float(extent.width) / float(extent.height));
  scene.addNode(pbr::Node {"camera"}).setCamera(std::make_shared<pbr::CameraUniform>(*allocator, camData));
  std::vector<float> head(reinterpret_cast<float const*>(&camData.view), reinterpret_cast<float const*>(&camData.view) + 16);
  head.insert(head.end(), reinterpret_cast<float const*>(&camData.proj), reinterpret_cast<float const*>(&camData.proj) + 16);
  head.insert(head.end(), matrices.begin(), matrices.end());
  std::ofstream(dir + "/matrices.bin", std::ios::binary).write(reinterpret_cast<char const*>(head.data()), static_cast<std::streamsize>(head.size() * 4));

  // ---- the reference's per-frame sequence (App.cpp:387-392) ----
  pbrSystem.render(cmd, scene, gBuffer, hdrImage.getImage(), extent);
  hdrImage.updateOutputTexture(outputImage.getImage(), outputImage.getImageView());
  tonemapper.run(cmd, hdrImage);
  pbr::AsyncSubmitter submitter {gpu};
  submitter.submit({.cmdBuffer = cmd});
  (void)submitter.wait();

  dumpImage(dir + "/pos.bin", gBuffer.getPositions().getImage(), extent.width, extent.height, 8, stream);
  dumpImage(dir + "/nrm.bin", gBuffer.getNormals().getImage(), extent.width, extent.height, 8, stream);
  dumpImage(dir + "/alb.bin", gBuffer.getAlbedo().getImage(), extent.width, extent.height, 8, stream);
  dumpImage(dir + "/depth.bin", gBuffer.getDepth().getImage(), extent.width, extent.height, 4, stream);
  dumpImage(dir + "/hdr.bin", hdrImage.getImage().getImage(), extent.width, extent.height, 8, stream);
  dumpImage(dir + "/ldr.bin", hdrImage.getOutputImage(), extent.width, extent.height, 4, stream);
  pbrcu_stream_destroy(stream);
  std::printf("facade_scene ok\n");
  return 0;
} catch (std::exception const& e) {
  std::fprintf(stderr, "facade_scene: %s\n", e.what());
  return 1;
}
