// One frame through the C++ facade exactly the way the reference's only client drives the systems
// (src/gltf_viewer/App.cpp:258-274 construction, :387-392 per frame):
//     pbrSystem.render(cmd, scene, gBuffer, hdrImage.getImage(), extent);
//     hdrImage.updateOutputTexture(outputImage, view);
//     tonemapper.run(cmd, hdrImage);
// usage: facade_frame <blinnphong|ggx> W H camx camy camz <dir> [envW envH cubeRes mips samples irrRes lutRes]
// Reads <dir>/{pos,nrm,alb}.bin (+ orm, emi, lights for ggx, env for IBL) as tightly packed planes,
// writes <dir>/hdr.bin (RGBA16F) and <dir>/ldr.bin (RGBA8).  Used by tests/test_gpu_facade.py, which
// compares the outputs with the oracle.
#include "pbr/AsyncSubmitter.hpp"
#include "pbr/PbrRenderSystem.hpp"
#include "pbr/TonemapperSystem.hpp"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

namespace {
auto readFile(std::string const& path) -> std::vector<char> {
  std::ifstream f(path, std::ios::binary | std::ios::ate);
  if (!f) throw std::runtime_error("cannot open " + path);
  std::vector<char> data(static_cast<std::size_t>(f.tellg()));
  f.seekg(0);
  f.read(data.data(), static_cast<std::streamsize>(data.size()));
  return data;
}
void writeFile(std::string const& path, std::vector<char> const& data) {
  std::ofstream f(path, std::ios::binary);
  f.write(data.data(), static_cast<std::streamsize>(data.size()));
}
// tight host rows <-> pitched device image
void upload(pbr::Image2D const& img, std::vector<char> const& host, void* stream) {
  auto const e = img.getExtent();
  std::size_t const row = std::size_t(e.width) * vk::texelBytes(img.getFormat());
  if (host.size() != row * e.height) throw std::runtime_error("plane size mismatch");
  for (std::uint32_t y = 0; y < e.height; ++y)
    pbr::core::check(pbrcu_memcpy_h2d(static_cast<char*>(img.getImage().data) + y * img.getImage().pitch, host.data() + y * row, row, stream), "upload");
}
auto download(vk::Image img, std::uint32_t w, std::uint32_t h, std::uint32_t texel, void* stream) -> std::vector<char> {
  std::vector<char> host(std::size_t(w) * h * texel);
  for (std::uint32_t y = 0; y < h; ++y)
    pbr::core::check(pbrcu_memcpy_d2h(host.data() + std::size_t(y) * w * texel, static_cast<char*>(img.data) + y * img.pitch, std::size_t(w) * texel, stream), "download");
  pbr::core::check(pbrcu_stream_synchronize(stream), "sync");
  return host;
}
}  // namespace

int main(int argc, char** argv) try {
  if (argc < 8) { std::fprintf(stderr, "usage: %s <blinnphong|ggx> W H cx cy cz dir [envW envH cubeRes mips samples irrRes lutRes]\n", argv[0]); return 2; }
  bool const ggx = std::strcmp(argv[1], "ggx") == 0;
  vk::Extent2D const extent {std::uint32_t(std::atoi(argv[2])), std::uint32_t(std::atoi(argv[3]))};
  glm::vec3 const camPos {float(std::atof(argv[4])), float(std::atof(argv[5])), float(std::atof(argv[6]))};
  std::string const dir = argv[7];

  auto gpu = pbr::core::makeGpuHandle({});
  pbr::MemoryAllocator allocator {gpu};
  pbr::PbrRenderSystem pbrSystem {gpu, {.shadingModel = ggx ? pbr::ShadingModel::MetallicRoughness : pbr::ShadingModel::BlinnPhong}};
  pbr::TonemapperSystem tonemapper {gpu, {}};
  auto gBuffer = pbrSystem.allocateGBuffer(allocator, extent);
  auto hdrImage = tonemapper.allocateHdrImage(allocator, extent);
  pbr::Image2D outputImage {allocator, vk::Format::eR8G8B8A8Unorm, extent};  // stands in for the swapchain image

  pbrcu_stream stream {};
  pbr::core::check(pbrcu_stream_create(&stream), "stream");
  vk::CommandBuffer cmd {stream};

  // the caller-filled G-buffer (the geometry pass's output)
  upload(gBuffer.getPositions(), readFile(dir + "/pos.bin"), stream);
  upload(gBuffer.getNormals(), readFile(dir + "/nrm.bin"), stream);
  upload(gBuffer.getAlbedo(), readFile(dir + "/alb.bin"), stream);

  pbr::Scene scene;
  auto camera = std::make_shared<pbr::CameraUniform>(allocator, pbr::makeCameraData(camPos, {0, 0, 0}, 1.5707963f, float(extent.width) / float(extent.height)));
  scene.addNode(pbr::Node {"camera"}).setCamera(camera);

  if (ggx) {
    upload(gBuffer.getOrm(), readFile(dir + "/orm.bin"), stream);
    upload(gBuffer.getEmissive(), readFile(dir + "/emi.bin"), stream);
    auto const lights = readFile(dir + "/lights.bin");
    for (std::size_t i = 0; i + sizeof(pbr::PointLight) <= lights.size(); i += sizeof(pbr::PointLight)) {
      pbr::PointLight l;
      std::memcpy(&l, lights.data() + i, sizeof(l));
      scene.addLight(l);
    }
    if (argc >= 15) {
      std::uint32_t const envW = std::atoi(argv[8]), envH = std::atoi(argv[9]);
      pbr::EnvironmentCreateInfo info {std::uint32_t(std::atoi(argv[10])), std::uint32_t(std::atoi(argv[11])), std::uint32_t(std::atoi(argv[12])),
                                       std::uint32_t(std::atoi(argv[13])), std::uint32_t(std::atoi(argv[14]))};
      auto const envHost = readFile(dir + "/env.bin");
      auto envDev = allocator.allocateBuffer(envHost.size(), {});
      pbr::core::check(pbrcu_memcpy_h2d(envDev.get(), envHost.data(), envHost.size(), stream), "env upload");
      auto env = std::make_shared<pbr::Environment>(gpu, allocator, info);
      env->precompute(cmd, static_cast<float const*>(envDev.get()), envW, envH);
      pbr::core::check(pbrcu_stream_synchronize(stream), "sync");
      pbrSystem.setEnvironment(env);
    }
  }

  // ---- the reference's per-frame sequence (App.cpp:387-392) ----
  pbrSystem.render(cmd, scene, gBuffer, hdrImage.getImage(), extent);
  hdrImage.updateOutputTexture(outputImage.getImage(), outputImage.getImageView());
  tonemapper.run(cmd, hdrImage);
  pbr::AsyncSubmitter submitter {gpu};
  submitter.submit({.cmdBuffer = cmd});
  (void)submitter.wait();

  writeFile(dir + "/hdr.bin", download(hdrImage.getImage().getImage(), extent.width, extent.height, 8, stream));
  writeFile(dir + "/ldr.bin", download(hdrImage.getOutputImage(), extent.width, extent.height, 4, stream));

  // error convention: a scene without a camera clears the target and draws nothing (PbrRenderSystem.cpp:486-499)
  pbr::Scene empty;
  pbrSystem.render(cmd, empty, gBuffer, hdrImage.getImage(), extent);
  auto const cleared = download(hdrImage.getImage().getImage(), extent.width, extent.height, 8, stream);
  for (char c : cleared)
    if (c != 0) { std::fprintf(stderr, "camera-less render did not clear the target\n"); return 1; }
  pbrcu_stream_destroy(stream);
  std::printf("facade_frame ok\n");
  return 0;
} catch (std::exception const& e) {
  std::fprintf(stderr, "facade_frame: %s\n", e.what());
  return 1;
}
