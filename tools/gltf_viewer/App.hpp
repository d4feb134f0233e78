// App -- the viewer application, headless.  Mirrors src/gltf_viewer/App.cpp: construction (:236-283: gpu handle,
// allocator, render systems, G-buffer, HDR image, asset load through gltf::Loader + TransferStager) and the frame
// loop (:285-318: controller update, camera upload, the first nine mesh nodes spun about +Y by deltaTime, then
// :384-392 render + tonemap).  What needed a window is replaced: the swapchain image is a device RGBA8 image that
// is written to a PNG at the end, the ImGui performance overlay ("Frame time %.3f ms", ui/PerformanceOverlay.cpp:15-19)
// is a JSON list of the same per-frame times, input events come from the command line (Options).
#pragma once
#include "FlyCamera.hpp"
#include "WriteImage.hpp"

#include "pbr/AsyncSubmitter.hpp"
#include "pbr/PbrRenderSystem.hpp"
#include "pbr/TonemapperSystem.hpp"
#include "pbr/TransferStager.hpp"
#include "pbr/gltf/Loader.hpp"

#include <chrono>
#include <cstdio>
#include <filesystem>
#include <optional>
#include <string>
#include <vector>

namespace app {
struct Options {
  std::filesystem::path asset;
  vk::Extent2D extent {1280, 720};
  std::uint32_t frames {60};
  std::optional<double> fixedDeltaTime;  // default: the measured frame duration, like the reference
  HeldKeys input {};
  double lookX {}, lookY {}, scroll {};  // per frame: cursor movement in pixels, scroll offset
  std::optional<glm::vec3> position;
  std::optional<float> yaw, pitch, fov;
  bool spin {true};
  bool ggx {false};  // Tier B: GGX + point lights + IBL instead of the reference's Blinn-Phong lighting pass
  std::string outPng, outPfm, outTimes, dumpDir;
  // CUDA addition (BASELINE configs[4], "256 camera views of the glTF scene"): views > 0 renders cameras viewFirst .. viewFirst + views - 1 of
  // an orbit of viewTotal cameras around the origin through PbrRenderSystem::renderViews instead of running the frame loop
  std::uint32_t views {}, viewFirst {}, viewTotal {};
  float orbitRadius {12.0f}, orbitHeight {-2.0f};
};

class App {
  Options _opt;
  pbr::core::SharedGpuHandle _gpu;
  std::shared_ptr<pbr::MemoryAllocator> _allocator;
  pbr::PbrRenderSystem _pbrSystem;
  pbr::TonemapperSystem _tonemapper;
  pbr::GBuffer _gBuffer;
  pbr::HdrImage _hdrImage;
  pbr::Image2D _outputImage;  // stands in for the swapchain image
  pbrcu_stream _stream {};
  FlyCamera _camera;
  pbr::Scene _scene;
  std::shared_ptr<pbr::Environment> _environment;
  pbr::Allocation _envSource;
  std::vector<double> _frameMs;

  static auto makeStream() -> pbrcu_stream {
    pbrcu_stream s {};
    pbr::core::check(pbrcu_stream_create(&s), "stream");
    return s;
  }
  auto download(vk::Image img, std::uint32_t texel) -> std::vector<std::uint8_t> {
    auto const w = _opt.extent.width, h = _opt.extent.height;
    std::vector<std::uint8_t> host(std::size_t(w) * h * texel);
    for (std::uint32_t y = 0; y < h; ++y)
      pbr::core::check(pbrcu_memcpy_d2h(host.data() + std::size_t(y) * w * texel, static_cast<char*>(img.data) + y * img.pitch, std::size_t(w) * texel, _stream), "download");
    pbr::core::check(pbrcu_stream_synchronize(_stream), "sync");
    return host;
  }
  static auto halfToFloat(std::uint16_t h) -> float {
    std::uint32_t const sign = std::uint32_t(h & 0x8000u) << 16, exp = (h >> 10) & 0x1Fu, man = h & 0x3FFu;
    std::uint32_t bits;
    if (exp == 0) {
      if (man == 0) bits = sign;
      else {
        int e = -1;
        std::uint32_t m = man;
        do { ++e; m <<= 1; } while ((m & 0x400u) == 0);
        bits = sign | std::uint32_t(127 - 15 - e) << 23 | (m & 0x3FFu) << 13;
      }
    } else if (exp == 31) bits = sign | 0x7F800000u | man << 13;
    else bits = sign | (exp + 112) << 23 | man << 13;
    float f;
    std::memcpy(&f, &bits, 4);
    return f;
  }

public:
  explicit App(Options opt)
      : _opt(std::move(opt))
      , _gpu(pbr::core::makeGpuHandle({}))
      , _allocator(std::make_shared<pbr::MemoryAllocator>(_gpu))
      , _pbrSystem(_gpu, {.shadingModel = _opt.ggx ? pbr::ShadingModel::MetallicRoughness : pbr::ShadingModel::BlinnPhong})
      , _tonemapper(_gpu, {})
      , _gBuffer(_pbrSystem.allocateGBuffer(*_allocator, _opt.extent))
      , _hdrImage(_tonemapper.allocateHdrImage(*_allocator, _opt.extent))
      , _outputImage(*_allocator, vk::Format::eR8G8B8A8Unorm, _opt.extent)
      , _stream(makeStream())
      , _camera(makeFlyCamera(_opt.extent.width, _opt.extent.height)) {
    vk::CommandBuffer cmd {_stream};
    {  // App.cpp:262-283: load the asset's first scene through one TransferStager batch
      pbr::TransferStager stager {_gpu, _allocator};
      pbr::gltf::Loader loader;
      auto asset = loader.loadAsset(_opt.asset, {.gpu = _gpu, .allocator = _allocator});
      _scene = asset.loadScene(stager, asset.getDocument().defaultScene.value_or(0));
      stager.submit(cmd);
      stager.wait();
    }
    if (_opt.ggx) {  // Tier B set-up: a procedural sky, its IBL precompute, four point lights around the origin
      pbr::EnvironmentCreateInfo const info {.cubeResolution = 256, .mipLevels = 9, .sampleCount = 256, .irradianceResolution = 32, .lutResolution = 256};
      _environment = std::make_shared<pbr::Environment>(_gpu, *_allocator, info);
      std::uint32_t const ew = 1024, eh = 512;
      _envSource = _allocator->allocateBuffer(std::size_t(ew) * eh * 16, {});
      pbr::core::check(pbrcu_synth_environment(_gpu->getContext(), static_cast<float*>(_envSource.get()), ew, eh, 0x5EED0002ull, _stream), "environment");
      _environment->precompute(cmd, static_cast<float const*>(_envSource.get()), ew, eh);
      _pbrSystem.setEnvironment(_environment);
      for (int i = 0; i < 4; ++i) {
        float const a = 1.5707963f * float(i);
        _scene.addLight({.position = {6.0f * std::cos(a), -4.0f, 6.0f * std::sin(a)}, .range = 0.0f, .color = {1.0f, 0.95f, 0.9f}, .intensity = 80.0f});
      }
    }
    place(_camera, _opt.position.value_or(glm::vec3 {0.0f}), _opt.yaw.value_or(Limits::startYaw), _opt.pitch.value_or(Limits::startPitch),
          _opt.fov.value_or(Limits::startFov));
  }
  App(App const&) = delete;
  auto operator=(App const&) -> App& = delete;
  ~App() noexcept { pbrcu_stream_destroy(_stream); }

  auto run() -> void {
    vk::CommandBuffer cmd {_stream};
    pbr::AsyncSubmitter submitter {_gpu};
    auto lastFrame = std::chrono::high_resolution_clock::now();
    double cursorX = 0.0, cursorY = 0.0;
    for (std::uint32_t frame = 0; frame < _opt.frames; ++frame) {
      auto const thisFrame = std::chrono::high_resolution_clock::now();
      auto const frameDuration = thisFrame - lastFrame;
      lastFrame = thisFrame;
      double const deltaTime = _opt.fixedDeltaTime.value_or(std::chrono::duration_cast<std::chrono::duration<double>>(frameDuration).count());
      if (frame > 0) _frameMs.push_back(std::chrono::duration_cast<std::chrono::duration<double, std::milli>>(frameDuration).count());

      // "pollEvents": the scripted per-frame input
      if (_opt.lookX != 0.0 || _opt.lookY != 0.0) {
        cursorX += _opt.lookX; cursorY += _opt.lookY;
        cursorTo(_camera, _opt.input, cursorX, cursorY);
      }
      if (_opt.scroll != 0.0) scrollBy(_camera, _opt.scroll);

      advance(_camera, _opt.input, deltaTime);
      if (auto const camera = _scene.findCamera(); camera) camera.value()->set(cameraData(_camera));
      if (_opt.spin) {
        int taken = 0;
        for (auto& node : _scene.iterateAllNodes()) {
          if (node.getMesh() == nullptr) continue;
          if (taken++ == 9) break;
          auto transform = node.getTransform();
          transform.rotation = glm::rotate(transform.rotation, static_cast<float>(deltaTime), {0.0f, 1.0f, 0.0f});
          node.setTransform(transform);
        }
      }
      // renderAndPresent / recordCommands (App.cpp:384-392)
      _pbrSystem.render(cmd, _scene, _gBuffer, _hdrImage.getImage(), _hdrImage.getExtent());
      _hdrImage.updateOutputTexture(_outputImage.getImage(), _outputImage.getImageView());
      _tonemapper.run(cmd, _hdrImage);
      submitter.submit({.cmdBuffer = cmd});
      (void)submitter.wait();  // the reference waits on the frame's fence before reusing its resources
    }
    _frameMs.push_back(std::chrono::duration_cast<std::chrono::duration<double, std::milli>>(std::chrono::high_resolution_clock::now() - lastFrame).count());
    writeOutputs();
  }

  [[nodiscard]] auto frameTimes() const noexcept -> std::vector<double> const& { return _frameMs; }

  // --views: one batched geometry pass + one batched lighting pass over a block of orbit cameras; timed after a warm-up call (scratch
  // growth, table uploads), geometry alone and geometry + lighting, host clock around enqueue + stream synchronise
  auto runViews() -> void {
    vk::CommandBuffer cmd {_stream};
    auto const n = _opt.views, total = _opt.viewTotal ? _opt.viewTotal : _opt.views;
    auto const block = _pbrSystem.allocateViewBlock(*_allocator, _opt.extent, n);
    std::vector<pbr::CameraData> cameras(n);
    float const aspect = float(_opt.extent.width) / float(_opt.extent.height);
    for (std::uint32_t v = 0; v < n; ++v) {
      float const a = 6.283185307179586f * float(_opt.viewFirst + v) / float(total);
      cameras[v] = pbr::makeCameraData({_opt.orbitRadius * std::cos(a), _opt.orbitHeight, _opt.orbitRadius * std::sin(a)}, {0.0f, 0.0f, 0.0f},
                                       _opt.fov.value_or(Limits::startFov), aspect);
    }
    auto const timed = [&](bool lighting) {
      auto const t0 = std::chrono::high_resolution_clock::now();
      _pbrSystem.renderViews(cmd, _scene, cameras, block, lighting);
      pbr::core::check(pbrcu_stream_synchronize(_stream), "sync");
      return std::chrono::duration_cast<std::chrono::duration<double, std::milli>>(std::chrono::high_resolution_clock::now() - t0).count();
    };
    (void)timed(true);
    double msGeometry = 1e300, msBoth = 1e300;
    for (std::uint32_t rep = 0; rep < (_opt.frames < 1 ? 1u : _opt.frames); ++rep) {
      msGeometry = std::min(msGeometry, timed(false));
      msBoth = std::min(msBoth, timed(true));
    }
    std::uint64_t triangles = 0, draws = 0;
    for (auto const& node : _scene.iterateAllNodes())
      if (node.getMesh())
        for (auto const& primitive : node.getMesh()->getPrimitives()) { triangles += primitive.indexCount / 3; ++draws; }
    if (!_opt.outTimes.empty()) {
      std::ofstream f(_opt.outTimes);
      f << "{\"views\": " << n << ", \"view_first\": " << _opt.viewFirst << ", \"view_total\": " << total << ", \"width\": " << _opt.extent.width
        << ", \"height\": " << _opt.extent.height << ", \"triangles_per_view\": " << triangles << ", \"draws\": " << draws << ", \"repeats\": "
        << (_opt.frames < 1 ? 1u : _opt.frames) << ", \"ms_geometry\": " << msGeometry << ", \"ms_geometry_plus_shading\": " << msBoth << "}\n";
    }
    if (!_opt.dumpDir.empty()) {  // for the parity tests: first, middle and last view -- planes, frame, and the matrices they were drawn with
      auto const w = _opt.extent.width, h = _opt.extent.height;
      for (std::uint32_t const v : {0u, n / 2, n - 1}) {
        auto const g = block.raw(v);
        auto const dump = [&](char const* name, void const* src, std::size_t texel) {
          std::vector<std::uint8_t> host(std::size_t(w) * h * texel);
          pbr::core::check(pbrcu_memcpy_d2h(host.data(), src, host.size(), _stream), "download");
          pbr::core::check(pbrcu_stream_synchronize(_stream), "sync");
          std::ofstream(_opt.dumpDir + "/view" + std::to_string(v) + "_" + name, std::ios::binary).write(reinterpret_cast<char const*>(host.data()), std::streamsize(host.size()));
        };
        dump("pos.bin", g.positions, 8);
        dump("nrm.bin", g.normals, 8);
        dump("alb.bin", g.albedo, 8);
        dump("depth.bin", g.depth, 4);
        dump("hdr.bin", block.getFrame(v).data, 8);
        auto const m = matricesOf(cameras[v]);
        std::ofstream(_opt.dumpDir + "/view" + std::to_string(v) + "_matrices.bin", std::ios::binary).write(reinterpret_cast<char const*>(m.data()), std::streamsize(m.size() * 4));
      }
    }
  }

private:
  auto writeOutputs() -> void {
    auto const w = _opt.extent.width, h = _opt.extent.height;
    if (!_opt.outPng.empty()) writePng(_opt.outPng, download(_hdrImage.getOutputImage(), 4).data(), w, h);
    if (!_opt.outPfm.empty()) {
      auto const hdr = download(_hdrImage.getImage().getImage(), 8);
      std::vector<float> rgb(std::size_t(w) * h * 3);
      for (std::size_t i = 0; i < std::size_t(w) * h; ++i)
        for (int c = 0; c < 3; ++c) {
          std::uint16_t v;
          std::memcpy(&v, hdr.data() + i * 8 + std::size_t(c) * 2, 2);
          rgb[i * 3 + std::size_t(c)] = halfToFloat(v);
        }
      writePfm(_opt.outPfm, rgb.data(), w, h);
    }
    if (!_opt.outTimes.empty()) {
      std::ofstream f(_opt.outTimes);
      double sum = 0.0;
      for (auto const t : _frameMs) sum += t;
      f << "{\"frames\": " << _frameMs.size() << ", \"width\": " << w << ", \"height\": " << h << ", \"mean_frame_time_ms\": " << (sum / double(_frameMs.size()))
        << ", \"frame_time_ms\": [";
      for (std::size_t i = 0; i < _frameMs.size(); ++i) f << (i ? ", " : "") << _frameMs[i];
      f << "]}\n";
    }
    if (!_opt.dumpDir.empty()) {  // for the parity tests: the last frame's G-buffer and the matrices it was drawn with
      auto const dump = [&](char const* name, vk::Image img, std::uint32_t texel) {
        auto const bytes = download(img, texel);
        std::ofstream(_opt.dumpDir + "/" + name, std::ios::binary).write(reinterpret_cast<char const*>(bytes.data()), std::streamsize(bytes.size()));
      };
      dump("pos.bin", _gBuffer.getPositions().getImage(), 8);
      dump("nrm.bin", _gBuffer.getNormals().getImage(), 8);
      dump("alb.bin", _gBuffer.getAlbedo().getImage(), 8);
      dump("depth.bin", _gBuffer.getDepth().getImage(), 4);
      dump("hdr.bin", _hdrImage.getImage().getImage(), 8);
      dump("ldr.bin", _hdrImage.getOutputImage(), 4);
      auto const m = matricesOf(_scene.findCamera().value()->get());
      std::ofstream(_opt.dumpDir + "/matrices.bin", std::ios::binary).write(reinterpret_cast<char const*>(m.data()), std::streamsize(m.size() * 4));
    }
  }
  // view, proj, eye (+ pad), then model (16) and normal model (12) of every draw in render() order
  auto matricesOf(pbr::CameraData const& cam) -> std::vector<float> {
    std::vector<float> m;
    m.insert(m.end(), reinterpret_cast<float const*>(&cam.view), reinterpret_cast<float const*>(&cam.view) + 16);
    m.insert(m.end(), reinterpret_cast<float const*>(&cam.proj), reinterpret_cast<float const*>(&cam.proj) + 16);
    m.insert(m.end(), {cam.position.x, cam.position.y, cam.position.z, 0.0f});
    for (auto const& node : _scene.iterateAllNodes()) {
      if (!node.getMesh()) continue;
      auto const tr = node.getTransform();
      auto const pc = pbr::makeModelPushConstant(tr.position, tr.rotation, tr.scale);
      for (std::size_t p = 0; p < node.getMesh()->getPrimitives().size(); ++p) {
        m.insert(m.end(), reinterpret_cast<float const*>(&pc.model), reinterpret_cast<float const*>(&pc.model) + 16);
        m.insert(m.end(), reinterpret_cast<float const*>(&pc.normalModel), reinterpret_cast<float const*>(&pc.normalModel) + 12);
      }
    }
    return m;
  }
};
}  // namespace app
