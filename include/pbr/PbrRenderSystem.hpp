// pbr/PbrRenderSystem.hpp -- the shading render system, same public surface as the reference's
// (src/pbr_engine/engine/pbr/PbrRenderSystem.hpp:13-55):
//     PbrRenderSystem(core::SharedGpuHandle gpu, PbrRenderSystemCreateInfo info);
//     auto allocateGBuffer(IAllocator&, vk::Extent2D) -> GBuffer;
//     auto render(vk::CommandBuffer, Scene const&, GBuffer const&, Image2D const& renderTarget,
//                 vk::Extent2D renderExtent) -> void;
// render() in the reference records a rasterised geometry pass that fills the G-buffer and a
// full-screen lighting pass that shades it (PbrRenderSystem.cpp:357-365).  Here the lighting pass is a
// CUDA kernel enqueued on the command buffer's stream:
//   ShadingModel::BlinnPhong  (default) = assets/shaders/pbr/lighting.glsl, bit-for-bit the reference
//   ShadingModel::MetallicRoughness     = GGX + point lights + IBL (Tier B, no reference counterpart)
// The geometry pass (K10, a CUDA visibility-buffer rasteriser behind pbrcu_geometry_pass) runs when the scene
// holds mesh nodes; a scene without meshes shades whatever the caller put into the G-buffer (the deferred hot
// path on its own).
// CUDA addition: renderViews() renders a block of camera views of one scene (BASELINE configs[4]) with ONE batched geometry pass
// and ONE batched lighting pass (pbrcu_geometry_pass_batched + pbrcu_shade_ggx_ibl_batched / pbrcu_lighting_blinnphong_batched).
#pragma once
#include "pbr/Environment.hpp"
#include "pbr/GBuffer.hpp"
#include "pbr/ModelPushConstant.hpp"
#include "pbr/Scene.hpp"
#include "pbr/ViewBlock.hpp"

#include <cstring>
#include <memory>
#include <span>
#include <stdexcept>
#include <string>
#include <vector>

namespace pbr {
enum class ShadingModel { BlinnPhong, MetallicRoughness };

struct PbrRenderSystemCreateInfo {
  vk::PipelineShaderStageCreateInfo geometryVertexShader {};
  vk::PipelineShaderStageCreateInfo geometryFragmentShader {};
  vk::PipelineShaderStageCreateInfo lightingVertexShader {};
  vk::PipelineShaderStageCreateInfo lightingFragmentShader {};
  ShadingModel shadingModel {ShadingModel::BlinnPhong};  // CUDA addition
};

class PbrRenderSystem {
public:
  static constexpr auto LIGHTING_PASS_OUTPUT_FORMAT = vk::Format::eR16G16B16A16Sfloat;

private:
  core::SharedGpuHandle _gpu;
  PbrRenderSystemCreateInfo _info;
  std::shared_ptr<Environment const> _environment;
  Allocation _lightsDevice;  // grown on demand (clustered shading only: short light lists travel in the kernel's parameters)
  std::size_t _lightsCapacity {};
  pbrcu_light_tiles _tiles {};  // per-tile light lists of the clustered pass, re-allocated when the extent or the light count grows
  std::uint32_t _tilesWidth {}, _tilesHeight {};

public:
  // more lights than this are culled per 16x16 tile first (K3c, pbrcu_cluster_lights) instead of being looped by every pixel
  static constexpr std::size_t CLUSTERED_LIGHTS_THRESHOLD = 8;

  PbrRenderSystem(core::SharedGpuHandle gpu, PbrRenderSystemCreateInfo info) : _gpu(std::move(gpu)), _info(info) {}
  PbrRenderSystem(PbrRenderSystem const&) = delete;
  auto operator=(PbrRenderSystem const&) -> PbrRenderSystem& = delete;
  ~PbrRenderSystem() { (void)pbrcu_light_tiles_free(_gpu ? _gpu->getContext() : nullptr, &_tiles); }

  [[nodiscard]] auto allocateGBuffer(IAllocator& allocator, vk::Extent2D extent) -> GBuffer {
    return GBuffer {allocator, extent, _info.shadingModel == ShadingModel::MetallicRoughness};
  }

  auto setEnvironment(std::shared_ptr<Environment const> env) -> void { _environment = std::move(env); }

  auto render(vk::CommandBuffer cmdBuffer, Scene const& scene, GBuffer const& gBuffer, Image2D const& renderTarget,
              vk::Extent2D renderExtent) -> void {
    recordGeometryPass(cmdBuffer, scene, gBuffer);
    recordLightingPass(cmdBuffer, scene, gBuffer, renderTarget, vk::Rect2D {.offset = {}, .extent = renderExtent});
  }

  [[nodiscard]] auto allocateViewBlock(IAllocator& allocator, vk::Extent2D extent, std::uint32_t views) -> ViewBlock {
    return ViewBlock {allocator, extent, views, _info.shadingModel == ShadingModel::MetallicRoughness};
  }

  // cameras.size() views of `scene` into the first cameras.size() views of `block`: what cameras.size() calls of render() with the
  // scene's camera set to cameras[v] would leave in G-buffer and render target v, from one batched geometry pass (draw tables and
  // camera-independent vertex work once) and one batched lighting pass.  The scene's own camera is not consulted.  MetallicRoughness
  // shading loops the scene's lights per pixel (no clustering: the tile lists depend on the view).  lightingPass = false stops after the
  // geometry pass (the G-buffers only).
  auto renderViews(vk::CommandBuffer cmdBuffer, Scene const& scene, std::span<CameraData const> cameras, ViewBlock const& block, bool lightingPass = true)
      -> void {
    if (cameras.empty()) return;
    if (cameras.size() > block.getViewCount()) throw std::runtime_error("pbr::PbrRenderSystem::renderViews: more cameras than the block has views");
    if (block.isExtended() != (_info.shadingModel == ShadingModel::MetallicRoughness))
      throw std::runtime_error("pbr::PbrRenderSystem::renderViews: the block was allocated for the other shading model");
    auto* ctx = _gpu->getContext();
    auto* stream = cmdBuffer.stream();
    auto const n = static_cast<std::uint32_t>(cameras.size());
    std::vector<float> views(16 * cameras.size()), projs(16 * cameras.size()), positions(3 * cameras.size());
    for (std::size_t v = 0; v < cameras.size(); ++v) {
      std::memcpy(views.data() + 16 * v, &cameras[v].view, 64);
      std::memcpy(projs.data() + 16 * v, &cameras[v].proj, 64);
      positions[3 * v] = cameras[v].position.x; positions[3 * v + 1] = cameras[v].position.y; positions[3 * v + 2] = cameras[v].position.z;
    }
    bool anyMesh = false;
    auto const draws = collectDraws(scene, anyMesh);
    auto const g0 = block.raw();
    if (anyMesh)
      core::check(pbrcu_geometry_pass_batched(ctx, &g0, block.viewStride8(), block.viewStride4(), n, views.data(), projs.data(), draws.data(),
                                              static_cast<std::uint32_t>(draws.size()), stream),
                  "PbrRenderSystem: batched geometry pass");
    if (!lightingPass) return;
    if (_info.shadingModel == ShadingModel::BlinnPhong) {
      core::check(pbrcu_lighting_blinnphong_batched(ctx, &g0, block.viewStride8(), block.viewStride4(), n, positions.data(), block.getFrames(), block.pitch8(),
                                                    block.frameStride(), stream),
                  "PbrRenderSystem: batched lighting pass");
      return;
    }
    auto const lights = scene.getLights();
    if (lights.size() > 128) throw std::runtime_error("pbr::PbrRenderSystem::renderViews: at most 128 lights (unclustered loop)");
    pbrcu_ibl ibl {};
    if (_environment) ibl = _environment->raw();
    core::check(pbrcu_shade_ggx_ibl_batched(ctx, &g0, block.viewStride8(), block.viewStride4(), n, positions.data(), reinterpret_cast<pbrcu_light const*>(lights.data()),
                                            static_cast<std::uint32_t>(lights.size()), _environment ? &ibl : nullptr, block.getFrames(), block.pitch8(),
                                            block.frameStride(), stream),
                "PbrRenderSystem: batched GGX shading pass");
  }

private:
  // one pbrcu_draw per primitive of every mesh node in iterateAllNodes() order, model / normal matrices from the node's own
  // transform (PbrRenderSystem.cpp:432-462, makeModelPushConstant :446-449)
  [[nodiscard]] static auto collectDraws(Scene const& scene, bool& anyMesh) -> std::vector<pbrcu_draw> {
    std::vector<pbrcu_draw> draws;
    anyMesh = false;
    for (auto const& node : scene.iterateAllNodes()) {
      auto const& mesh = node.getMesh();
      if (!mesh) continue;
      anyMesh = true;
      auto const transform = node.getTransform();
      auto const pushConst = makeModelPushConstant(transform.position, transform.rotation, transform.scale);
      for (auto const& primitive : mesh->getPrimitives()) {
        if (!primitive.material) throw std::runtime_error("pbr::PbrRenderSystem::render: primitive without a material");
        pbrcu_draw d {};
        d.vertices = mesh->getVertexBuffer().get();
        d.indices = static_cast<std::uint16_t const*>(mesh->getIndexBuffer().get());
        d.first_index = primitive.firstIndex;
        d.index_count = primitive.indexCount;
        d.vertex_offset = static_cast<std::int32_t>(primitive.firstVertex);
        d.vertex_count = primitive.vertexCount;
        static_assert(sizeof(pushConst.model) == sizeof(d.model) && sizeof(pushConst.normalModel) == sizeof(d.normal_model));
        std::memcpy(d.model, &pushConst.model, sizeof(d.model));
        std::memcpy(d.normal_model, &pushConst.normalModel, sizeof(d.normal_model));
        auto const color = primitive.material->getMaterialData().color;
        d.material.color[0] = color.x; d.material.color[1] = color.y; d.material.color[2] = color.z; d.material.color[3] = color.w;
        auto const texture = [](std::shared_ptr<Image2D> const& img, char const* what) {
          if (!img || img->getFormat() != vk::Format::eR8G8B8A8Unorm)
            throw std::runtime_error(std::string("pbr::PbrRenderSystem::render: material needs an R8G8B8A8_UNORM ") + what + " texture");
          return pbrcu_texture {img->getImage().data, img->getExtent().width, img->getExtent().height, img->getImage().pitch};
        };
        d.material.color_tex = texture(primitive.material->getColorTexture(), "colour");
        d.material.normal_tex = texture(primitive.material->getNormalTexture(), "normal");
        auto const mr = primitive.material->getMetallicRoughness();
        d.material.occlusion = mr.occlusion; d.material.roughness = mr.roughness; d.material.metallic = mr.metallic;
        d.material.emissive[0] = mr.emissive.x; d.material.emissive[1] = mr.emissive.y; d.material.emissive[2] = mr.emissive.z;
        draws.push_back(d);
      }
    }
    return draws;
  }

  // PbrRenderSystem.cpp:367-465: clear, then one drawIndexed per primitive.
  // Runs when the scene holds mesh nodes; a scene without meshes keeps the caller-filled G-buffer.
  auto recordGeometryPass(vk::CommandBuffer cmdBuffer, Scene const& scene, GBuffer const& gBuffer) -> void {
    bool anyMesh = false;
    auto draws = collectDraws(scene, anyMesh);
    if (!anyMesh) return;
    auto const camera = scene.findCamera();
    if (!camera) draws.clear();  // no camera bound: the pass still clears the attachments
    CameraData const cam = camera ? camera.value()->get() : CameraData {};
    auto const g = gBuffer.raw();
    core::check(pbrcu_geometry_pass(_gpu->getContext(), &g, reinterpret_cast<float const*>(&cam.view), reinterpret_cast<float const*>(&cam.proj), draws.data(),
                                    static_cast<std::uint32_t>(draws.size()), nullptr, cmdBuffer.stream()),
                "PbrRenderSystem: geometry pass");
  }

  auto recordLightingPass(vk::CommandBuffer cmdBuffer, Scene const& scene, GBuffer const& gBuffer,
                          Image2D const& renderTo, vk::Rect2D renderArea) -> void {
    if (renderTo.getFormat() != LIGHTING_PASS_OUTPUT_FORMAT)
      throw std::runtime_error("pbr::PbrRenderSystem::render: render target must be R16G16B16A16_SFLOAT");
    auto* ctx = _gpu->getContext();
    auto* stream = cmdBuffer.stream();
    auto const target = renderTo.getImage();
    auto const ext = gBuffer.getExtent();
    pbrcu_rect rect {0, 0, renderArea.extent.width < ext.width ? renderArea.extent.width : ext.width,
                     renderArea.extent.height < ext.height ? renderArea.extent.height : ext.height};
    // loadOp = clear to 0 over the render area (PbrRenderSystem.cpp:470-480) -- and nothing else is
    // drawn when the scene has no camera (:486-499)
    auto const camera = scene.findCamera();
    if (!camera) {
      core::check(pbrcu_memset_2d(target.data, target.pitch, 0, std::size_t(rect.w) * 8, rect.h, stream), "PbrRenderSystem: clear");
      return;
    }
    if (rect.w == 0 || rect.h == 0) return;
    auto const g = gBuffer.raw();
    auto const pos = camera.value()->get().position;
    float const cam[3] = {pos.x, pos.y, pos.z};
    if (_info.shadingModel == ShadingModel::BlinnPhong) {
      core::check(pbrcu_lighting_blinnphong(ctx, &g, cam, rect, target.data, target.pitch, stream), "PbrRenderSystem: lighting pass");
      return;
    }
    if (!gBuffer.isExtended())
      throw std::runtime_error("pbr::PbrRenderSystem::render: MetallicRoughness shading needs an extended G-buffer");
    auto const lights = scene.getLights();
    auto const nLights = static_cast<std::uint32_t>(lights.size());
    pbrcu_ibl ibl {};
    if (_environment) ibl = _environment->raw();
    if (lights.size() <= CLUSTERED_LIGHTS_THRESHOLD) {
      // a short list goes to the kernel as HOST data (it travels in the kernel's parameter block: no upload, no staging)
      core::check(pbrcu_shade_ggx_ibl(ctx, &g, cam, reinterpret_cast<pbrcu_light const*>(lights.data()), nLights, nullptr, _environment ? &ibl : nullptr, rect, target.data, target.pitch, stream),
                  "PbrRenderSystem: GGX shading pass");
      return;
    }
    // many lights: per-tile light lists first (K3c), then every pixel loops its tile's list only
    if (lights.size() > _lightsCapacity) {
      _lightsDevice = Allocation {};
      void* p = nullptr;
      core::check(pbrcu_malloc(ctx, lights.size() * sizeof(pbrcu_light), &p), "PbrRenderSystem: light buffer");
      _lightsDevice = Allocation {_gpu, p, lights.size() * sizeof(pbrcu_light), false};
      _lightsCapacity = lights.size();
    }
    core::check(pbrcu_memcpy_h2d(_lightsDevice.get(), lights.data(), lights.size() * sizeof(pbrcu_light), stream), "PbrRenderSystem: light upload");
    if (_tilesWidth != ext.width || _tilesHeight != ext.height || _tiles.max_per_tile < nLights) {
      core::check(pbrcu_light_tiles_free(ctx, &_tiles), "PbrRenderSystem: light tiles");
      core::check(pbrcu_light_tiles_alloc(ctx, ext.width, ext.height, nLights < 1024 ? nLights : 1024, &_tiles), "PbrRenderSystem: light tiles");
      _tilesWidth = ext.width; _tilesHeight = ext.height;
    }
    auto const* lightsDev = static_cast<pbrcu_light const*>(_lightsDevice.get());
    core::check(pbrcu_cluster_lights(ctx, &g, lightsDev, nLights, &_tiles, stream), "PbrRenderSystem: light clustering");
    core::check(pbrcu_shade_ggx_ibl(ctx, &g, cam, lightsDev, nLights, &_tiles, _environment ? &ibl : nullptr, rect, target.data, target.pitch, stream),
                "PbrRenderSystem: GGX shading pass");
  }
};
}  // namespace pbr
