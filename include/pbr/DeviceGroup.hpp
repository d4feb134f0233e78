// pbr/DeviceGroup.hpp -- the GPUs of one box driven from one host thread (CUDA addition; no reference counterpart: the
// reference owns one device with one queue, src/pbr_engine/core/pbr/core/GpuHandle.cpp:77-81, and its only client is one
// frame loop, src/gltf_viewer/App.cpp:444-483).  A thin RAII face of pbrcu_mg_* (csrc/mg.cu): every member has its own
// pbrcu_ctx and stream, peer access is enabled between all pairs, the calls ENQUEUE on the members' streams and end with an
// event barrier, so work enqueued afterwards on any member's stream sees the whole frame.  Errors become the exception the
// rest of the facade throws (std::runtime_error).
//
//     pbr::DeviceGroup group({0, 1, 2, 3});
//     auto bands = group.rowBands(height, previousFrameTileRowCosts);            // cost-balanced, tile-aligned
//     group.shadeFrame(gbuffers, cameraPosition, scene.getLights(), tiles, ibls, bands, frames, pitch, pbr::Gather::P2P);
//     group.synchronize();                                                        // every member's frame holds the whole image
#pragma once
#include "pbr/Scene.hpp"
#include "pbr/core/GpuHandle.hpp"

#include <array>
#include <cstdint>
#include <span>
#include <vector>

namespace pbr {
enum class Gather : int {
  None = PBRCU_MG_GATHER_NONE,  // member i's frame holds its own row band
  P2P = PBRCU_MG_GATHER_P2P,    // fused into the shading kernel's stores: every texel also goes to the other members' frames (NVLink)
  Nccl = PBRCU_MG_GATHER_NCCL   // one grouped ncclBroadcast per band after the shading
};

struct RowBand { std::uint32_t y0, y1; };  // rows [y0, y1) of the frame

class DeviceGroup {
  pbrcu_mg* _mg {};

public:
  explicit DeviceGroup(std::span<int const> devices) {
    core::check(pbrcu_mg_create(devices.data(), static_cast<int>(devices.size()), &_mg), "pbr::DeviceGroup");
  }
  DeviceGroup(std::initializer_list<int> devices) : DeviceGroup(std::span<int const>(devices.begin(), devices.size())) {}
  DeviceGroup(DeviceGroup const&) = delete;
  auto operator=(DeviceGroup const&) -> DeviceGroup& = delete;
  ~DeviceGroup() { pbrcu_mg_destroy(_mg); }

  [[nodiscard]] auto size() const noexcept -> std::size_t { return static_cast<std::size_t>(pbrcu_mg_size(_mg)); }
  [[nodiscard]] auto getContext(std::size_t member) const noexcept -> pbrcu_ctx* { return pbrcu_mg_ctx(_mg, static_cast<int>(member)); }
  [[nodiscard]] auto getStream(std::size_t member) const noexcept -> pbrcu_stream { return pbrcu_mg_stream(_mg, static_cast<int>(member)); }
  [[nodiscard]] auto getDeviceOrdinal(std::size_t member) const noexcept -> int { return pbrcu_mg_device(_mg, static_cast<int>(member)); }
  // makes the member's device current: allocate its buffers (pbrcu_malloc, pbrcu_gbuffer_alloc ...) after this
  auto makeCurrent(std::size_t member) const -> void { core::check(pbrcu_mg_set_device(_mg, static_cast<int>(member)), "pbr::DeviceGroup::makeCurrent"); }

  // Row bands on light-tile boundaries; tileRowCosts (one per 16-row tile row, e.g. the previous frame's sum over the row's tiles of
  // 9 + lights in the tile) balances them by cost, an empty span cuts equal numbers of tile rows.
  [[nodiscard]] auto rowBands(std::uint32_t height, std::span<float const> tileRowCosts = {}) const -> std::vector<RowBand> {
    if (!tileRowCosts.empty() && tileRowCosts.size() != (height + PBRCU_TILE - 1) / PBRCU_TILE)
      throw std::runtime_error("pbr::DeviceGroup::rowBands: one cost per tile row expected");
    std::vector<std::uint32_t> raw(2 * size());
    core::check(pbrcu_mg_row_bands(height, static_cast<int>(size()), tileRowCosts.empty() ? nullptr : tileRowCosts.data(), raw.data()), "pbr::DeviceGroup::rowBands");
    std::vector<RowBand> bands(size());
    for (std::size_t i = 0; i < size(); ++i) bands[i] = {raw[2 * i], raw[2 * i + 1]};
    return bands;
  }

  // One frame over the group.  gBuffers / tiles / ibls / frames: one entry per member, in that member's memory; tiles or ibls may be
  // empty (no clustering / no IBL).  lights: the scene's list (host memory).
  auto shadeFrame(std::span<pbrcu_gbuffer const> gBuffers, glm::vec3 cameraPosition, std::span<PointLight const> lights,
                  std::span<pbrcu_light_tiles const> tiles, std::span<pbrcu_ibl const> ibls, std::span<RowBand const> bands,
                  std::span<void* const> frames, std::size_t pitch, Gather gather) -> void {
    if (gBuffers.size() != size() || bands.size() != size() || frames.size() != size() || (!tiles.empty() && tiles.size() != size()) ||
        (!ibls.empty() && ibls.size() != size()))
      throw std::runtime_error("pbr::DeviceGroup::shadeFrame: one entry per member expected");
    static_assert(sizeof(RowBand) == 2 * sizeof(std::uint32_t));
    float const cam[3] = {cameraPosition.x, cameraPosition.y, cameraPosition.z};
    core::check(pbrcu_mg_shade_frame(_mg, gBuffers.data(), cam, reinterpret_cast<pbrcu_light const*>(lights.data()), static_cast<std::uint32_t>(lights.size()),
                                     tiles.empty() ? nullptr : tiles.data(), ibls.empty() ? nullptr : ibls.data(), &bands.data()->y0, frames.data(), pitch,
                                     static_cast<int>(gather)),
                "pbr::DeviceGroup::shadeFrame");
  }

  // The same frame display-ready: RGBA8 frames, the ACES + gamma tonemap fused into the shading kernel's stores (half the gather's bytes).
  auto shadeFrameLdr(std::span<pbrcu_gbuffer const> gBuffers, glm::vec3 cameraPosition, std::span<PointLight const> lights,
                     std::span<pbrcu_light_tiles const> tiles, std::span<pbrcu_ibl const> ibls, std::span<RowBand const> bands,
                     std::span<void* const> ldrFrames, std::size_t pitch, Gather gather) -> void {
    if (gBuffers.size() != size() || bands.size() != size() || ldrFrames.size() != size() || (!tiles.empty() && tiles.size() != size()) ||
        (!ibls.empty() && ibls.size() != size()))
      throw std::runtime_error("pbr::DeviceGroup::shadeFrameLdr: one entry per member expected");
    float const cam[3] = {cameraPosition.x, cameraPosition.y, cameraPosition.z};
    core::check(pbrcu_mg_shade_frame_ldr(_mg, gBuffers.data(), cam, reinterpret_cast<pbrcu_light const*>(lights.data()), static_cast<std::uint32_t>(lights.size()),
                                         tiles.empty() ? nullptr : tiles.data(), ibls.empty() ? nullptr : ibls.data(), &bands.data()->y0, ldrFrames.data(), pitch,
                                         static_cast<int>(gather)),
                "pbr::DeviceGroup::shadeFrameLdr");
  }

  // Batched views: member i shades viewCounts[i] views laid out viewStride8 / viewStride4 bytes apart from firstViews[i]'s planes.
  auto shadeViews(std::span<pbrcu_gbuffer const> firstViews, std::size_t viewStride8, std::size_t viewStride4, std::span<std::uint32_t const> viewCounts,
                  std::span<float const* const> cameraPositions, std::span<PointLight const> lights, std::span<pbrcu_ibl const> ibls,
                  std::span<void* const> outputs, std::size_t outPitch, std::size_t outViewStride) -> void {
    if (firstViews.size() != size() || viewCounts.size() != size() || cameraPositions.size() != size() || outputs.size() != size() ||
        (!ibls.empty() && ibls.size() != size()))
      throw std::runtime_error("pbr::DeviceGroup::shadeViews: one entry per member expected");
    core::check(pbrcu_mg_shade_views(_mg, firstViews.data(), viewStride8, viewStride4, viewCounts.data(), cameraPositions.data(),
                                     reinterpret_cast<pbrcu_light const*>(lights.data()), static_cast<std::uint32_t>(lights.size()),
                                     ibls.empty() ? nullptr : ibls.data(), outputs.data(), outPitch, outViewStride),
                "pbr::DeviceGroup::shadeViews");
  }

  // The geometry pass of view blocks: member i rasterises viewCounts[i] cameras over its own copy of the scene (draws[i]) into the G-buffers
  // that start at firstViews[i]; follow with shadeViews on the same blocks.
  auto geometryViews(std::span<pbrcu_gbuffer const> firstViews, std::size_t viewStride8, std::size_t viewStride4, std::span<std::uint32_t const> viewCounts,
                     std::span<float const* const> viewMatrices, std::span<float const* const> projMatrices, std::span<pbrcu_draw const* const> draws,
                     std::span<std::uint32_t const> drawCounts) -> void {
    if (firstViews.size() != size() || viewCounts.size() != size() || viewMatrices.size() != size() || projMatrices.size() != size() || draws.size() != size() ||
        drawCounts.size() != size())
      throw std::runtime_error("pbr::DeviceGroup::geometryViews: one entry per member expected");
    core::check(pbrcu_mg_geometry_views(_mg, firstViews.data(), viewStride8, viewStride4, viewCounts.data(), viewMatrices.data(), projMatrices.data(), draws.data(),
                                        drawCounts.data()),
                "pbr::DeviceGroup::geometryViews");
  }

  // GGX prefilter sharded over the members; gather = true leaves the whole chain in every member's destination.
  auto prefilter(std::span<pbrcu_ibl_source const* const> sources, std::uint32_t mipLevels, std::uint32_t sampleCount, std::span<void* const> chains,
                 bool gather) -> void {
    if (sources.size() != size() || chains.size() != size()) throw std::runtime_error("pbr::DeviceGroup::prefilter: one entry per member expected");
    core::check(pbrcu_mg_prefilter(_mg, sources.data(), mipLevels, sampleCount, chains.data(), gather ? 1 : 0), "pbr::DeviceGroup::prefilter");
  }

  auto barrier() -> void { core::check(pbrcu_mg_barrier(_mg), "pbr::DeviceGroup::barrier"); }
  auto synchronize() -> void { core::check(pbrcu_mg_synchronize(_mg), "pbr::DeviceGroup::synchronize"); }
};
}  // namespace pbr
