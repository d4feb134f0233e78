// pbr/Environment.hpp -- image-based-lighting set (Tier B; NO reference counterpart: the reference has
// no cube-map type and no IBL shaders, cmake/CompileShaders.cmake:17-33 is the full shader list).
// Owns the device textures the GGX shading pass samples and runs the precompute K4..K9 through the
// C ABI: equirect -> cube, mip chain, GGX-prefiltered specular chain, cosine irradiance, split-sum LUT.
#pragma once
#include "pbr/memory/IAllocator.hpp"

#include <cstdint>

namespace pbr {
struct EnvironmentCreateInfo {
  std::uint32_t cubeResolution {2048};
  std::uint32_t mipLevels {12};
  std::uint32_t sampleCount {1024};
  std::uint32_t irradianceResolution {64};
  std::uint32_t lutResolution {512};
};

class Environment {
  core::SharedGpuHandle _gpu;
  EnvironmentCreateInfo _info;
  Allocation _source, _prefiltered, _irradiance, _lut;

public:
  Environment(core::SharedGpuHandle gpu, IAllocator& allocator, EnvironmentCreateInfo info)
      : _gpu(std::move(gpu)), _info(info) {
    auto const chainBytes = pbrcu_cube_chain_texels(info.cubeResolution, info.mipLevels) * 8;
    _source = allocator.allocateBuffer(chainBytes, {});
    _prefiltered = allocator.allocateBuffer(chainBytes, {});
    _irradiance = allocator.allocateBuffer(std::size_t(6) * info.irradianceResolution * info.irradianceResolution * 8, {});
    _lut = allocator.allocateBuffer(std::size_t(info.lutResolution) * info.lutResolution * 8, {});
  }

  // Records the whole precompute from a lat-long RGBA32F panorama already in device memory.
  auto precompute(vk::CommandBuffer cmd, float const* equirectRgba32fDevice, std::uint32_t width, std::uint32_t height)
      -> void {
    auto* ctx = _gpu->getContext();
    auto* s = cmd.stream();
    core::check(pbrcu_ibl_equirect_to_cube(ctx, equirectRgba32fDevice, width, height, _info.cubeResolution, _source.get(), s), "Environment: equirect_to_cube");
    core::check(pbrcu_ibl_cube_mipgen(ctx, _source.get(), _info.cubeResolution, _info.mipLevels, s), "Environment: cube_mipgen");
    core::check(pbrcu_ibl_prefilter_ggx(ctx, _source.get(), _info.cubeResolution, _info.mipLevels, _info.sampleCount, _prefiltered.get(), 0, _info.mipLevels - 1, 0, 1, s), "Environment: prefilter_ggx");
    core::check(pbrcu_ibl_irradiance(ctx, _source.get(), _info.cubeResolution, _info.mipLevels, _info.irradianceResolution, _info.sampleCount, _irradiance.get(), s), "Environment: irradiance");
    core::check(pbrcu_ibl_brdf_lut(ctx, _info.lutResolution, _info.sampleCount, static_cast<float*>(_lut.get()), s), "Environment: brdf_lut");
  }

  [[nodiscard]] auto getInfo() const noexcept -> EnvironmentCreateInfo const& { return _info; }
  [[nodiscard]] auto getSourceChain() const noexcept -> void* { return _source.get(); }
  [[nodiscard]] auto getPrefiltered() const noexcept -> void* { return _prefiltered.get(); }
  [[nodiscard]] auto getIrradiance() const noexcept -> void* { return _irradiance.get(); }
  [[nodiscard]] auto getLut() const noexcept -> float* { return static_cast<float*>(_lut.get()); }
  [[nodiscard]] auto raw() const noexcept -> pbrcu_ibl {
    pbrcu_ibl i {};
    i.irradiance = _irradiance.get(); i.irr_res = _info.irradianceResolution;
    i.prefiltered = _prefiltered.get(); i.pref_res = _info.cubeResolution; i.pref_mips = _info.mipLevels;
    i.lut = static_cast<float const*>(_lut.get()); i.lut_res = _info.lutResolution;
    return i;
  }
};
}  // namespace pbr
