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
  bool exactFiltering {false};  // fp32 filter weights for every ambient tap of the shading pass (pbrcu_ibl.flags)
};

class Environment {
  core::SharedGpuHandle _gpu;
  EnvironmentCreateInfo _info;
  Allocation _source, _prefiltered, _irradiance, _lut;
  pbrcu_ibl_accel* _accel {};  // footprint records + texture copies the fast shading kernels sample (built by precompute)

public:
  Environment(Environment const&) = delete;
  auto operator=(Environment const&) -> Environment& = delete;
  ~Environment() { (void)pbrcu_ibl_accel_free(_gpu ? _gpu->getContext() : nullptr, _accel); }

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
    // K6 and K7 convolve the same environment: its texture copy is prepared once for both
    pbrcu_ibl_source* source = nullptr;
    core::check(pbrcu_ibl_source_create(ctx, _source.get(), _info.cubeResolution, _info.mipLevels, s, &source), "Environment: source texture");
    int const rcPrefilter = pbrcu_ibl_prefilter_ggx_prepared(ctx, source, _info.sampleCount, _prefiltered.get(), 0, _info.mipLevels - 1, 0, 1, nullptr, 0, s);
    int const rcIrradiance = pbrcu_ibl_irradiance_prepared(ctx, source, _info.irradianceResolution, _info.sampleCount, _irradiance.get(), s);
    (void)pbrcu_stream_synchronize(s);  // the source object must outlive the work that samples it
    (void)pbrcu_ibl_source_free(ctx, source);
    core::check(rcPrefilter, "Environment: prefilter_ggx");
    core::check(rcIrradiance, "Environment: irradiance");
    core::check(pbrcu_ibl_brdf_lut(ctx, _info.lutResolution, _info.sampleCount, static_cast<float*>(_lut.get()), s), "Environment: brdf_lut");
    // the records / texture copies of the finished textures (a snapshot: rebuilt by every precompute)
    (void)pbrcu_ibl_accel_free(ctx, _accel);
    _accel = nullptr;
    auto const plain = raw();
    core::check(pbrcu_ibl_accel_build(ctx, &plain, &_accel, s), "Environment: footprint records");
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
    i.accel = _accel;
    i.flags = _info.exactFiltering ? PBRCU_IBL_EXACT : 0u;
    return i;
  }
};
}  // namespace pbr
