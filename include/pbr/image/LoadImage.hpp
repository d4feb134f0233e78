// pbr/image/LoadImage.hpp -- texture files / embedded images -> device Image2D (R8G8B8A8_UNORM), staged through the
// TransferStager.  Mirrors src/pbr_engine/image/pbr/image/LoadImage.hpp:12-18 and LoadImage.cpp:20-73: decode to four
// 8-bit channels (stbi_load / stbi_load_from_memory with req_comp 4 there; pbr/image/Decode.hpp here, PNG only),
// copy into a byte vector and hand it to TransferStager::addTransfer with a 2-D, one-level, one-layer image info.
// Difference: an undecodable file throws std::runtime_error (the reference passes stb's null pointer on).
#pragma once
#include "pbr/TransferStager.hpp"
#include "pbr/image/Decode.hpp"

#include <filesystem>
#include <fstream>
#include <iterator>

namespace pbr::image {
namespace detail {
[[nodiscard]] inline auto stageImage2D(TransferStager& stager, DecodedImage const& image) -> Image2D {
  std::vector<std::byte> pixels(image.rgba.size());
  std::memcpy(pixels.data(), image.rgba.data(), pixels.size());
  return stager.addTransfer(std::move(pixels), vk::ImageCreateInfo {.format = vk::Format::eR8G8B8A8Unorm,
                                                                    .extent {.width = image.width, .height = image.height, .depth = 1},
                                                                    .mipLevels = 1,
                                                                    .arrayLayers = 1});
}
}  // namespace detail

[[nodiscard]] inline auto loadImage2D(core::GpuHandle const& /*gpu*/, TransferStager& stager, std::span<std::uint8_t const> buffer) -> Image2D {
  return detail::stageImage2D(stager, decode(buffer));
}
[[nodiscard]] inline auto loadImage2D(core::GpuHandle const& gpu, TransferStager& stager, std::filesystem::path const& path) -> Image2D {
  std::ifstream in(path, std::ios::binary);
  if (!in) throw std::runtime_error("pbr::image::loadImage2D: cannot open " + path.string());
  std::vector<std::uint8_t> const bytes((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
  return loadImage2D(gpu, stager, std::span<std::uint8_t const>(bytes));
}
}  // namespace pbr::image
