// WriteImage.hpp -- frame outputs of the headless viewer (in place of the swapchain present): an 8-bit RGBA PNG
// (stored-deflate blocks: no compressor needed, any PNG reader opens it) and a PFM of the linear HDR frame.
#pragma once
#include <cstdint>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace app {
namespace detail {
inline auto crc32(std::uint8_t const* data, std::size_t n, std::uint32_t crc = 0) -> std::uint32_t {
  static std::uint32_t const* table = [] {
    static std::uint32_t t[256];
    for (std::uint32_t i = 0; i < 256; ++i) {
      std::uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
      t[i] = c;
    }
    return t;
  }();
  crc = ~crc;
  for (std::size_t i = 0; i < n; ++i) crc = table[(crc ^ data[i]) & 0xFFu] ^ (crc >> 8);
  return ~crc;
}
inline auto be32(std::vector<std::uint8_t>& out, std::uint32_t v) -> void {
  out.push_back(std::uint8_t(v >> 24)); out.push_back(std::uint8_t(v >> 16)); out.push_back(std::uint8_t(v >> 8)); out.push_back(std::uint8_t(v));
}
inline auto chunk(std::vector<std::uint8_t>& out, char const* type, std::vector<std::uint8_t> const& body) -> void {
  be32(out, std::uint32_t(body.size()));
  std::size_t const start = out.size();
  out.insert(out.end(), type, type + 4);
  out.insert(out.end(), body.begin(), body.end());
  be32(out, crc32(out.data() + start, out.size() - start));
}
}  // namespace detail

// rgba: width * height * 4 bytes, top row first
inline auto writePng(std::string const& path, std::uint8_t const* rgba, std::uint32_t width, std::uint32_t height) -> void {
  using namespace detail;
  std::vector<std::uint8_t> raw;  // filter byte 0 + row
  raw.reserve((std::size_t(width) * 4 + 1) * height);
  for (std::uint32_t y = 0; y < height; ++y) {
    raw.push_back(0);
    raw.insert(raw.end(), rgba + std::size_t(y) * width * 4, rgba + std::size_t(y + 1) * width * 4);
  }
  std::vector<std::uint8_t> z {0x78, 0x01};  // zlib header, then stored blocks of at most 65535 bytes
  std::uint32_t a = 1, b = 0;                 // Adler-32
  for (std::size_t pos = 0; pos < raw.size() || pos == 0;) {
    std::size_t const n = std::min<std::size_t>(65535, raw.size() - pos);
    z.push_back(pos + n >= raw.size() ? 1 : 0);
    z.push_back(std::uint8_t(n)); z.push_back(std::uint8_t(n >> 8)); z.push_back(std::uint8_t(~n)); z.push_back(std::uint8_t(~n >> 8));
    z.insert(z.end(), raw.begin() + std::ptrdiff_t(pos), raw.begin() + std::ptrdiff_t(pos + n));
    for (std::size_t i = pos; i < pos + n; ++i) { a = (a + raw[i]) % 65521u; b = (b + a) % 65521u; }
    pos += n;
    if (n == 0) break;
  }
  be32(z, (b << 16) | a);
  std::vector<std::uint8_t> out {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  std::vector<std::uint8_t> ihdr;
  be32(ihdr, width); be32(ihdr, height);
  ihdr.insert(ihdr.end(), {8, 6, 0, 0, 0});
  chunk(out, "IHDR", ihdr);
  chunk(out, "IDAT", z);
  chunk(out, "IEND", {});
  std::ofstream f(path, std::ios::binary);
  if (!f) throw std::runtime_error("cannot write " + path);
  f.write(reinterpret_cast<char const*>(out.data()), std::streamsize(out.size()));
}

// rgb: width * height * 3 floats, top row first (PFM stores bottom-up; negative scale = little-endian)
inline auto writePfm(std::string const& path, float const* rgb, std::uint32_t width, std::uint32_t height) -> void {
  std::ofstream f(path, std::ios::binary);
  if (!f) throw std::runtime_error("cannot write " + path);
  f << "PF\n" << width << " " << height << "\n-1.0\n";
  for (std::uint32_t y = height; y-- > 0;) f.write(reinterpret_cast<char const*>(rgb + std::size_t(y) * width * 3), std::streamsize(width) * 12);
}
}  // namespace app
