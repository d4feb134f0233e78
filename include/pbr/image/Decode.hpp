// pbr/image/Decode.hpp -- PNG and JPEG -> tightly packed R8G8B8A8 on the host.
// The reference decodes textures with the vendored stb_image (src/pbr_engine/image/pbr/image/LoadImage.cpp:58-73:
// stbi_load / stbi_load_from_memory with req_comp = 4); decode() is an own implementation of that call for the two
// containers glTF allows (PNG here, JPEG in DecodeJpeg.hpp).  The PNG decoder is written from the PNG (ISO/IEC 15948)
// and DEFLATE (RFC 1951) specifications, with stb's conversion rules to 8-bit RGBA where the specification leaves a choice:
//   * 16-bit samples keep their HIGH byte (stb_image.h: stbi__convert_16_to_8);
//   * 1/2/4-bit grey samples are scaled by 0xFF / 0x55 / 0x11; palette indices are not scaled;
//   * grey / RGB images gain alpha 255; a tRNS colour key gives alpha 0 to the matching pixels (compared at the
//     file's bit depth); a palette's tRNS entries are per-index alpha;
//   * gamma / colour-profile chunks are ignored, checksums (CRC-32, Adler-32) are not verified, rows are top-down.
// Adam7-interlaced files are supported.  The other containers stb accepts (BMP, TGA, GIF, PSD, HDR, PIC, PNM) are NOT
// decoded: decode() throws, where the reference would go on with a null pixel pointer.
// CPU-only header: tests/cpp/gltf_dump.cpp pins it against the reference's own stb_image build (oracle/_ref).
#pragma once
#include "pbr/image/DecodeJpeg.hpp"

#include <array>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <span>
#include <stdexcept>
#include <string>
#include <vector>

namespace pbr::image {

struct DecodedImage {
  std::uint32_t width {}, height {};
  std::vector<std::uint8_t> rgba;  // width * height * 4, row-major, top row first
};

namespace detail {

[[noreturn]] inline void fail(char const* what) { throw std::runtime_error(std::string("pbr::image::decode: ") + what); }

// ---- RFC 1951 inflate (inside an RFC 1950 zlib wrapper) ----
class Inflater {
  std::uint8_t const* _in;
  std::size_t _size, _pos = 0;
  std::uint64_t _bits = 0;
  int _nbits = 0;
  std::vector<std::uint8_t>& _out;

  struct Huffman {
    // canonical code: count of codes per length, symbols in code order, and a 10-bit direct lookup
    // (entry = symbol << 4 | length; 0 = longer than the table: walk the canonical ranges)
    std::array<std::uint16_t, 16> count {};
    std::array<std::uint16_t, 288> symbol {};
    std::array<std::uint16_t, 1024> fast {};
  };
  static constexpr int kFastBits = 10;

  auto refill() -> void {
    while (_nbits <= 56 && _pos < _size) {
      _bits |= std::uint64_t(_in[_pos++]) << _nbits;
      _nbits += 8;
    }
  }
  auto take(int n) -> std::uint32_t {
    if (_nbits < n) {
      refill();
      if (_nbits < n) fail("truncated deflate stream");
    }
    auto const v = std::uint32_t(_bits & ((std::uint64_t(1) << n) - 1));
    _bits >>= n;
    _nbits -= n;
    return v;
  }
  static auto reverse(std::uint32_t code, int len) -> std::uint32_t {
    std::uint32_t r = 0;
    for (int i = 0; i < len; ++i) r |= ((code >> i) & 1u) << (len - 1 - i);
    return r;
  }
  static auto build(Huffman& h, std::uint8_t const* lengths, int n) -> void {
    h.count.fill(0);
    h.fast.fill(0);
    for (int i = 0; i < n; ++i) ++h.count[lengths[i]];
    h.count[0] = 0;
    std::array<std::uint16_t, 16> offs {};
    std::array<std::uint32_t, 16> next {};
    std::uint32_t code = 0;
    int left = 1;
    for (int len = 1; len < 16; ++len) {
      left = (left << 1) - h.count[len];
      if (left < 0) fail("over-subscribed Huffman code");
      offs[len] = std::uint16_t(len == 1 ? 0 : offs[len - 1] + h.count[len - 1]);
      code = (code + (len == 1 ? 0 : h.count[len - 1])) << 1;
      next[len] = code;
    }
    for (int sym = 0; sym < n; ++sym) {
      int const len = lengths[sym];
      if (len == 0) continue;
      h.symbol[offs[len]++] = std::uint16_t(sym);
      std::uint32_t const c = next[len]++;
      if (len <= kFastBits) {
        std::uint32_t const r = reverse(c, len);
        for (std::uint32_t k = r; k < (1u << kFastBits); k += 1u << len) h.fast[k] = std::uint16_t((sym << 4) | len);
      }
    }
  }
  auto decode(Huffman const& h) -> int {
    if (_nbits < 15) refill();
    if (auto const e = h.fast[_bits & ((1u << kFastBits) - 1)]; e != 0) {
      int const len = e & 15;
      if (len > _nbits) fail("truncated deflate stream");
      _bits >>= len;
      _nbits -= len;
      return e >> 4;
    }
    // canonical walk, one bit at a time (codes longer than the lookup; rare)
    int code = 0, first = 0, index = 0;
    for (int len = 1; len < 16; ++len) {
      code |= int(take(1));
      int const cnt = h.count[len];
      if (code - cnt < first) return h.symbol[index + (code - first)];
      index += cnt;
      first = (first + cnt) << 1;
      code <<= 1;
    }
    fail("invalid Huffman code");
  }

  auto stored() -> void {
    _bits >>= _nbits & 7;  // to the next byte boundary
    _nbits -= _nbits & 7;
    std::uint32_t const len = take(16), nlen = take(16);
    if ((len ^ 0xFFFFu) != nlen) fail("corrupt stored block");
    for (std::uint32_t i = 0; i < len; ++i) _out.push_back(std::uint8_t(take(8)));
  }
  auto codes(Huffman const& lit, Huffman const& dist) -> void {
    static constexpr std::uint16_t lbase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
    static constexpr std::uint8_t lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
    static constexpr std::uint16_t dbase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
    static constexpr std::uint8_t dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
    for (;;) {
      int const sym = decode(lit);
      if (sym < 256) {
        _out.push_back(std::uint8_t(sym));
      } else if (sym == 256) {
        return;
      } else {
        if (sym > 285) fail("invalid length symbol");
        std::size_t const len = lbase[sym - 257] + take(lext[sym - 257]);
        int const ds = decode(dist);
        if (ds > 29) fail("invalid distance symbol");
        std::size_t const d = dbase[ds] + take(dext[ds]);
        if (d > _out.size()) fail("distance beyond the start of the output");
        std::size_t const from = _out.size() - d;
        for (std::size_t i = 0; i < len; ++i) _out.push_back(_out[from + i]);  // may overlap: byte by byte
      }
    }
  }
  auto fixed() -> void {
    static Huffman const* tables = [] {
      static Huffman t[2];
      std::uint8_t l[288];
      for (int i = 0; i < 144; ++i) l[i] = 8;
      for (int i = 144; i < 256; ++i) l[i] = 9;
      for (int i = 256; i < 280; ++i) l[i] = 7;
      for (int i = 280; i < 288; ++i) l[i] = 8;
      build(t[0], l, 288);
      for (int i = 0; i < 30; ++i) l[i] = 5;
      build(t[1], l, 30);
      return t;
    }();
    codes(tables[0], tables[1]);
  }
  auto dynamic() -> void {
    static constexpr std::uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    int const nlen = int(take(5)) + 257, ndist = int(take(5)) + 1, ncode = int(take(4)) + 4;
    if (nlen > 286 || ndist > 30) fail("bad dynamic block header");
    std::uint8_t lengths[320] = {};
    for (int i = 0; i < ncode; ++i) lengths[order[i]] = std::uint8_t(take(3));
    Huffman lencode;
    build(lencode, lengths, 19);
    std::memset(lengths, 0, sizeof(lengths));
    for (int i = 0; i < nlen + ndist;) {
      int const sym = decode(lencode);
      if (sym < 16) {
        lengths[i++] = std::uint8_t(sym);
      } else {
        int rep = 0;
        std::uint8_t val = 0;
        if (sym == 16) {
          if (i == 0) fail("repeat without a previous length");
          val = lengths[i - 1];
          rep = 3 + int(take(2));
        } else if (sym == 17) {
          rep = 3 + int(take(3));
        } else {
          rep = 11 + int(take(7));
        }
        if (i + rep > nlen + ndist) fail("too many code lengths");
        while (rep--) lengths[i++] = val;
      }
    }
    if (lengths[256] == 0) fail("no end-of-block code");
    Huffman lit, dist;
    build(lit, lengths, nlen);
    build(dist, lengths + nlen, ndist);
    codes(lit, dist);
  }

public:
  Inflater(std::span<std::uint8_t const> in, std::vector<std::uint8_t>& out) : _in(in.data()), _size(in.size()), _out(out) {}

  auto run() -> void {
    std::uint32_t const cmf = take(8), flg = take(8);
    if ((cmf & 15u) != 8u || ((cmf << 8) | flg) % 31u != 0u) fail("bad zlib header");
    if (flg & 0x20u) fail("zlib preset dictionary");
    for (bool last = false; !last;) {
      last = take(1) != 0;
      switch (take(2)) {
        case 0: stored(); break;
        case 1: fixed(); break;
        case 2: dynamic(); break;
        default: fail("reserved deflate block type");
      }
    }
  }
};

inline auto be32(std::uint8_t const* p) -> std::uint32_t { return (std::uint32_t(p[0]) << 24) | (std::uint32_t(p[1]) << 16) | (std::uint32_t(p[2]) << 8) | p[3]; }
inline auto paeth(int a, int b, int c) -> int {
  int const p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// Reverses the per-scanline filters of one (sub-)image in place; returns the bytes consumed.
inline auto unfilter(std::uint8_t* data, std::size_t avail, std::uint32_t w, std::uint32_t h, int bitsPerPixel) -> std::size_t {
  std::size_t const stride = (std::size_t(w) * bitsPerPixel + 7) / 8, bpp = std::size_t(bitsPerPixel + 7) / 8;
  if (w == 0 || h == 0) return 0;
  if (avail < (stride + 1) * h) fail("not enough pixel data");
  std::vector<std::uint8_t> zero(stride, 0);
  std::uint8_t const* prev = zero.data();
  for (std::uint32_t y = 0; y < h; ++y) {
    std::uint8_t* row = data + y * (stride + 1);
    int const filter = row[0];
    std::uint8_t* cur = row + 1;
    for (std::size_t i = 0; i < stride; ++i) {
      int const a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
      int add = 0;
      switch (filter) {
        case 0: break;
        case 1: add = a; break;
        case 2: add = b; break;
        case 3: add = (a + b) >> 1; break;
        case 4: add = paeth(a, b, c); break;
        default: fail("invalid scanline filter");
      }
      cur[i] = std::uint8_t(cur[i] + add);
    }
    prev = cur;
  }
  return (stride + 1) * h;
}

struct PngHeader {
  std::uint32_t width, height;
  int depth, color, interlace;
};

}  // namespace detail

[[nodiscard]] inline auto isPng(std::span<std::uint8_t const> bytes) -> bool {
  static constexpr std::uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  return bytes.size() >= 8 && std::memcmp(bytes.data(), sig, 8) == 0;
}

// The counterpart of stbi_load_from_memory(buffer, size, &w, &h, &channels, 4) for PNG and JPEG input.
[[nodiscard]] inline auto decode(std::span<std::uint8_t const> bytes) -> DecodedImage {
  using namespace detail;
  if (jpeg::isJpeg(bytes)) {
    auto px = jpeg::decode(bytes);
    return {px.width, px.height, std::move(px.rgba)};
  }
  if (!isPng(bytes)) fail("neither a PNG nor a JPEG file (the containers this build decodes)");
  PngHeader hdr {};
  bool haveHeader = false, havePalette = false;
  std::uint8_t palette[256][4];
  for (auto& e : palette) { e[0] = e[1] = e[2] = 0; e[3] = 255; }
  bool haveKey = false;
  std::uint16_t key[3] = {};
  std::vector<std::uint8_t> idat;
  for (std::size_t pos = 8;;) {
    if (pos + 8 > bytes.size()) fail("truncated PNG (no IEND)");
    std::uint32_t const len = be32(bytes.data() + pos);
    std::uint8_t const* type = bytes.data() + pos + 4;
    std::uint8_t const* body = type + 4;
    if (std::size_t(len) + 12 > bytes.size() - pos) fail("truncated PNG chunk");
    auto const is = [&](char const* t) { return std::memcmp(type, t, 4) == 0; };
    if (is("IHDR")) {
      if (len != 13 || haveHeader) fail("bad IHDR");
      hdr = {be32(body), be32(body + 4), body[8], body[9], body[12]};
      if (body[10] != 0 || body[11] != 0) fail("unknown compression / filter method");
      if (hdr.width == 0 || hdr.height == 0 || hdr.width > (1u << 24) || hdr.height > (1u << 24)) fail("bad image size");
      bool const depthOk = hdr.depth == 1 || hdr.depth == 2 || hdr.depth == 4 || hdr.depth == 8 || hdr.depth == 16;
      bool const colorOk = hdr.color == 0 || hdr.color == 2 || hdr.color == 3 || hdr.color == 4 || hdr.color == 6;
      if (!depthOk || !colorOk || hdr.interlace > 1) fail("bad IHDR fields");
      if (hdr.color == 3 && hdr.depth == 16) fail("16-bit palette");
      if ((hdr.color == 2 || hdr.color == 4 || hdr.color == 6) && hdr.depth < 8) fail("bad bit depth for the colour type");
      haveHeader = true;
    } else if (!haveHeader) {
      fail("first chunk is not IHDR");
    } else if (is("PLTE")) {
      if (len > 768 || len % 3 != 0) fail("bad PLTE");
      for (std::uint32_t i = 0; i < len / 3; ++i) { palette[i][0] = body[3 * i]; palette[i][1] = body[3 * i + 1]; palette[i][2] = body[3 * i + 2]; }
      havePalette = true;
    } else if (is("tRNS")) {
      if (hdr.color == 3) {
        if (!havePalette || len > 256) fail("bad tRNS");
        for (std::uint32_t i = 0; i < len; ++i) palette[i][3] = body[i];
      } else if (hdr.color == 0 || hdr.color == 2) {
        std::uint32_t const n = hdr.color == 0 ? 1 : 3;
        if (len != 2 * n) fail("bad tRNS");
        for (std::uint32_t i = 0; i < n; ++i) key[i] = std::uint16_t((body[2 * i] << 8) | body[2 * i + 1]);
        haveKey = true;
      } else {
        fail("tRNS with an alpha colour type");
      }
    } else if (is("IDAT")) {
      idat.insert(idat.end(), body, body + len);
    } else if (is("IEND")) {
      break;
    } else if ((type[0] & 0x20) == 0) {
      fail("unknown critical chunk");
    }
    pos += std::size_t(len) + 12;
  }
  if (hdr.color == 3 && !havePalette) fail("palette image without PLTE");

  int const channels = hdr.color == 0 ? 1 : hdr.color == 2 ? 3 : hdr.color == 3 ? 1 : hdr.color == 4 ? 2 : 4;
  int const bitsPerPixel = channels * hdr.depth;
  std::vector<std::uint8_t> raw;
  raw.reserve((std::size_t(hdr.width) * bitsPerPixel / 8 + 2) * hdr.height);
  Inflater(idat, raw).run();

  DecodedImage img {hdr.width, hdr.height, std::vector<std::uint8_t>(std::size_t(hdr.width) * hdr.height * 4)};
  int const maxv = (1 << hdr.depth) - 1;
  int const scale = hdr.depth == 1 ? 0xFF : hdr.depth == 2 ? 0x55 : hdr.depth == 4 ? 0x11 : 1;
  // one decoded sample of a (sub-)image row, as the file's integer
  auto const sample = [&](std::uint8_t const* row, std::size_t index) -> int {
    if (hdr.depth == 8) return row[index];
    if (hdr.depth == 16) return (row[2 * index] << 8) | row[2 * index + 1];
    std::size_t const bit = index * std::size_t(hdr.depth);
    return (row[bit >> 3] >> (8 - hdr.depth - int(bit & 7))) & maxv;
  };
  auto const to8 = [&](int v) -> std::uint8_t { return std::uint8_t(hdr.depth == 16 ? v >> 8 : v); };
  auto const emit = [&](std::uint8_t const* row, std::uint32_t sx, std::uint32_t x, std::uint32_t y) {
    std::uint8_t* px = img.rgba.data() + (std::size_t(y) * hdr.width + x) * 4;
    switch (hdr.color) {
      case 0: {
        int const g = sample(row, sx);
        px[0] = px[1] = px[2] = std::uint8_t(hdr.depth < 8 ? g * scale : to8(g));
        px[3] = haveKey && g == key[0] ? 0 : 255;
        break;
      }
      case 2: {
        int const r = sample(row, 3 * std::size_t(sx)), g = sample(row, 3 * std::size_t(sx) + 1), b = sample(row, 3 * std::size_t(sx) + 2);
        px[0] = to8(r); px[1] = to8(g); px[2] = to8(b);
        px[3] = haveKey && r == key[0] && g == key[1] && b == key[2] ? 0 : 255;
        break;
      }
      case 3: {
        auto const* e = palette[sample(row, sx)];
        px[0] = e[0]; px[1] = e[1]; px[2] = e[2]; px[3] = e[3];
        break;
      }
      case 4: {
        px[0] = px[1] = px[2] = to8(sample(row, 2 * std::size_t(sx)));
        px[3] = to8(sample(row, 2 * std::size_t(sx) + 1));
        break;
      }
      default: {
        for (int c = 0; c < 4; ++c) px[c] = to8(sample(row, 4 * std::size_t(sx) + c));
        break;
      }
    }
  };
  if (hdr.interlace == 0) {
    std::size_t const stride = (std::size_t(hdr.width) * bitsPerPixel + 7) / 8;
    unfilter(raw.data(), raw.size(), hdr.width, hdr.height, bitsPerPixel);
    for (std::uint32_t y = 0; y < hdr.height; ++y) {
      std::uint8_t const* row = raw.data() + y * (stride + 1) + 1;
      for (std::uint32_t x = 0; x < hdr.width; ++x) emit(row, x, x, y);
    }
  } else {
    static constexpr int x0[7] = {0, 4, 0, 2, 0, 1, 0}, y0[7] = {0, 0, 4, 0, 2, 0, 1}, dx[7] = {8, 8, 4, 4, 2, 2, 1}, dy[7] = {8, 8, 8, 4, 4, 2, 2};
    std::size_t offset = 0;
    for (int pass = 0; pass < 7; ++pass) {
      std::uint32_t const pw = (hdr.width - x0[pass] + dx[pass] - 1) / dx[pass], ph = (hdr.height - y0[pass] + dy[pass] - 1) / dy[pass];
      if (hdr.width <= std::uint32_t(x0[pass]) || hdr.height <= std::uint32_t(y0[pass]) || pw == 0 || ph == 0) continue;
      std::size_t const stride = (std::size_t(pw) * bitsPerPixel + 7) / 8;
      if (offset > raw.size()) fail("not enough pixel data");
      unfilter(raw.data() + offset, raw.size() - offset, pw, ph, bitsPerPixel);
      for (std::uint32_t j = 0; j < ph; ++j) {
        std::uint8_t const* row = raw.data() + offset + j * (stride + 1) + 1;
        for (std::uint32_t i = 0; i < pw; ++i) emit(row, i, x0[pass] + i * dx[pass], y0[pass] + j * dy[pass]);
      }
      offset += (stride + 1) * ph;
    }
  }
  return img;
}

}  // namespace pbr::image
