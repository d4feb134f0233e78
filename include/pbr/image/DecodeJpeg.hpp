// pbr/image/DecodeJpeg.hpp -- baseline and progressive JPEG (ITU-T T.81, Huffman-coded, 8-bit) -> R8G8B8A8.
// The second container the reference's stbi_load_from_memory(..., 4) call meets in glTF assets
// (src/pbr_engine/image/pbr/image/LoadImage.cpp:58-73).  An own decoder, written from T.81 / JFIF, that makes the
// numerical choices of the vendored stb_image v2.30 wherever the standard leaves them open, so that the pixels are the
// SAME BYTES the reference would upload:
//   * inverse DCT: the 13-bit fixed-point "islow" factorisation (column pass keeps two extra bits: +512 >> 10,
//     row pass +65536 + (128 << 17) >> 17, clamp to 0..255); a column whose AC terms are all zero is dc * 4;
//   * coefficients are de-quantised into 16-bit integers (the product wraps like a C `short`);
//   * chroma upsampling: 2x horizontal = (3 near + far + 2) >> 2 with replicated ends, 2x vertical likewise between
//     the two nearest rows, 2x2 = the separable 3:1 filter >> 4 evaluated rows first; any other ratio replicates;
//   * YCbCr -> RGB in 20-bit fixed point with the constants rounded to 12 bits (1.402, 0.71414, 0.34414, 1.772), the
//     Cb term of green truncated to its upper 16 bits; alpha = 255;
//   * three components whose ids are 'R','G','B', or an Adobe APP14 segment with transform 0 and no JFIF header, are
//     taken as RGB; four components are CMYK / YCCK per the Adobe transform flag, merged with (x*y + 128) * 257 >> 16.
// Arithmetic-coded, lossless, hierarchical and 12-bit files are rejected (stb rejects them too).  Damaged streams throw
// (stb keeps whatever it decoded).  Pinned bit-exact against the reference's stb_image build on files written by Pillow and
// OpenCV (tests/test_image_decode_cpu.py).
#pragma once
#include <array>
#include <cstdint>
#include <cstring>
#include <span>
#include <stdexcept>
#include <string>
#include <vector>

namespace pbr::image::jpeg {

struct Pixels {
  std::uint32_t width {}, height {};
  std::vector<std::uint8_t> rgba;
};

namespace detail {

[[noreturn]] inline void fail(char const* what) { throw std::runtime_error(std::string("pbr::image::decode: JPEG: ") + what); }

constexpr std::uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                      41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                      30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// canonical Huffman table (T.81 Annex C): symbols in code order, first code / first index per length
struct Huffman {
  std::array<std::uint8_t, 256> symbols {};
  std::array<int, 18> firstCode {}, firstIndex {}, count {};
  bool defined = false;
  auto build(std::uint8_t const* counts, std::uint8_t const* values, int n) -> void {
    int code = 0, index = 0;
    for (int len = 1; len <= 16; ++len) {
      count[std::size_t(len)] = counts[len - 1];
      firstCode[std::size_t(len)] = code;
      firstIndex[std::size_t(len)] = index;
      code += counts[len - 1];
      index += counts[len - 1];
      if (code > (1 << len)) fail("over-subscribed Huffman table");
      code <<= 1;
    }
    std::memcpy(symbols.data(), values, std::size_t(n));
    defined = true;
  }
};

// MSB-first reader of one entropy-coded segment: 0xFF00 is a stuffed 0xFF; any other marker ends the data (zeros follow)
class BitReader {
  std::uint8_t const* _p;
  std::uint8_t const* _end;
  std::uint32_t _acc = 0;
  int _n = 0;
  int _marker = -1;

public:
  BitReader(std::uint8_t const* p, std::uint8_t const* end) : _p(p), _end(end) {}
  [[nodiscard]] auto position() const -> std::uint8_t const* { return _p; }
  [[nodiscard]] auto marker() const -> int { return _marker; }
  auto bit() -> int {
    if (_n == 0) {
      std::uint32_t b = 0;
      if (_marker < 0 && _p < _end) {
        b = *_p++;
        if (b == 0xFF) {
          int c = _p < _end ? *_p++ : 0xD9;
          while (c == 0xFF && _p < _end) c = *_p++;
          if (c != 0) { _marker = c; b = 0; }
        }
      }
      _acc = b;
      _n = 8;
    }
    --_n;
    return int((_acc >> _n) & 1u);
  }
  auto bits(int n) -> int {
    int v = 0;
    for (int i = 0; i < n; ++i) v = (v << 1) | bit();
    return v;
  }
  // T.81 F.2.2.1 RECEIVE + EXTEND
  auto receiveExtend(int n) -> int {
    if (n == 0) return 0;
    int const v = bits(n);
    return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
  }
  auto decode(Huffman const& h) -> int {
    int code = 0;
    for (int len = 1; len <= 16; ++len) {
      code = (code << 1) | bit();
      int const off = code - h.firstCode[std::size_t(len)];
      if (off >= 0 && off < h.count[std::size_t(len)]) return h.symbols[std::size_t(h.firstIndex[std::size_t(len)] + off)];
    }
    fail("invalid Huffman code");
  }
  // drop the rest of the interval and step over the RSTn marker that must follow; false = something else follows
  auto restart() -> bool {
    _n = 0;
    if (_marker < 0) {  // the marker has not been met yet: it is the next two bytes (after optional fill)
      while (_p < _end && *_p != 0xFF) ++_p;
      while (_p < _end && *_p == 0xFF) ++_p;
      if (_p >= _end) return false;
      _marker = *_p++;
    }
    if (_marker < 0xD0 || _marker > 0xD7) return false;
    _marker = -1;
    return true;
  }
};

struct Component {
  int id {}, h {1}, v {1}, tq {}, hd {}, ha {};
  int x {}, y {};    // samples of this component in the image
  int w2 {}, h2 {};  // MCU-padded plane
  int dcPred {};
  std::vector<std::uint8_t> plane;
  std::vector<std::int16_t> coeff;  // progressive: 64 per block, (w2 / 8) blocks per row
};

inline auto clamp255(int x) -> std::uint8_t { return std::uint8_t(x < 0 ? 0 : x > 255 ? 255 : x); }

// one 1-D pass of the fixed-point inverse DCT (even part x0..x3, odd part t0..t3), constants scaled by 4096
struct Idct1D {
  int x0, x1, x2, x3, t0, t1, t2, t3;
  static constexpr int f(float c) { return int(c * 4096 + 0.5); }  // float constant, scaled exactly, rounded in double
  Idct1D(int s0, int s1, int s2, int s3, int s4, int s5, int s6, int s7) {
    int p2 = s2, p3 = s6;
    int p1 = (p2 + p3) * f(0.5411961f);
    int const e2 = p1 + p3 * f(-1.847759065f), e3 = p1 + p2 * f(0.765366865f);
    int const e0 = (s0 + s4) * 4096, e1 = (s0 - s4) * 4096;
    x0 = e0 + e3; x3 = e0 - e3; x1 = e1 + e2; x2 = e1 - e2;
    t0 = s7; t1 = s5; t2 = s3; t3 = s1;
    p3 = t0 + t2;
    int p4 = t1 + t3;
    p1 = t0 + t3;
    p2 = t1 + t2;
    int const p5 = (p3 + p4) * f(1.175875602f);
    t0 *= f(0.298631336f); t1 *= f(2.053119869f); t2 *= f(3.072711026f); t3 *= f(1.501321110f);
    p1 = p5 + p1 * f(-0.899976223f);
    p2 = p5 + p2 * f(-2.562915447f);
    p3 *= f(-1.961570560f);
    p4 *= f(-0.390180644f);
    t3 += p1 + p4; t2 += p2 + p3; t1 += p2 + p4; t0 += p1 + p3;
  }
};

inline auto idctBlock(std::uint8_t* out, int stride, std::int16_t const* d) -> void {
  int v[64];
  for (int i = 0; i < 8; ++i) {
    if (d[i + 8] == 0 && d[i + 16] == 0 && d[i + 24] == 0 && d[i + 32] == 0 && d[i + 40] == 0 && d[i + 48] == 0 && d[i + 56] == 0) {
      int const dc = d[i] * 4;
      for (int r = 0; r < 8; ++r) v[i + 8 * r] = dc;
      continue;
    }
    Idct1D const c(d[i], d[i + 8], d[i + 16], d[i + 24], d[i + 32], d[i + 40], d[i + 48], d[i + 56]);
    int const x0 = c.x0 + 512, x1 = c.x1 + 512, x2 = c.x2 + 512, x3 = c.x3 + 512;
    v[i] = (x0 + c.t3) >> 10; v[i + 56] = (x0 - c.t3) >> 10;
    v[i + 8] = (x1 + c.t2) >> 10; v[i + 48] = (x1 - c.t2) >> 10;
    v[i + 16] = (x2 + c.t1) >> 10; v[i + 40] = (x2 - c.t1) >> 10;
    v[i + 24] = (x3 + c.t0) >> 10; v[i + 32] = (x3 - c.t0) >> 10;
  }
  for (int r = 0; r < 8; ++r) {
    int const* w = v + 8 * r;
    std::uint8_t* o = out + r * stride;
    Idct1D const c(w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7]);
    int const bias = 65536 + (128 << 17);
    int const x0 = c.x0 + bias, x1 = c.x1 + bias, x2 = c.x2 + bias, x3 = c.x3 + bias;
    o[0] = clamp255((x0 + c.t3) >> 17); o[7] = clamp255((x0 - c.t3) >> 17);
    o[1] = clamp255((x1 + c.t2) >> 17); o[6] = clamp255((x1 - c.t2) >> 17);
    o[2] = clamp255((x2 + c.t1) >> 17); o[5] = clamp255((x2 - c.t1) >> 17);
    o[3] = clamp255((x3 + c.t0) >> 17); o[4] = clamp255((x3 - c.t0) >> 17);
  }
}

class Decoder {
  std::span<std::uint8_t const> _in;
  std::size_t _pos = 0;
  std::array<std::array<std::uint16_t, 64>, 4> _quant {};  // natural (row-major) order
  std::array<Huffman, 4> _dc, _ac;
  std::vector<Component> _comp;
  int _width = 0, _height = 0, _hMax = 1, _vMax = 1, _mcuX = 0, _mcuY = 0;
  bool _progressive = false, _haveFrame = false, _jfif = false;
  int _adobeTransform = -1, _rgbIds = 0, _restartInterval = 0;
  // current scan
  std::vector<int> _order;
  int _specStart = 0, _specEnd = 63, _succHigh = 0, _succLow = 0, _eobRun = 0;

  auto u8() -> int {
    if (_pos >= _in.size()) fail("truncated file");
    return _in[_pos++];
  }
  auto u16() -> int {
    int const hi = u8();
    return (hi << 8) | u8();
  }
  auto nextMarker() -> int {
    int x = u8();
    if (x != 0xFF) fail("expected a marker");
    while (x == 0xFF) x = u8();
    return x;
  }

  auto tables(int marker) -> void {
    int len = u16() - 2;
    if (len < 0) fail("bad segment length");
    if (marker == 0xDB) {
      while (len > 0) {
        int const q = u8(), precision = q >> 4, t = q & 15;
        if (precision > 1 || t > 3) fail("bad DQT");
        for (int i = 0; i < 64; ++i) _quant[std::size_t(t)][kZigzag[i]] = std::uint16_t(precision ? u16() : u8());
        len -= precision ? 129 : 65;
      }
      if (len != 0) fail("bad DQT length");
    } else if (marker == 0xC4) {
      while (len > 0) {
        int const q = u8(), tc = q >> 4, th = q & 15;
        if (tc > 1 || th > 3) fail("bad DHT");
        std::uint8_t counts[16], values[256];
        int n = 0;
        for (auto& c : counts) { c = std::uint8_t(u8()); n += c; }
        if (n > 256) fail("bad DHT");
        for (int i = 0; i < n; ++i) values[i] = std::uint8_t(u8());
        (tc == 0 ? _dc : _ac)[std::size_t(th)].build(counts, values, n);
        len -= 17 + n;
      }
      if (len != 0) fail("bad DHT length");
    } else if (marker == 0xDD) {
      if (len != 2) fail("bad DRI length");
      _restartInterval = u16();
    } else if ((marker >= 0xE0 && marker <= 0xEF) || marker == 0xFE) {
      std::size_t const end = _pos + std::size_t(len);
      if (end > _in.size()) fail("truncated segment");
      if (marker == 0xE0 && len >= 5 && std::memcmp(_in.data() + _pos, "JFIF\0", 5) == 0) _jfif = true;
      if (marker == 0xEE && len >= 12 && std::memcmp(_in.data() + _pos, "Adobe\0", 6) == 0) _adobeTransform = _in[_pos + 11];
      _pos = end;
    } else {
      fail("unsupported marker (arithmetic coding, lossless or hierarchical JPEG?)");
    }
  }

  auto frame(int marker) -> void {
    if (_haveFrame) fail("more than one frame");
    _progressive = marker == 0xC2;
    int const len = u16();
    if (u8() != 8) fail("only 8-bit samples are supported");
    _height = u16();
    _width = u16();
    int const n = u8();
    if (_width == 0 || _height == 0) fail("empty image");
    if (n != 1 && n != 3 && n != 4) fail("bad component count");
    if (len != 8 + 3 * n) fail("bad SOF length");
    _comp.resize(std::size_t(n));
    for (int i = 0; i < n; ++i) {
      auto& c = _comp[std::size_t(i)];
      c.id = u8();
      if (n == 3 && c.id == "RGB"[i]) ++_rgbIds;
      int const q = u8();
      c.h = q >> 4; c.v = q & 15; c.tq = u8();
      if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4 || c.tq > 3) fail("bad component parameters");
      _hMax = c.h > _hMax ? c.h : _hMax;
      _vMax = c.v > _vMax ? c.v : _vMax;
    }
    for (auto const& c : _comp)
      if (_hMax % c.h != 0 || _vMax % c.v != 0) fail("unsupported sampling factors");
    _mcuX = (_width + 8 * _hMax - 1) / (8 * _hMax);
    _mcuY = (_height + 8 * _vMax - 1) / (8 * _vMax);
    for (auto& c : _comp) {
      c.x = (_width * c.h + _hMax - 1) / _hMax;
      c.y = (_height * c.v + _vMax - 1) / _vMax;
      c.w2 = _mcuX * c.h * 8;
      c.h2 = _mcuY * c.v * 8;
      c.plane.assign(std::size_t(c.w2) * std::size_t(c.h2), 0);
      if (_progressive) c.coeff.assign(std::size_t(c.w2) * std::size_t(c.h2), 0);
    }
    _haveFrame = true;
  }

  // ---- sequential (baseline / extended) block: decode, de-quantise, transform ----
  auto blockSequential(BitReader& br, Component& c, std::uint8_t* out) -> void {
    std::int16_t data[64] = {};
    auto const& q = _quant[std::size_t(c.tq)];
    int const t = br.decode(_dc[std::size_t(c.hd)]);
    if (t > 15) fail("bad DC category");
    c.dcPred += br.receiveExtend(t);
    data[0] = std::int16_t(c.dcPred * q[0]);
    for (int k = 1; k < 64;) {
      int const rs = br.decode(_ac[std::size_t(c.ha)]), r = rs >> 4, s = rs & 15;
      if (s == 0) {
        if (rs != 0xF0) break;
        k += 16;
      } else {
        k += r;
        if (k > 63) fail("coefficient index out of range");
        int const zig = kZigzag[k++];
        data[zig] = std::int16_t(br.receiveExtend(s) * q[std::size_t(zig)]);
      }
    }
    idctBlock(out, c.w2, data);
  }
  // ---- progressive (T.81 Annex G) ----
  auto blockProgressiveDc(BitReader& br, Component& c, std::int16_t* data) -> void {
    if (_specEnd != 0) fail("DC scan with AC coefficients");
    if (_succHigh == 0) {
      int const t = br.decode(_dc[std::size_t(c.hd)]);
      if (t > 15) fail("bad DC category");
      c.dcPred += br.receiveExtend(t);
      data[0] = std::int16_t(c.dcPred * (1 << _succLow));
    } else if (br.bit()) {
      data[0] = std::int16_t(data[0] + (1 << _succLow));
    }
  }
  auto refine(BitReader& br, std::int16_t& coefficient, int bit) -> void {
    if (br.bit() && (coefficient & bit) == 0) coefficient = std::int16_t(coefficient > 0 ? coefficient + bit : coefficient - bit);
  }
  auto blockProgressiveAc(BitReader& br, Component& c, std::int16_t* data) -> void {
    if (_specStart == 0) fail("AC scan that starts at the DC coefficient");
    auto const& table = _ac[std::size_t(c.ha)];
    if (_succHigh == 0) {
      if (_eobRun > 0) { --_eobRun; return; }
      for (int k = _specStart; k <= _specEnd;) {
        int const rs = br.decode(table), r = rs >> 4, s = rs & 15;
        if (s == 0) {
          if (r < 15) {
            _eobRun = (1 << r) - 1 + (r ? br.bits(r) : 0);
            break;
          }
          k += 16;
        } else {
          k += r;
          if (k > 63) fail("coefficient index out of range");
          data[kZigzag[k++]] = std::int16_t(br.receiveExtend(s) * (1 << _succLow));
        }
      }
      return;
    }
    int const bit = 1 << _succLow;
    if (_eobRun > 0) {
      --_eobRun;
      for (int k = _specStart; k <= _specEnd; ++k)
        if (auto& p = data[kZigzag[k]]; p != 0) refine(br, p, bit);
      return;
    }
    for (int k = _specStart; k <= _specEnd;) {
      int const rs = br.decode(table);
      int r = rs >> 4, s = rs & 15;
      if (s == 0) {
        if (r < 15) {
          _eobRun = (1 << r) - 1 + (r ? br.bits(r) : 0);
          r = 64;  // the rest of the band holds only refinements
        }
      } else {
        if (s != 1) fail("bad refinement code");
        s = br.bit() ? bit : -bit;
      }
      while (k <= _specEnd) {
        auto& p = data[kZigzag[k++]];
        if (p != 0) {
          refine(br, p, bit);
        } else {
          if (r == 0) { p = std::int16_t(s); break; }
          --r;
        }
      }
    }
  }

  auto scan() -> void {
    int const len = u16(), n = u8();
    if (n < 1 || n > int(_comp.size()) || len != 6 + 2 * n) fail("bad SOS");
    _order.clear();
    for (int i = 0; i < n; ++i) {
      int const id = u8(), q = u8();
      int which = -1;
      for (std::size_t k = 0; k < _comp.size(); ++k)
        if (_comp[k].id == id) { which = int(k); break; }
      if (which < 0 || (q >> 4) > 3 || (q & 15) > 3) fail("bad SOS component");
      _comp[std::size_t(which)].hd = q >> 4;
      _comp[std::size_t(which)].ha = q & 15;
      _order.push_back(which);
    }
    _specStart = u8();
    _specEnd = u8();
    int const a = u8();
    _succHigh = a >> 4;
    _succLow = a & 15;
    if (_progressive) {
      if (_specStart > 63 || _specEnd > 63 || _specStart > _specEnd || _succHigh > 13 || _succLow > 13) fail("bad SOS");
    } else {
      if (_specStart != 0 || _succHigh != 0 || _succLow != 0) fail("bad SOS");
      _specEnd = 63;
    }

    BitReader br(_in.data() + _pos, _in.data() + _in.size());
    int todo = _restartInterval ? _restartInterval : 0x7FFFFFFF;
    _eobRun = 0;
    for (auto& c : _comp) c.dcPred = 0;
    bool stop = false;
    auto const endOfMcu = [&] {
      if (--todo > 0) return;
      if (!br.restart()) { stop = true; return; }  // no RSTn where one is due: keep what has been decoded
      todo = _restartInterval ? _restartInterval : 0x7FFFFFFF;
      _eobRun = 0;
      for (auto& c : _comp) c.dcPred = 0;
    };
    if (_order.size() == 1) {  // non-interleaved: the component's own block grid, row by row
      auto& c = _comp[std::size_t(_order[0])];
      int const bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3;
      for (int j = 0; j < bh && !stop; ++j)
        for (int i = 0; i < bw && !stop; ++i) {
          if (!_progressive) blockSequential(br, c, c.plane.data() + std::size_t(c.w2) * std::size_t(j) * 8 + std::size_t(i) * 8);
          else {
            std::int16_t* data = c.coeff.data() + 64 * (std::size_t(i) + std::size_t(j) * std::size_t(c.w2 / 8));
            if (_specStart == 0) blockProgressiveDc(br, c, data); else blockProgressiveAc(br, c, data);
          }
          endOfMcu();
        }
    } else {  // interleaved: MCU by MCU, h x v blocks of every component of the scan
      for (int j = 0; j < _mcuY && !stop; ++j)
        for (int i = 0; i < _mcuX && !stop; ++i) {
          for (int const which : _order) {
            auto& c = _comp[std::size_t(which)];
            for (int y = 0; y < c.v; ++y)
              for (int x = 0; x < c.h; ++x) {
                int const bx = i * c.h + x, by = j * c.v + y;
                if (!_progressive) blockSequential(br, c, c.plane.data() + std::size_t(c.w2) * std::size_t(by) * 8 + std::size_t(bx) * 8);
                else blockProgressiveDc(br, c, c.coeff.data() + 64 * (std::size_t(bx) + std::size_t(by) * std::size_t(c.w2 / 8)));
              }
          }
          endOfMcu();
        }
    }
    // continue after the entropy-coded data: at the marker the reader met, else at the next marker in the stream
    _pos = std::size_t(br.position() - _in.data());
    if (br.marker() >= 0) {
      _pos -= 2;  // re-read it as a marker (fill bytes, if any, are skipped again by nextMarker)
      while (_pos > 0 && _in[_pos] != 0xFF) --_pos;
    } else {
      while (_pos + 1 < _in.size() && !(_in[_pos] == 0xFF && _in[_pos + 1] != 0x00 && _in[_pos + 1] != 0xFF)) ++_pos;
    }
  }

  auto finishProgressive() -> void {
    for (auto& c : _comp) {
      auto const& q = _quant[std::size_t(c.tq)];
      int const bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3;
      for (int j = 0; j < bh; ++j)
        for (int i = 0; i < bw; ++i) {
          std::int16_t* data = c.coeff.data() + 64 * (std::size_t(i) + std::size_t(j) * std::size_t(c.w2 / 8));
          for (int k = 0; k < 64; ++k) data[k] = std::int16_t(data[k] * q[std::size_t(k)]);
          idctBlock(c.plane.data() + std::size_t(c.w2) * std::size_t(j) * 8 + std::size_t(i) * 8, c.w2, data);
        }
    }
  }

  // ---- upsampling of one output row of a component to full width ----
  struct Resampler {
    int hs {1}, vs {1}, ystep {}, wLores {}, ypos {};
    std::uint8_t const* line0 {};
    std::uint8_t const* line1 {};
    std::vector<std::uint8_t> buffer;
  };
  static auto resampleRow(Resampler& r, std::uint8_t const* nearRow, std::uint8_t const* farRow) -> std::uint8_t const* {
    int const w = r.wLores;
    std::uint8_t* out = r.buffer.data();
    if (r.hs == 1 && r.vs == 1) return nearRow;
    if (r.hs == 1 && r.vs == 2) {
      for (int i = 0; i < w; ++i) out[i] = std::uint8_t((3 * nearRow[i] + farRow[i] + 2) >> 2);
    } else if (r.hs == 2 && r.vs == 1) {
      if (w == 1) {
        out[0] = out[1] = nearRow[0];
      } else {
        out[0] = nearRow[0];
        out[1] = std::uint8_t((nearRow[0] * 3 + nearRow[1] + 2) >> 2);
        int i = 1;
        for (; i < w - 1; ++i) {
          int const n = 3 * nearRow[i] + 2;
          out[2 * i] = std::uint8_t((n + nearRow[i - 1]) >> 2);
          out[2 * i + 1] = std::uint8_t((n + nearRow[i + 1]) >> 2);
        }
        out[2 * i] = std::uint8_t((nearRow[w - 2] * 3 + nearRow[w - 1] + 2) >> 2);
        out[2 * i + 1] = nearRow[w - 1];
      }
    } else if (r.hs == 2 && r.vs == 2) {
      if (w == 1) {
        out[0] = out[1] = std::uint8_t((3 * nearRow[0] + farRow[0] + 2) >> 2);
      } else {
        int t1 = 3 * nearRow[0] + farRow[0];
        out[0] = std::uint8_t((t1 + 2) >> 2);
        for (int i = 1; i < w; ++i) {
          int const t0 = t1;
          t1 = 3 * nearRow[i] + farRow[i];
          out[2 * i - 1] = std::uint8_t((3 * t0 + t1 + 8) >> 4);
          out[2 * i] = std::uint8_t((3 * t1 + t0 + 8) >> 4);
        }
        out[2 * w - 1] = std::uint8_t((t1 + 2) >> 2);
      }
    } else {
      for (int i = 0; i < w; ++i)
        for (int j = 0; j < r.hs; ++j) out[i * r.hs + j] = nearRow[i];
    }
    return out;
  }

  static auto blinn(int x, int y) -> std::uint8_t {
    unsigned const t = unsigned(x * y + 128);
    return std::uint8_t((t + (t >> 8)) >> 8);
  }
  static auto ycc(std::uint8_t* out, int y, int cb, int cr) -> void {
    constexpr auto fx = [](float c) { return int(c * 4096.0f + 0.5f) << 8; };
    int const yf = (y << 20) + (1 << 19);
    cr -= 128; cb -= 128;
    int const r = (yf + cr * fx(1.40200f)) >> 20;
    int const g = (yf + cr * -fx(0.71414f) + int(unsigned(cb * -fx(0.34414f)) & 0xFFFF0000u)) >> 20;
    int const b = (yf + cb * fx(1.77200f)) >> 20;
    out[0] = clamp255(r); out[1] = clamp255(g); out[2] = clamp255(b); out[3] = 255;
  }

  auto output() -> Pixels {
    Pixels px {std::uint32_t(_width), std::uint32_t(_height), std::vector<std::uint8_t>(std::size_t(_width) * std::size_t(_height) * 4)};
    std::size_t const n = _comp.size();
    bool const isRgb = n == 3 && (_rgbIds == 3 || (_adobeTransform == 0 && !_jfif));
    std::vector<Resampler> rs(n);
    for (std::size_t k = 0; k < n; ++k) {
      auto& r = rs[k];
      r.hs = _hMax / _comp[k].h; r.vs = _vMax / _comp[k].v;
      r.ystep = r.vs >> 1;
      r.wLores = (_width + r.hs - 1) / r.hs;
      r.line0 = r.line1 = _comp[k].plane.data();
      r.buffer.resize(std::size_t(_width) + 8);
    }
    std::uint8_t const* row[4] = {};
    for (int j = 0; j < _height; ++j) {
      for (std::size_t k = 0; k < n; ++k) {
        auto& r = rs[k];
        bool const bottom = r.ystep >= (r.vs >> 1);
        row[k] = resampleRow(r, bottom ? r.line1 : r.line0, bottom ? r.line0 : r.line1);
        if (++r.ystep >= r.vs) {
          r.ystep = 0;
          r.line0 = r.line1;
          if (++r.ypos < _comp[k].y) r.line1 += _comp[k].w2;
        }
      }
      std::uint8_t* out = px.rgba.data() + std::size_t(j) * std::size_t(_width) * 4;
      for (int i = 0; i < _width; ++i, out += 4) {
        if (n == 1) {
          out[0] = out[1] = out[2] = row[0][i]; out[3] = 255;
        } else if (n == 3 && isRgb) {
          out[0] = row[0][i]; out[1] = row[1][i]; out[2] = row[2][i]; out[3] = 255;
        } else if (n == 3) {
          ycc(out, row[0][i], row[1][i], row[2][i]);
        } else if (_adobeTransform == 0) {  // CMYK
          int const m = row[3][i];
          out[0] = blinn(row[0][i], m); out[1] = blinn(row[1][i], m); out[2] = blinn(row[2][i], m); out[3] = 255;
        } else if (_adobeTransform == 2) {  // YCCK
          ycc(out, row[0][i], row[1][i], row[2][i]);
          int const m = row[3][i];
          out[0] = blinn(255 - out[0], m); out[1] = blinn(255 - out[1], m); out[2] = blinn(255 - out[2], m);
        } else {  // YCbCr + an ignored fourth channel
          ycc(out, row[0][i], row[1][i], row[2][i]);
        }
      }
    }
    return px;
  }

public:
  explicit Decoder(std::span<std::uint8_t const> bytes) : _in(bytes) {}

  auto run() -> Pixels {
    if (nextMarker() != 0xD8) fail("no SOI");
    for (;;) {
      int const m = nextMarker();
      if (m == 0xD9) break;
      if (m == 0xC0 || m == 0xC1 || m == 0xC2) frame(m);
      else if (m == 0xDA) {
        if (!_haveFrame) fail("scan before the frame header");
        scan();
        if (_pos + 1 >= _in.size()) break;  // a file that ends without EOI: take what is there
      } else if (m >= 0xD0 && m <= 0xD7) {
        continue;  // a restart marker left over after the last interval
      } else if (m == 0xDC) {
        if (u16() != 4 || u16() != _height) fail("bad DNL");
      } else {
        tables(m);
      }
    }
    if (!_haveFrame) fail("no frame");
    if (_progressive) finishProgressive();
    return output();
  }
};

}  // namespace detail

[[nodiscard]] inline auto isJpeg(std::span<std::uint8_t const> bytes) -> bool { return bytes.size() >= 3 && bytes[0] == 0xFF && bytes[1] == 0xD8 && bytes[2] == 0xFF; }
[[nodiscard]] inline auto decode(std::span<std::uint8_t const> bytes) -> Pixels { return detail::Decoder(bytes).run(); }

}  // namespace pbr::image::jpeg
