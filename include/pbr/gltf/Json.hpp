// pbr/gltf/Json.hpp -- a small RFC 8259 JSON reader for the glTF import path (the reference parses glTF with
// fastgltf 0.8.0 + simdjson, cmake/Dependencies.cmake:8, neither of which is available here).  CPU-only.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <memory>
#include <stdexcept>
#include <string>
#include <string_view>
#include <utility>
#include <vector>

namespace pbr::gltf::json {

class Value {
public:
  enum class Kind { Null, Bool, Number, String, Array, Object };
  using Member = std::pair<std::string, Value>;

private:
  Kind _kind = Kind::Null;
  bool _bool = false;
  double _number = 0.0;
  std::string _string;
  std::vector<Value> _array;
  std::vector<Member> _object;  // insertion order kept; duplicate keys: the first wins on lookup
  friend class Parser;

public:
  [[nodiscard]] auto kind() const noexcept -> Kind { return _kind; }
  [[nodiscard]] auto isNull() const noexcept -> bool { return _kind == Kind::Null; }
  [[nodiscard]] auto isObject() const noexcept -> bool { return _kind == Kind::Object; }
  [[nodiscard]] auto isArray() const noexcept -> bool { return _kind == Kind::Array; }

  // object member or nullptr
  [[nodiscard]] auto find(std::string_view key) const noexcept -> Value const* {
    if (_kind != Kind::Object) return nullptr;
    for (auto const& m : _object)
      if (m.first == key) return &m.second;
    return nullptr;
  }
  [[nodiscard]] auto members() const -> std::vector<Member> const& { return _object; }
  [[nodiscard]] auto items() const -> std::vector<Value> const& {
    static std::vector<Value> const none;
    return _kind == Kind::Array ? _array : none;
  }
  [[nodiscard]] auto size() const noexcept -> std::size_t { return _kind == Kind::Array ? _array.size() : 0; }
  [[nodiscard]] auto at(std::size_t i) const -> Value const& {
    if (_kind != Kind::Array || i >= _array.size()) throw std::out_of_range("pbr::gltf: JSON array index out of range");
    return _array[i];
  }
  [[nodiscard]] auto number() const -> double {
    if (_kind != Kind::Number) throw std::runtime_error("pbr::gltf: JSON value is not a number");
    return _number;
  }
  [[nodiscard]] auto index() const -> std::size_t {
    double const v = number();
    if (!(v >= 0.0) || v != std::floor(v) || v > 9.0e15) throw std::runtime_error("pbr::gltf: JSON value is not a non-negative integer");
    return static_cast<std::size_t>(v);
  }
  [[nodiscard]] auto string() const -> std::string const& {
    if (_kind != Kind::String) throw std::runtime_error("pbr::gltf: JSON value is not a string");
    return _string;
  }
  [[nodiscard]] auto boolean() const -> bool {
    if (_kind != Kind::Bool) throw std::runtime_error("pbr::gltf: JSON value is not a boolean");
    return _bool;
  }
  // helpers with defaults, for optional properties
  [[nodiscard]] auto numberOr(std::string_view key, double def) const -> double {
    auto const* v = find(key);
    return v ? v->number() : def;
  }
  [[nodiscard]] auto indexOr(std::string_view key, std::size_t def) const -> std::size_t {
    auto const* v = find(key);
    return v ? v->index() : def;
  }
  [[nodiscard]] auto stringOr(std::string_view key, std::string def) const -> std::string {
    auto const* v = find(key);
    return v ? v->string() : def;
  }
};

class Parser {
  std::string_view _s;
  std::size_t _p = 0;
  int _depth = 0;

  [[noreturn]] auto fail(char const* what) const -> void {
    throw std::runtime_error("pbr::gltf: JSON parse error at byte " + std::to_string(_p) + ": " + what);
  }
  auto ws() -> void {
    while (_p < _s.size() && (_s[_p] == ' ' || _s[_p] == '\t' || _s[_p] == '\n' || _s[_p] == '\r')) ++_p;
  }
  auto literal(std::string_view word) -> bool {
    if (_s.substr(_p, word.size()) != word) return false;
    _p += word.size();
    return true;
  }
  auto hex4() -> unsigned {
    if (_p + 4 > _s.size()) fail("truncated \\u escape");
    unsigned v = 0;
    for (int i = 0; i < 4; ++i) {
      char const c = _s[_p++];
      v <<= 4;
      if (c >= '0' && c <= '9') v |= unsigned(c - '0');
      else if (c >= 'a' && c <= 'f') v |= unsigned(c - 'a' + 10);
      else if (c >= 'A' && c <= 'F') v |= unsigned(c - 'A' + 10);
      else fail("bad \\u escape");
    }
    return v;
  }
  static auto utf8(std::string& out, unsigned cp) -> void {
    if (cp < 0x80) {
      out += char(cp);
    } else if (cp < 0x800) {
      out += char(0xC0 | (cp >> 6)); out += char(0x80 | (cp & 0x3F));
    } else if (cp < 0x10000) {
      out += char(0xE0 | (cp >> 12)); out += char(0x80 | ((cp >> 6) & 0x3F)); out += char(0x80 | (cp & 0x3F));
    } else {
      out += char(0xF0 | (cp >> 18)); out += char(0x80 | ((cp >> 12) & 0x3F)); out += char(0x80 | ((cp >> 6) & 0x3F)); out += char(0x80 | (cp & 0x3F));
    }
  }
  auto string() -> std::string {
    std::string out;
    ++_p;  // opening quote
    for (;;) {
      if (_p >= _s.size()) fail("unterminated string");
      unsigned char const c = static_cast<unsigned char>(_s[_p++]);
      if (c == '"') return out;
      if (c < 0x20) fail("control character in string");
      if (c != '\\') { out += char(c); continue; }
      if (_p >= _s.size()) fail("unterminated escape");
      switch (_s[_p++]) {
        case '"': out += '"'; break;
        case '\\': out += '\\'; break;
        case '/': out += '/'; break;
        case 'b': out += '\b'; break;
        case 'f': out += '\f'; break;
        case 'n': out += '\n'; break;
        case 'r': out += '\r'; break;
        case 't': out += '\t'; break;
        case 'u': {
          unsigned cp = hex4();
          if (cp >= 0xD800 && cp < 0xDC00 && _s.substr(_p, 2) == "\\u") {  // surrogate pair
            _p += 2;
            unsigned const lo = hex4();
            if (lo < 0xDC00 || lo > 0xDFFF) fail("bad surrogate pair");
            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
          }
          utf8(out, cp);
          break;
        }
        default: fail("unknown escape");
      }
    }
  }
  auto number() -> double {
    std::size_t const start = _p;
    if (_p < _s.size() && _s[_p] == '-') ++_p;
    if (_p >= _s.size() || _s[_p] < '0' || _s[_p] > '9') fail("bad number");
    if (_s[_p] == '0') ++_p; else while (_p < _s.size() && _s[_p] >= '0' && _s[_p] <= '9') ++_p;
    if (_p < _s.size() && _s[_p] == '.') {
      ++_p;
      if (_p >= _s.size() || _s[_p] < '0' || _s[_p] > '9') fail("bad fraction");
      while (_p < _s.size() && _s[_p] >= '0' && _s[_p] <= '9') ++_p;
    }
    if (_p < _s.size() && (_s[_p] == 'e' || _s[_p] == 'E')) {
      ++_p;
      if (_p < _s.size() && (_s[_p] == '+' || _s[_p] == '-')) ++_p;
      if (_p >= _s.size() || _s[_p] < '0' || _s[_p] > '9') fail("bad exponent");
      while (_p < _s.size() && _s[_p] >= '0' && _s[_p] <= '9') ++_p;
    }
    return std::strtod(std::string(_s.substr(start, _p - start)).c_str(), nullptr);  // correctly rounded, locale "C" digits only
  }
  auto value() -> Value {
    if (++_depth > 256) fail("nesting too deep");
    ws();
    if (_p >= _s.size()) fail("unexpected end of input");
    Value v;
    char const c = _s[_p];
    if (c == '{') {
      v._kind = Value::Kind::Object;
      ++_p;
      ws();
      if (_p < _s.size() && _s[_p] == '}') {
        ++_p;
      } else {
        for (;;) {
          ws();
          if (_p >= _s.size() || _s[_p] != '"') fail("expected a member name");
          std::string key = string();
          ws();
          if (_p >= _s.size() || _s[_p] != ':') fail("expected ':'");
          ++_p;
          v._object.emplace_back(std::move(key), value());
          ws();
          if (_p < _s.size() && _s[_p] == ',') { ++_p; continue; }
          if (_p < _s.size() && _s[_p] == '}') { ++_p; break; }
          fail("expected ',' or '}'");
        }
      }
    } else if (c == '[') {
      v._kind = Value::Kind::Array;
      ++_p;
      ws();
      if (_p < _s.size() && _s[_p] == ']') {
        ++_p;
      } else {
        for (;;) {
          v._array.push_back(value());
          ws();
          if (_p < _s.size() && _s[_p] == ',') { ++_p; continue; }
          if (_p < _s.size() && _s[_p] == ']') { ++_p; break; }
          fail("expected ',' or ']'");
        }
      }
    } else if (c == '"') {
      v._kind = Value::Kind::String;
      v._string = string();
    } else if (c == 't' && literal("true")) {
      v._kind = Value::Kind::Bool; v._bool = true;
    } else if (c == 'f' && literal("false")) {
      v._kind = Value::Kind::Bool; v._bool = false;
    } else if (c == 'n' && literal("null")) {
      v._kind = Value::Kind::Null;
    } else if (c == '-' || (c >= '0' && c <= '9')) {
      v._kind = Value::Kind::Number;
      v._number = number();
    } else {
      fail("unexpected character");
    }
    --_depth;
    return v;
  }

public:
  explicit Parser(std::string_view text) : _s(text) {}
  auto parse() -> Value {
    if (_s.substr(0, 3) == "\xEF\xBB\xBF") _p = 3;  // UTF-8 byte order mark
    Value v = value();
    ws();
    if (_p != _s.size()) fail("trailing characters");
    return v;
  }
};

[[nodiscard]] inline auto parse(std::string_view text) -> Value { return Parser(text).parse(); }

}  // namespace pbr::gltf::json
