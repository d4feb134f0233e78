// TEST INFRASTRUCTURE ONLY -- never linked into the product library (libpbrcu.so).
//
// glsl_cpu_shim.hpp: the minimum of GLSL 4.60 needed to compile the reference's own shader
// sources, unmodified and where they lie under /root/reference/assets/shaders, as C++20 for
// the host CPU.  It is what turns `oracle/_ref/libpbr_ref_shaders.so` into "the reference's
// own code for the path, run on the host cores" (the graphics stack that would normally run
// these shaders is unavailable on this box).
//
// Semantics pinned here, because GLSL/SPIR-V leaves them open:
//   * max/min/clamp are IEEE-754 maxNum/minNum (fmaxf/fminf): max(NaN, 0) == 0.
//   * pow is libm powf (pow(x<0, y) -> NaN, which clamp() then maps to 0).
//   * normalize(v) = v * (1/sqrt(dot(v,v)))  (normalize(0) -> NaN).
//   * texture() on a G-buffer plane is a point fetch of texel floor(uv*extent): the reference
//     samples with linear/clamp-to-edge (PbrRenderSystem.cpp:68-77) but always at exact texel
//     centres (G-buffer extent == render extent, App.cpp:271-274), where linear == point.
//   * image stores convert with round-to-nearest-even (unorm8: rint(x*255)).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace glsl {

struct vec2 {
  union { struct { float x, y; }; struct { float r, g; }; };
  vec2() = default;
  constexpr vec2(float a, float b) : x(a), y(b) {}
  constexpr explicit vec2(float a) : x(a), y(a) {}
  // ivec/uvec style construction used by fullscreen_vertex.glsl
  constexpr vec2(int a, int b) : x(float(a)), y(float(b)) {}
};
struct uvec2;
struct ivec2 {
  int x, y;
  ivec2() = default;
  constexpr ivec2(int a, int b) : x(a), y(b) {}
  explicit ivec2(uvec2 const& u);
};
// gl_GlobalInvocationID.xy -> ivec2(...)
struct uvec2 { unsigned x, y; };
inline ivec2::ivec2(uvec2 const& u) : x(int(u.x)), y(int(u.y)) {}

struct vec3 {
  union { struct { float x, y, z; }; struct { float r, g, b; }; };
  vec3() = default;
  constexpr vec3(float a, float b, float c) : x(a), y(b), z(c) {}
  constexpr explicit vec3(float a) : x(a), y(a), z(a) {}
};

// plain-old-data alias used for the .xyz / .rgb swizzle members of vec4
struct vec3pod {
  float x, y, z;
  operator vec3() const { return vec3(x, y, z); }
};

struct vec4 {
  union {
    struct { float x, y, z, w; };
    struct { float r, g, b, a; };
    vec3pod xyz;
    vec3pod rgb;
  };
  vec4() = default;
  constexpr vec4(float a_, float b_, float c_, float d_) : x(a_), y(b_), z(c_), w(d_) {}
  constexpr explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
  constexpr vec4(vec3 v, float d_) : x(v.x), y(v.y), z(v.z), w(d_) {}
  constexpr vec4(vec2 v, float c_, float d_) : x(v.x), y(v.y), z(c_), w(d_) {}
};

// ---- arithmetic ------------------------------------------------------------------------
inline vec2 operator*(vec2 a, float s) { return {a.x * s, a.y * s}; }
inline vec2 operator+(vec2 a, float s) { return {a.x + s, a.y + s}; }

inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline vec3 operator+(vec3 a, float s) { return {a.x + s, a.y + s, a.z + s}; }
inline vec3 operator-(vec3 a, float s) { return {a.x - s, a.y - s, a.z - s}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator*(float s, vec3 a) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
// GLSL literals such as 0.0245786 are floats; C++ reads them as double -> narrow here
inline vec3 operator+(vec3 a, double s) { return a + float(s); }
inline vec3 operator-(vec3 a, double s) { return a - float(s); }
inline vec3 operator*(vec3 a, double s) { return a * float(s); }
inline vec3 operator*(double s, vec3 a) { return a * float(s); }
inline vec3 operator+(vec3pod a, vec3 b) { return vec3(a) + b; }
inline vec3 operator-(vec3 a, vec3pod b) { return a - vec3(b); }
inline vec3 operator*(vec3pod a, double s) { return vec3(a) * float(s); }

inline vec4 operator+(vec4 a, vec4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline vec4 operator*(vec4 a, vec4 b) { return {a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w}; }
inline vec4 operator*(vec4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline vec4 operator*(float s, vec4 a) { return a * s; }

// ---- matrices (column-major, exactly as GLSL constructs them) ------------------------------
struct mat3x3 {
  vec3 c[3];
  mat3x3() = default;
  // mat3(a,b,c, d,e,f, g,h,i): columns (a,b,c), (d,e,f), (g,h,i)
  constexpr mat3x3(double a, double b, double c_, double d, double e, double f, double g, double h,
                   double i)
      : c{vec3(float(a), float(b), float(c_)), vec3(float(d), float(e), float(f)),
          vec3(float(g), float(h), float(i))} {}
  constexpr mat3x3(vec3 c0, vec3 c1, vec3 c2) : c{c0, c1, c2} {}
};
using mat3 = mat3x3;
inline vec3 operator*(mat3x3 const& m, vec3 v) {
  // GLSL M*v = sum_j column_j * v[j]; evaluated left to right like glslang's lowering
  return vec3(m.c[0].x * v.x + m.c[1].x * v.y + m.c[2].x * v.z,
              m.c[0].y * v.x + m.c[1].y * v.y + m.c[2].y * v.z,
              m.c[0].z * v.x + m.c[1].z * v.y + m.c[2].z * v.z);
}
inline vec3 operator*(mat3x3 const& m, vec3pod v) { return m * vec3(v); }

struct mat4x4 {
  vec4 c[4];
  mat4x4() = default;
};
using mat4 = mat4x4;
inline vec4 operator*(mat4x4 const& m, vec4 v) {
  vec4 r;
  r.x = m.c[0].x * v.x + m.c[1].x * v.y + m.c[2].x * v.z + m.c[3].x * v.w;
  r.y = m.c[0].y * v.x + m.c[1].y * v.y + m.c[2].y * v.z + m.c[3].y * v.w;
  r.z = m.c[0].z * v.x + m.c[1].z * v.y + m.c[2].z * v.z + m.c[3].z * v.w;
  r.w = m.c[0].w * v.x + m.c[1].w * v.y + m.c[2].w * v.z + m.c[3].w * v.w;
  return r;
}
inline mat4x4 operator*(mat4x4 const& a, mat4x4 const& b) {
  mat4x4 r;
  for (int j = 0; j < 4; ++j) r.c[j] = a * b.c[j];
  return r;
}

// ---- built-ins ------------------------------------------------------------------------------
inline float max(float a, float b) { return std::fmax(a, b); }
inline float max(float a, double b) { return std::fmax(a, float(b)); }
inline float min(float a, float b) { return std::fmin(a, b); }
inline float clamp(float x, float lo, float hi) { return std::fmin(std::fmax(x, lo), hi); }
inline vec4 clamp(vec4 v, vec4 lo, vec4 hi) {
  return {clamp(v.x, lo.x, hi.x), clamp(v.y, lo.y, hi.y), clamp(v.z, lo.z, hi.z),
          clamp(v.w, lo.w, hi.w)};
}
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(vec3 a, vec3 b) {
  return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
inline float inversesqrt(float x) { return 1.0f / std::sqrt(x); }
inline vec3 normalize(vec3 v) { return v * inversesqrt(dot(v, v)); }
inline float pow(float x, float y) { return std::pow(x, y); }
inline float pow(float x, int y) { return std::pow(x, float(y)); }
inline vec3 pow(vec3 a, vec3 b) { return {std::pow(a.x, b.x), std::pow(a.y, b.y), std::pow(a.z, b.z)}; }
inline vec3 pow(vec3pod a, vec3 b) { return pow(vec3(a), b); }

// ---- fp16 / unorm conversions (RNE), shared by the image types below --------------------
inline float half_to_float(std::uint16_t h) {
  std::uint32_t const s = (std::uint32_t(h) & 0x8000u) << 16;
  std::uint32_t e = (h >> 10) & 0x1Fu;
  std::uint32_t m = h & 0x3FFu;
  std::uint32_t bits;
  if (e == 0) {
    if (m == 0) {
      bits = s;
    } else {  // subnormal: normalise
      int sh = 0;
      while (!(m & 0x400u)) { m <<= 1; ++sh; }
      m &= 0x3FFu;
      bits = s | (std::uint32_t(127 - 15 - sh + 1) << 23) | (m << 13);
    }
  } else if (e == 31) {
    bits = s | 0x7F800000u | (m << 13);
  } else {
    bits = s | ((e + 112u) << 23) | (m << 13);
  }
  float f;
  std::memcpy(&f, &bits, 4);
  return f;
}
inline std::uint16_t float_to_half(float f) {
  std::uint32_t x;
  std::memcpy(&x, &f, 4);
  std::uint32_t const s = (x >> 16) & 0x8000u;
  x &= 0x7FFFFFFFu;
  if (x >= 0x7F800000u) return std::uint16_t(s | 0x7C00u | ((x > 0x7F800000u) ? (0x200u | ((x >> 13) & 0x3FFu)) : 0u));
  if (x >= 0x477FF000u) return std::uint16_t(s | 0x7C00u);  // rounds to >= 65520 -> inf
  if (x < 0x33000001u) return std::uint16_t(s);              // < 2^-25 (or == 2^-25 tie -> 0)
  std::uint32_t e = x >> 23;
  std::uint32_t m = (x & 0x7FFFFFu) | 0x800000u;
  if (e < 113) {  // subnormal half
    std::uint32_t const shift = 126 - e;  // 14..24
    std::uint32_t const q = m >> shift;
    std::uint32_t const rem = m & ((1u << shift) - 1u);
    std::uint32_t const halfway = 1u << (shift - 1);
    std::uint32_t r = q;
    if (rem > halfway || (rem == halfway && (q & 1u))) ++r;
    return std::uint16_t(s | r);
  }
  std::uint32_t h = ((e - 112u) << 10) | ((m >> 13) & 0x3FFu);
  std::uint32_t const rem = m & 0x1FFFu;
  if (rem > 0x1000u || (rem == 0x1000u && (h & 1u))) ++h;  // carry into exponent is correct
  return std::uint16_t(s | h);
}

// ---- resources: a 2-D plane in host memory ----------------------------------------------
enum class Fmt { RGBA16F, RGBA16UNORM, R32F, RGBA8 };
struct Plane {
  void const* data = nullptr;
  int width = 0, height = 0;
  std::size_t pitch = 0;  // bytes
  Fmt fmt = Fmt::RGBA16F;
  bool repeat = false;    // material textures: all-zero VkSamplerCreateInfo (gltf/Asset.cpp:116-117) = nearest + repeat
};
inline vec4 fetch(Plane const& p, int x, int y) {
  auto const* row = static_cast<unsigned char const*>(p.data) + std::size_t(y) * p.pitch;
  switch (p.fmt) {
    case Fmt::RGBA16F: {
      auto const* t = reinterpret_cast<std::uint16_t const*>(row) + 4 * x;
      return {half_to_float(t[0]), half_to_float(t[1]), half_to_float(t[2]), half_to_float(t[3])};
    }
    case Fmt::RGBA16UNORM: {
      auto const* t = reinterpret_cast<std::uint16_t const*>(row) + 4 * x;
      return {float(t[0]) / 65535.0f, float(t[1]) / 65535.0f, float(t[2]) / 65535.0f,
              float(t[3]) / 65535.0f};
    }
    case Fmt::R32F: {
      auto const* t = reinterpret_cast<float const*>(row) + x;
      return {t[0], 0.0f, 0.0f, 1.0f};
    }
    case Fmt::RGBA8: {
      auto const* t = row + 4 * x;
      return {float(t[0]) / 255.0f, float(t[1]) / 255.0f, float(t[2]) / 255.0f, float(t[3]) / 255.0f};
    }
  }
  return vec4(0.0f);
}

struct texture2D { Plane plane; };
struct sampler {};
struct sampler2D {
  Plane plane;
  sampler2D() = default;
  sampler2D(texture2D const& t, sampler const&) : plane(t.plane) {}
};
inline vec4 texture(sampler2D const& s, vec2 uv) {
  int x = int(std::floor(uv.x * float(s.plane.width)));
  int y = int(std::floor(uv.y * float(s.plane.height)));
  if (s.plane.repeat) {
    x %= s.plane.width; if (x < 0) x += s.plane.width;
    y %= s.plane.height; if (y < 0) y += s.plane.height;
    return fetch(s.plane, x, y);
  }
  x = x < 0 ? 0 : (x >= s.plane.width ? s.plane.width - 1 : x);   // clamp-to-edge
  y = y < 0 ? 0 : (y >= s.plane.height ? s.plane.height - 1 : y);
  return fetch(s.plane, x, y);
}

struct image2D {
  Plane plane;       // readonly access
  void* dst = nullptr;  // writeonly access (rgba8)
  std::size_t dstPitch = 0;
};
inline ivec2 imageSize(image2D const& i) { return {i.plane.width, i.plane.height}; }
inline vec4 imageLoad(image2D const& i, ivec2 c) { return fetch(i.plane, c.x, c.y); }
inline std::uint8_t to_unorm8(float v) {
  float const c = std::fmin(std::fmax(v, 0.0f), 1.0f);  // NaN -> 0
  return std::uint8_t(std::lrint(c * 255.0f));          // RNE (default rounding mode)
}
inline void imageStore(image2D const& i, ivec2 c, vec4 v) {
  auto* t = static_cast<std::uint8_t*>(i.dst) + std::size_t(c.y) * i.dstPitch + 4 * std::size_t(c.x);
  t[0] = to_unorm8(v.x); t[1] = to_unorm8(v.y); t[2] = to_unorm8(v.z); t[3] = to_unorm8(v.w);
}

struct GlobalInvocation {
  uvec2 xy;
  unsigned z;
};

}  // namespace glsl

// ---- the storage/interface qualifiers become thread-local globals of the including TU ---------
using namespace glsl;
#define layout(...)
#define uniform static thread_local
#define in static thread_local
#define out static thread_local
#define readonly
#define writeonly
// `layout(set=0,binding=0) uniform SceneUBO { Camera cam; };` -> static thread_local anonymous union
#define SceneUBO union
#define MaterialUBO union
// `layout(push_constant) uniform PC { mat4x4 model; mat3x3 normalModel; } pc;` -> a thread-local struct variable
#define PC struct
