// pbr/Math.hpp -- the handful of glm types and functions the kept pbr:: interfaces use.
//
// The reference takes its vector maths from glm 1.0.1 (cmake/Dependencies.cmake:6, fetched from the
// network at configure time, not vendored).  glm is not on this image, so this header provides the
// subset the facade needs under the same names (namespace glm): vec2/3/4, mat4, mat3, mat3x4, quat,
// lookAtRH, perspectiveRH_NO, translate, scale, toMat4, inverse, transpose.  Matrices are column-major
// like glm (m[c][r]).  If a real glm is on the include path, define PBR_USE_SYSTEM_GLM to use it.
#pragma once
#if defined(PBR_USE_SYSTEM_GLM)
#define GLM_ENABLE_EXPERIMENTAL
#include <glm/glm.hpp>
#include <glm/gtc/matrix_transform.hpp>
#include <glm/gtx/quaternion.hpp>
#else
#include <cmath>
#include <cstddef>

namespace glm {
struct vec2 {
  float x {}, y {};
  constexpr vec2() = default;
  constexpr vec2(float a, float b) : x(a), y(b) {}
  constexpr explicit vec2(float s) : x(s), y(s) {}
};
struct vec3 {
  float x {}, y {}, z {};
  constexpr vec3() = default;
  constexpr vec3(float a, float b, float c) : x(a), y(b), z(c) {}
  constexpr explicit vec3(float s) : x(s), y(s), z(s) {}
  constexpr float& operator[](std::size_t i) { return i == 0 ? x : (i == 1 ? y : z); }
  constexpr float operator[](std::size_t i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
struct vec4 {
  float x {}, y {}, z {}, w {};
  constexpr vec4() = default;
  constexpr vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
  constexpr vec4(vec3 v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
  constexpr explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
  constexpr float& operator[](std::size_t i) { return i == 0 ? x : (i == 1 ? y : (i == 2 ? z : w)); }
  constexpr float operator[](std::size_t i) const { return i == 0 ? x : (i == 1 ? y : (i == 2 ? z : w)); }
};
constexpr vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
constexpr vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
constexpr vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
constexpr vec3 operator*(float s, vec3 a) { return a * s; }
constexpr vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
constexpr bool operator==(vec3 a, vec3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
constexpr vec4 operator+(vec4 a, vec4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
constexpr vec4 operator*(vec4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
constexpr float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
constexpr vec3 cross(vec3 a, vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float length(vec3 a) { return std::sqrt(dot(a, a)); }
inline vec3 normalize(vec3 a) { return a * (1.0f / length(a)); }

struct quat {  // w + xi + yj + zk
  float w {1.0f}, x {}, y {}, z {};
  constexpr quat() = default;
  constexpr quat(float w_, float x_, float y_, float z_) : w(w_), x(x_), y(y_), z(z_) {}
};
template <class T> constexpr T identity();
template <> constexpr quat identity<quat>() { return quat {1.0f, 0.0f, 0.0f, 0.0f}; }
inline quat angleAxis(float angle, vec3 axis) {
  const float s = std::sin(angle * 0.5f);
  return {std::cos(angle * 0.5f), axis.x * s, axis.y * s, axis.z * s};
}

struct mat3 {
  vec3 c[3] {};
  constexpr mat3() = default;
  constexpr explicit mat3(float d) : c {{d, 0, 0}, {0, d, 0}, {0, 0, d}} {}
  constexpr vec3& operator[](std::size_t i) { return c[i]; }
  constexpr vec3 const& operator[](std::size_t i) const { return c[i]; }
};
struct mat4 {
  vec4 c[4] {};
  constexpr mat4() = default;
  constexpr explicit mat4(float d) : c {{d, 0, 0, 0}, {0, d, 0, 0}, {0, 0, d, 0}, {0, 0, 0, d}} {}
  constexpr vec4& operator[](std::size_t i) { return c[i]; }
  constexpr vec4 const& operator[](std::size_t i) const { return c[i]; }
};
using mat4x4 = mat4;
using mat3x3 = mat3;
// three columns of vec4 (the std430/push-constant padding of a mat3: ModelPushConstant.hpp:15-16)
struct mat3x4 {
  vec4 c[3] {};
  constexpr mat3x4() = default;
  constexpr explicit mat3x4(mat3 const& m) : c {{m[0], 0.0f}, {m[1], 0.0f}, {m[2], 0.0f}} {}
  constexpr vec4& operator[](std::size_t i) { return c[i]; }
  constexpr vec4 const& operator[](std::size_t i) const { return c[i]; }
};
constexpr vec4 operator*(mat4 const& m, vec4 v) { return m[0] * v.x + m[1] * v.y + m[2] * v.z + m[3] * v.w; }
constexpr vec3 operator*(mat3 const& m, vec3 v) { return m[0] * v.x + m[1] * v.y + m[2] * v.z; }
constexpr mat4 operator*(mat4 const& a, mat4 const& b) {
  mat4 r;
  for (std::size_t j = 0; j < 4; ++j) r[j] = a * b[j];
  return r;
}
constexpr mat4 transpose(mat4 const& m) {
  mat4 r;
  for (std::size_t i = 0; i < 4; ++i)
    for (std::size_t j = 0; j < 4; ++j) r[i][j] = m[j][i];
  return r;
}
constexpr mat3 toMat3(mat4 const& m) {
  mat3 r;
  for (std::size_t i = 0; i < 3; ++i) r[i] = vec3 {m[i].x, m[i].y, m[i].z};
  return r;
}
// general 4x4 inverse by cofactors (what glm::inverse computes)
constexpr mat4 inverse(mat4 const& m) {
  const float a00 = m[0][0], a01 = m[0][1], a02 = m[0][2], a03 = m[0][3];
  const float a10 = m[1][0], a11 = m[1][1], a12 = m[1][2], a13 = m[1][3];
  const float a20 = m[2][0], a21 = m[2][1], a22 = m[2][2], a23 = m[2][3];
  const float a30 = m[3][0], a31 = m[3][1], a32 = m[3][2], a33 = m[3][3];
  const float b00 = a00 * a11 - a01 * a10, b01 = a00 * a12 - a02 * a10, b02 = a00 * a13 - a03 * a10;
  const float b03 = a01 * a12 - a02 * a11, b04 = a01 * a13 - a03 * a11, b05 = a02 * a13 - a03 * a12;
  const float b06 = a20 * a31 - a21 * a30, b07 = a20 * a32 - a22 * a30, b08 = a20 * a33 - a23 * a30;
  const float b09 = a21 * a32 - a22 * a31, b10 = a21 * a33 - a23 * a31, b11 = a22 * a33 - a23 * a32;
  const float det = b00 * b11 - b01 * b10 + b02 * b09 + b03 * b08 - b04 * b07 + b05 * b06;
  const float id = 1.0f / det;
  mat4 r;
  r[0] = vec4 {(a11 * b11 - a12 * b10 + a13 * b09) * id, (a02 * b10 - a01 * b11 - a03 * b09) * id,
               (a31 * b05 - a32 * b04 + a33 * b03) * id, (a22 * b04 - a21 * b05 - a23 * b03) * id};
  r[1] = vec4 {(a12 * b08 - a10 * b11 - a13 * b07) * id, (a00 * b11 - a02 * b08 + a03 * b07) * id,
               (a32 * b02 - a30 * b05 - a33 * b01) * id, (a20 * b05 - a22 * b02 + a23 * b01) * id};
  r[2] = vec4 {(a10 * b10 - a11 * b08 + a13 * b06) * id, (a01 * b08 - a00 * b10 - a03 * b06) * id,
               (a30 * b04 - a31 * b02 + a33 * b00) * id, (a21 * b02 - a20 * b04 - a23 * b00) * id};
  r[3] = vec4 {(a11 * b07 - a10 * b09 - a12 * b06) * id, (a00 * b09 - a01 * b07 + a02 * b06) * id,
               (a31 * b01 - a30 * b03 - a32 * b00) * id, (a20 * b03 - a21 * b01 + a22 * b00) * id};
  return r;
}
constexpr mat4 translate(mat4 const& m, vec3 v) {
  mat4 r = m;
  r[3] = m[0] * v.x + m[1] * v.y + m[2] * v.z + m[3];
  return r;
}
constexpr mat4 scale(mat4 const& m, vec3 v) {
  mat4 r;
  r[0] = m[0] * v.x; r[1] = m[1] * v.y; r[2] = m[2] * v.z; r[3] = m[3];
  return r;
}
constexpr mat4 toMat4(quat q) {
  const float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z, xy = q.x * q.y, xz = q.x * q.z, yz = q.y * q.z;
  const float wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z;
  mat4 r {1.0f};
  r[0] = vec4 {1 - 2 * (yy + zz), 2 * (xy + wz), 2 * (xz - wy), 0};
  r[1] = vec4 {2 * (xy - wz), 1 - 2 * (xx + zz), 2 * (yz + wx), 0};
  r[2] = vec4 {2 * (xz + wy), 2 * (yz - wx), 1 - 2 * (xx + yy), 0};
  return r;
}
// right-handed look-at: f = normalize(center - eye), s = normalize(f x up), u = s x f
inline mat4 lookAtRH(vec3 eye, vec3 center, vec3 up) {
  const vec3 f = normalize(center - eye), s = normalize(cross(f, up)), u = cross(s, f);
  mat4 r {1.0f};
  r[0] = vec4 {s.x, u.x, -f.x, 0};
  r[1] = vec4 {s.y, u.y, -f.y, 0};
  r[2] = vec4 {s.z, u.z, -f.z, 0};
  r[3] = vec4 {-dot(s, eye), -dot(u, eye), dot(f, eye), 1};
  return r;
}
// right-handed perspective, clip z in [-1, 1] ("NO"): CameraData.hpp:29 calls exactly this variant
inline mat4 perspectiveRH_NO(float fovy, float aspect, float zNear, float zFar) {
  const float t = std::tan(fovy * 0.5f);
  mat4 r;
  r[0][0] = 1.0f / (aspect * t);
  r[1][1] = 1.0f / t;
  r[2][2] = -(zFar + zNear) / (zFar - zNear);
  r[2][3] = -1.0f;
  r[3][2] = -(2.0f * zFar * zNear) / (zFar - zNear);
  return r;
}
// ---- what the viewer's CameraController and frame loop use (src/gltf_viewer/CameraController.hpp, App.cpp:301-308) ----
constexpr vec2 operator-(vec2 a, vec2 b) { return {a.x - b.x, a.y - b.y}; }
constexpr vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }
template <class T> constexpr T half_pi();
template <> constexpr float half_pi<float>() { return 1.5707963267948966192313216916397514420985846996875529104874722961f; }
// x (1 - a) + y a, component-wise
constexpr vec3 mix(vec3 x, vec3 y, float a) { return x * (1.0f - a) + y * a; }
constexpr quat operator*(quat p, quat q) {
  return {p.w * q.w - p.x * q.x - p.y * q.y - p.z * q.z, p.w * q.x + p.x * q.w + p.y * q.z - p.z * q.y,
          p.w * q.y + p.y * q.w + p.z * q.x - p.x * q.z, p.w * q.z + p.z * q.w + p.x * q.y - p.y * q.x};
}
// q * angleAxis(angle, normalised v): the axis is re-normalised only when its length is off by more than 0.001
inline quat rotate(quat q, float angle, vec3 v) {
  vec3 tmp = v;
  const float len = length(tmp);
  if (std::fabs(len - 1.0f) > 0.001f) tmp = tmp * (1.0f / len);
  const float s = std::sin(angle * 0.5f);
  return q * quat {std::cos(angle * 0.5f), tmp.x * s, tmp.y * s, tmp.z * s};
}
// m * (rotation by angle about normalize(v)), Rodrigues form
inline mat4 rotate(mat4 const& m, float angle, vec3 v) {
  const float c = std::cos(angle), s = std::sin(angle);
  const vec3 axis = normalize(v), temp = axis * (1.0f - c);
  mat4 rot;
  rot[0][0] = c + temp.x * axis.x; rot[0][1] = temp.x * axis.y + s * axis.z; rot[0][2] = temp.x * axis.z - s * axis.y;
  rot[1][0] = temp.y * axis.x - s * axis.z; rot[1][1] = c + temp.y * axis.y; rot[1][2] = temp.y * axis.z + s * axis.x;
  rot[2][0] = temp.z * axis.x + s * axis.y; rot[2][1] = temp.z * axis.y - s * axis.x; rot[2][2] = c + temp.z * axis.z;
  mat4 r;
  r[0] = m[0] * rot[0][0] + m[1] * rot[0][1] + m[2] * rot[0][2];
  r[1] = m[0] * rot[1][0] + m[1] * rot[1][1] + m[2] * rot[1][2];
  r[2] = m[0] * rot[2][0] + m[1] * rot[2][1] + m[2] * rot[2][2];
  r[3] = m[3];
  return r;
}
}  // namespace glm
#endif
