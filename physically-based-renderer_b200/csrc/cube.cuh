// cube.cuh -- cube-map addressing and exact-weight (fp32) filtering shared by the IBL kernels and the
// GGX shading kernel.  Conventions (DESIGN.md "Cube maps"): faces +X,-X,+Y,-Y,+Z,-Z with the
// OpenGL/Vulkan (sc, tc, ma) mapping; chain = for mip, for face, row-major RGBA16F texels; bilinear
// inside a face with clamp-to-edge; trilinear = lerp of two bilinears, lerp(a,b,t) = a + t (b - a).
#pragma once
#include "common.cuh"

namespace pbrcu {

__host__ __device__ inline size_t cube_mip_offset(uint32_t base_res, uint32_t mip) {
  size_t off = 0;
  for (uint32_t m = 0; m < mip; ++m) {
    size_t r = size_t(base_res >> m);
    off += 6 * r * r;
  }
  return off;
}

__device__ __forceinline__ float3 cube_texel_dir(int face, int x, int y, int res) {
  const float inv = 1.0f / float(res);
  const float sc = (2.0f * float(x) + 1.0f) * inv - 1.0f;
  const float tc = (2.0f * float(y) + 1.0f) * inv - 1.0f;
  switch (face) {
    case 0: return make_float3(1.0f, -tc, -sc);
    case 1: return make_float3(-1.0f, -tc, sc);
    case 2: return make_float3(sc, 1.0f, tc);
    case 3: return make_float3(sc, -1.0f, -tc);
    case 4: return make_float3(sc, -tc, 1.0f);
    default: return make_float3(-sc, -tc, -1.0f);
  }
}

__device__ __forceinline__ float3 normalize3(float3 v) {
  const float inv = rsqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
  return make_float3(v.x * inv, v.y * inv, v.z * inv);
}

__device__ __forceinline__ void cube_dir_to_face_uv(float3 d, int& face, float& u, float& v) {
  // branchless (select-only) form of the OpenGL/Vulkan major-axis rule; ties go X, then Y, then Z
  const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
  const bool isx = ax >= ay && ax >= az;
  const bool isy = !isx && ay >= az;
  const float ma = isx ? ax : (isy ? ay : az);
  const float major = isx ? d.x : (isy ? d.y : d.z);
  const bool neg = major < 0.0f;
  face = (isx ? 0 : (isy ? 2 : 4)) + (neg ? 1 : 0);
  // +X: sc=-z  -X: sc=+z  +-Y: sc=+x  +Z: sc=+x  -Z: sc=-x ;  +Y: tc=+z  -Y: tc=-z  others: tc=-y
  const float sc = isx ? (neg ? d.z : -d.z) : ((isy || !neg) ? d.x : -d.x);
  const float tc = isy ? (neg ? -d.z : d.z) : -d.y;
  const float inv = __fdividef(0.5f, ma);
  u = fmaf(sc, inv, 0.5f);
  v = fmaf(tc, inv, 0.5f);
}

__device__ __forceinline__ float lerpf(float a, float b, float t) { return fmaf(t, b - a, a); }

struct RGB { float r, g, b; };

__device__ __forceinline__ RGB texel_rgb(const uint2* p) {
  const uint2 t = __ldg(p);
  const float2 a = h2_to_f2(t.x), b = h2_to_f2(t.y);
  return RGB{a.x, a.y, b.x};
}

// bilinear in one face of one level (level = pointer to that level's first texel)
__device__ __forceinline__ RGB face_bilinear(const uint2* level, int res, int face, float u, float v) {
  const float fx = u * float(res) - 0.5f, fy = v * float(res) - 0.5f;
  const float x0f = floorf(fx), y0f = floorf(fy);
  const float tx = fx - x0f, ty = fy - y0f;
  const int hi = res - 1;
  const int x0 = min(max(int(x0f), 0), hi), x1 = min(max(int(x0f) + 1, 0), hi);
  const int y0 = min(max(int(y0f), 0), hi), y1 = min(max(int(y0f) + 1, 0), hi);
  const uint2* f = level + size_t(face) * res * res;
  const RGB t00 = texel_rgb(f + y0 * res + x0), t10 = texel_rgb(f + y0 * res + x1);
  const RGB t01 = texel_rgb(f + y1 * res + x0), t11 = texel_rgb(f + y1 * res + x1);
  RGB o;
  o.r = lerpf(lerpf(t00.r, t10.r, tx), lerpf(t01.r, t11.r, tx), ty);
  o.g = lerpf(lerpf(t00.g, t10.g, tx), lerpf(t01.g, t11.g, tx), ty);
  o.b = lerpf(lerpf(t00.b, t10.b, tx), lerpf(t01.b, t11.b, tx), ty);
  return o;
}

// Offsets (in texels) of every level of a chain, so sampling needs no loop: filled on the host.
struct CubeChain {
  const uint2* texels;
  uint32_t base_res;
  uint32_t nmips;
  uint32_t mip_offset[16];
};

inline CubeChain make_chain(const void* texels, uint32_t base_res, uint32_t nmips) {
  CubeChain c{};
  c.texels = static_cast<const uint2*>(texels);
  c.base_res = base_res;
  c.nmips = nmips;
  for (uint32_t m = 0; m < 16; ++m) c.mip_offset[m] = uint32_t(cube_mip_offset(base_res, m < nmips ? m : nmips));
  return c;
}

// textureLod(cube, dir, lod) with exact fp32 weights
__device__ __forceinline__ RGB cube_sample(const CubeChain& c, float3 dir, float lod) {
  int face;
  float u, v;
  cube_dir_to_face_uv(dir, face, u, v);
  lod = fminf(fmaxf(lod, 0.0f), float(c.nmips - 1));
  const float l0f = floorf(lod);
  const float f = lod - l0f;
  const int l0 = int(l0f);
  RGB c0 = face_bilinear(c.texels + c.mip_offset[l0], int(c.base_res >> l0), face, u, v);
  if (f > 0.0f) {
    const int l1 = min(l0 + 1, int(c.nmips) - 1);
    const RGB c1 = face_bilinear(c.texels + c.mip_offset[l1], int(c.base_res >> l1), face, u, v);
    c0.r = lerpf(c0.r, c1.r, f);
    c0.g = lerpf(c0.g, c1.g, f);
    c0.b = lerpf(c0.b, c1.b, f);
  }
  return c0;
}

}  // namespace pbrcu
