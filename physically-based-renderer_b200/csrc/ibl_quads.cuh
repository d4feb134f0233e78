// ibl_quads.cuh -- "footprint record" layout of the IBL textures for the GGX shading kernel (K3).
// Record of texel (x, y) of a face = the four texels t00 = (x,y), t10 = (x+1,y), t01 = (x,y+1), t11 = (x+1,y+1)
// with x+1 / y+1 clamped to the face edge, one 32-byte sector.  Cube maps (RGBA16F, alpha unused) store
// {rg00, rg10, rg01, rg11, b00|b10, b01|b11, 0, 0} so that the 24 bytes a tap needs are six registers with
// the halves already where the filter wants them (a plain copy of the four texels made ptxas repack the
// blue halves right after the load, i.e. before the light loop that should cover the load's latency);
// the LUT stores its 4 x float2.  A bilinear tap with clamp-to-edge is then one sector: with fx = max(u*res - 0.5, 0) the footprint origin is
// floor(fx) in [0, res-1] and the weights frac(fx) are 0 where the oracle's clamp makes both taps equal,
// so the result is bit-identical to cube.cuh's four-load face_bilinear.
#pragma once
#include "cube_tex.cuh"

namespace pbrcu {
// A texture the shading kernels read with GATHERS (tld4: the four texels of a bilinear footprint, one channel per
// instruction, converted to fp32 by the texture unit) and filter with their own fp32 weights -- the unit does the
// addressing, the edge clamp and the fp16 unpack, the result is bit-identical to cube.cuh's four-load face_bilinear.
// Layout ("atlas", one 2-D array because tld4 has neither layers nor levels): every face of every level is stored with
// a one-texel gutter that repeats its edge texels (clamp-to-edge without any clamp arithmetic), i.e. in a slot of
// (res + 2)^2 texels; level 0's six slots are stacked in a first column, the slots of levels 1.. in a second one.
// Texel (i, j), i, j in [-1, res], of face f of level l lives at (ax_l + 1 + i, ay_l + f (res_l + 2) + 1 + j); the
// footprint whose origin is (i0, j0) is gathered at the corner its four texels share, (ax_l + i0 + 2, ay_l + f (res_l + 2) + j0 + 2).
struct GatherTex {
  cudaArray_t array;
  cudaTextureObject_t tex;
  uint32_t width, height;
};
struct AtlasLevel { uint32_t res, ax, ay; };
inline AtlasLevel atlas_level(uint32_t base_res, uint32_t level) {
  AtlasLevel a{base_res >> level, 0u, 0u};
  if (level >= 1) {
    a.ax = base_res + 2;
    for (uint32_t k = 1; k < level; ++k) a.ay += 6 * ((base_res >> k) + 2);
  }
  return a;
}
inline void atlas_extent(uint32_t base_res, uint32_t nmips, uint32_t faces, uint32_t& w, uint32_t& h) {
  w = base_res + 2 + (nmips > 1 ? (base_res >> 1) + 2 : 0);
  uint32_t h1 = 0;
  for (uint32_t k = 1; k < nmips; ++k) h1 += faces * ((base_res >> k) + 2);
  h = faces * (base_res + 2);
  if (h1 > h) h = h1;
}
}  // namespace pbrcu

struct pbrcu_ibl_accel {
  // gather atlases of the prefiltered chain, the irradiance map (both RGBA16F) and the LUT (RG32F, one face, one level)
  pbrcu::GatherTex g_pref, g_irr, g_lut;
  uint4* pref;      // chain_texels(pref_res, pref_mips) records, same level/face/row order as the chain
  uint4* irr;       // 6*irr_res^2 records
  uint4* lut;       // lut_res^2 records of 4 x float2
  uint32_t pref_res, pref_mips, irr_res, lut_res;
  // the prefiltered chain and the irradiance map as flagged hardware textures (cube_tex.cuh): K3 takes
  // one TEX fetch per map and falls back to the records above where the filtered flag is set
  pbrcu::CubeTex pref_tex, irr_tex;
  // the LUT as a bilinear, clamp-to-edge 2-D float2 texture (K3's single-list kernel: the LUT is smooth, the unit's
  // 8-bit weights cost at most 1/512 of the texel-to-texel range)
  cudaArray_t lut_array;
  cudaTextureObject_t lut_tex;
  size_t bytes;
};

namespace pbrcu {

struct Quad8 { uint32_t w[8]; };  // LUT record
struct Quad6 { uint32_t w[6]; };  // cube-map record (the 24 bytes in use)

// one 256-bit load; the records are re-read by neighbouring pixels, so let them allocate in L1
__device__ __forceinline__ Quad8 ldg_quad(const uint4* p) {
  Quad8 q;
  asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(q.w[0]), "=r"(q.w[1]), "=r"(q.w[2]), "=r"(q.w[3]), "=r"(q.w[4]), "=r"(q.w[5]), "=r"(q.w[6]), "=r"(q.w[7])
               : "l"(p));
  return q;
}

// the 24 bytes of a cube-map record: same sector, a 128-bit and a 64-bit load
__device__ __forceinline__ Quad6 ldg_rec(const uint4* p) {
  Quad6 q;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q.w[0]), "=r"(q.w[1]), "=r"(q.w[2]), "=r"(q.w[3]) : "l"(p));
  asm volatile("ld.global.nc.v2.u32 {%0,%1}, [%2+16];" : "=r"(q.w[4]), "=r"(q.w[5]) : "l"(p));
  return q;
}

// f16 halves of a packed word as f32, and the mixed-precision difference b(f16) - a(f32) in ONE
// instruction (sm_100 FHADD): the rounding is the same single rounding as cvt + sub.
__device__ __forceinline__ float h_lo(uint32_t w) {
  float f;
  asm("{.reg .b16 l,h; mov.b32 {l,h},%1; cvt.f32.f16 %0,l;}" : "=f"(f) : "r"(w));
  return f;
}
__device__ __forceinline__ float h_hi(uint32_t w) {
  float f;
  asm("{.reg .b16 l,h; mov.b32 {l,h},%1; cvt.f32.f16 %0,h;}" : "=f"(f) : "r"(w));
  return f;
}
__device__ __forceinline__ float hsub_lo(uint32_t w, float a) {
  float f;
  asm("{.reg .b16 l,h; mov.b32 {l,h},%1; sub.rn.f32.f16 %0,l,%2;}" : "=f"(f) : "r"(w), "f"(a));
  return f;
}
__device__ __forceinline__ float hsub_hi(uint32_t w, float a) {
  float f;
  asm("{.reg .b16 l,h; mov.b32 {l,h},%1; sub.rn.f32.f16 %0,h,%2;}" : "=f"(f) : "r"(w), "f"(a));
  return f;
}

// footprint origin and weights of a bilinear tap at normalised (u, v) in a res x res face.
// floor() and the float -> int conversion are done on the FMA pipe: for 0 <= f < 2^22, f + 1.5 * 2^23 rounded DOWN is
// exactly floor(f) + 1.5 * 2^23, whose bit pattern is 0x4B400000 + floor(f) -- FRND and F2I run on the quarter-rate XU pipe
// and sit on the dependency chain that leads to the record loads.
struct Tap { float tx, ty; int idx; };
__device__ __forceinline__ Tap make_tap(float u, float v, int res) {
  const float resf = float(res);
  constexpr float kMagic = 12582912.0f;  // 1.5 * 2^23
  const float fx = fmaxf(fmaf(u, resf, -0.5f), 0.0f), fy = fmaxf(fmaf(v, resf, -0.5f), 0.0f);
  const float bx = __fadd_rd(fx, kMagic), by = __fadd_rd(fy, kMagic);
  const float x0 = bx - kMagic, y0 = by - kMagic;  // exact
  Tap t;
  t.tx = fx - x0;
  t.ty = fy - y0;
  const int ix = __float_as_int(bx) - 0x4B400000, iy = __float_as_int(by) - 0x4B400000;
  t.idx = iy * res + ix;
  return t;
}

// bilinear of the RGB channels of one cube-map record: lerp(lerp(t00,t10,tx), lerp(t01,t11,tx), ty), lerp(a,b,t) = fma(t,b-a,a)
__device__ __forceinline__ RGB quad_bilinear_rgb(const Quad6& q, float tx, float ty) {
  RGB o;
  {
    const float a = h_lo(q.w[0]), c = h_lo(q.w[2]);
    const float top = fmaf(tx, hsub_lo(q.w[1], a), a), bot = fmaf(tx, hsub_lo(q.w[3], c), c);
    o.r = fmaf(ty, bot - top, top);
  }
  {
    const float a = h_hi(q.w[0]), c = h_hi(q.w[2]);
    const float top = fmaf(tx, hsub_hi(q.w[1], a), a), bot = fmaf(tx, hsub_hi(q.w[3], c), c);
    o.g = fmaf(ty, bot - top, top);
  }
  {
    const float a = h_lo(q.w[4]), c = h_lo(q.w[5]);
    const float top = fmaf(tx, hsub_hi(q.w[4], a), a), bot = fmaf(tx, hsub_hi(q.w[5], c), c);
    o.b = fmaf(ty, bot - top, top);
  }
  return o;
}

// bilinear of a LUT record (4 x float2)
__device__ __forceinline__ float2 quad_bilinear_lut(const Quad8& q, float tx, float ty) {
  const float ax = __uint_as_float(q.w[0]), ay = __uint_as_float(q.w[1]), cx = __uint_as_float(q.w[4]), cy = __uint_as_float(q.w[5]);
  return make_float2(lerpf(lerpf(ax, __uint_as_float(q.w[2]), tx), lerpf(cx, __uint_as_float(q.w[6]), tx), ty),
                     lerpf(lerpf(ay, __uint_as_float(q.w[3]), tx), lerpf(cy, __uint_as_float(q.w[7]), tx), ty));
}

}  // namespace pbrcu
