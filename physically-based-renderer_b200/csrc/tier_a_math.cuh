// tier_a_math.cuh -- per-texel device math of the reference's two passes.
//   blinnphong_texel : assets/shaders/pbr/lighting.glsl:19-29 + BlinnPhong.lib.glsl:4-10
//   tonemap_texel    : assets/shaders/tonemappers/aces+gamma.glsl:10-40 + Gamma.lib.glsl:4-6
// All arithmetic fp32; one RNE rounding at each store (fp16 / unorm8).
#pragma once
#include "common.cuh"
#include "f32x2.cuh"

namespace pbrcu {

struct F4 { float x, y, z, w; };

// Single MUFU ops with flush-to-zero: the libdevice forms (rsqrtf, __fdividef, __log2f, exp2f) wrap each MUFU in a
// compare and two predicated rescales for subnormal arguments / results, a fifth of the fused pass's instructions.
// What flushing changes: a subnormal squared distance (camera within 1e-19 of a texel's position without coinciding
// with it) normalises like a zero vector, and colours below 2^-126 map to exactly 0 -- both allowed to the reference's
// own shaders (SPIR-V without DenormPreserve) and far below the LDR quantum.
__device__ __forceinline__ float mufu_rsq(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mufu_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mufu_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mufu_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

// R16G16B16A16_UNORM decode without the conversion unit (I2F shares the MUFU pipe, the fused pass's busiest):
// the 16 bits become the low mantissa of 2^23, so (2^23 + u) - 2^23 = u exactly.
__device__ __forceinline__ float u16_lo_f(uint32_t packed) { return __uint_as_float(__byte_perm(packed, 0x4B000000u, 0x7610)) - 8388608.0f; }
__device__ __forceinline__ float u16_hi_f(uint32_t packed) { return __uint_as_float(__byte_perm(packed, 0x4B000000u, 0x7632)) - 8388608.0f; }

// UNORM8 encode of four channels, again without F2I: clamp (NaN -> 0), scale, and add 1.5 * 2^23 -- the sum's
// unit in the last place is 1, so the add IS the round-to-nearest-even and the byte sits in the low mantissa bits
// (same two roundings as lrintf(c * 255.0f)).
__device__ __forceinline__ uint32_t unorm8_bits(float v) { return __float_as_uint(__fadd_rn(__fmul_rn(__saturatef(v), 255.0f), 12582912.0f)); }
__device__ __forceinline__ uint32_t pack_unorm8x4(float r, float g, float b, float a) {
  const uint32_t rg = __byte_perm(unorm8_bits(r), unorm8_bits(g), 0x0040);
  const uint32_t ba = __byte_perm(unorm8_bits(b), unorm8_bits(a), 0x0040);
  return __byte_perm(rg, ba, 0x5410);
}

// lighting.glsl:19-29.  pos/nrm: RGBA16F texel (two packed half2 words); alb: RGBA16_UNORM texel.
// With L = V (:26), H = normalize(V+V); the shader's operation order is kept so NaN propagation for
// normalize(0) (cleared texel seen from a camera at the origin) ends in max(NaN,0) = 0 like the
// oracle: fmaxf is IEEE maxNum.
__device__ __forceinline__ F4 blinnphong_texel(uint2 pos, uint2 nrm, uint2 alb, float cx, float cy,
                                               float cz) {
  float2 pxy = h2_to_f2(pos.x), pzw = h2_to_f2(pos.y);
  float2 nxy = h2_to_f2(nrm.x), nzw = h2_to_f2(nrm.y);
  float dx = cx - pxy.x, dy = cy - pxy.y, dz = cz - pzw.x;
  float inv = mufu_rsq(dx * dx + dy * dy + dz * dz);                // :25 normalize(cam - P)
  float vx = dx * inv, vy = dy * inv, vz = dz * inv;
  float ndv = fmaxf(nxy.x * vx + nxy.y * vy + nzw.x * vz, 0.0f);   // BlinnPhong.lib.glsl:5
  float hx = vx + vx, hy = vy + vy, hz = vz + vz;                  // :6 H = normalize(V + L)
  float hinv = mufu_rsq(hx * hx + hy * hy + hz * hz);
  hx *= hinv; hy *= hinv; hz *= hinv;
  float ndh = fmaxf(nxy.x * hx + nxy.y * hy + nzw.x * hz, 0.0f);   // :7
  float s = ndh * ndh;                                             // :8 pow(NdotH, 64) = 6 squarings
  s *= s; s *= s; s *= s; s *= s; s *= s;
  F4 o;                                                            // :9 all four channels
  const float k = 1.0f / 65535.0f;
  o.x = (u16_lo_f(alb.x) * k) * ndv + s;
  o.y = (u16_hi_f(alb.x) * k) * ndv + s;
  o.z = (u16_lo_f(alb.y) * k) * ndv + s;
  o.w = (u16_hi_f(alb.y) * k) * ndv + s;
  return o;
}

__device__ __forceinline__ uint2 pack_rgba16f(F4 c) {
  return make_uint2(f2_to_h2(c.x, c.y), f2_to_h2(c.z, c.w));
}
__device__ __forceinline__ F4 unpack_rgba16f(uint2 t) {
  float2 a = h2_to_f2(t.x), b = h2_to_f2(t.y);
  return F4{a.x, a.y, b.x, b.y};
}

// aces+gamma.glsl:20-24.  Every multiply-add below is an EXPLICIT fma in the order tonemap_pair (the packed two-texel form further down) uses,
// so a texel gets the same bits whichever form shades it -- the scalar tail of a row, the vector body, or K3's fused LDR store (shade_ggx.cuh).
__device__ __forceinline__ float rrt_odt_fit(float c) {
  const float num = fmaf(c, c + 0.0245786f, -0.000090537f);
  const float den = fmaf(c, fmaf(0.983729f, c, 0.4329510f), 0.238081f);
  return num * mufu_rcp(den);  // den >= 0.238 for c >= 0 and at most ~4e9 for fp16 inputs: no range fix-up needed
}
// Gamma.lib.glsl:4-6 with the pinned negative-base behaviour: pow(max(x,0), 1/2.2).
// x = 0 -> lg2 = -inf -> ex2(-inf) = 0.
__device__ __forceinline__ float gamma22(float x) {
  return mufu_ex2(mufu_lg2(fmaxf(x, 0.0f)) * (1.0f / 2.2f));
}
// aces+gamma.glsl:30-40.  mat3x3(a,b,c,d,e,f,g,h,i) fills COLUMNS: (M v).x = a v.x + d v.y + g v.z.
__device__ __forceinline__ uint32_t tonemap_texel_scalar(F4 hdr) {
  const float ax = fmaf(0.02840f, hdr.z, fmaf(0.07600f, hdr.y, 0.59719f * hdr.x));
  const float ay = fmaf(0.13383f, hdr.z, fmaf(0.90834f, hdr.y, 0.35458f * hdr.x));
  const float az = fmaf(0.83777f, hdr.z, fmaf(0.01566f, hdr.y, 0.04823f * hdr.x));
  const float fx = rrt_odt_fit(ax), fy = rrt_odt_fit(ay), fz = rrt_odt_fit(az);
  const float ox = fmaf(-0.00327f, fz, fmaf(-0.10208f, fy, 1.60475f * fx));
  const float oy = fmaf(-0.07276f, fz, fmaf(1.10813f, fy, -0.53108f * fx));
  const float oz = fmaf(1.07602f, fz, fmaf(-0.00605f, fy, -0.07367f * fx));
  return pack_unorm8x4(gamma22(ox), gamma22(oy), gamma22(oz), hdr.w);
}

// ---- the same two passes for a PAIR of neighbouring texels, texel 0 in the low and texel 1 in the high lane of
// packed f32x2 instructions (f32x2.cuh): each lane is rounded like the scalar instruction, and the pair costs
// little more than half the issue slots of two scalar texels -- the fused pass is bound by issue slots and the
// MUFU pipe, not by HBM, until this is done.  Operation order as in blinnphong_texel / tonemap_texel.
struct F4x2 { f2 x, y, z, w; };

__device__ __forceinline__ F4x2 blinnphong_pair(uint4 pos, uint4 nrm, uint4 alb, float cx, float cy, float cz) {
  const float2 p0xy = h2_to_f2(pos.x), p0zw = h2_to_f2(pos.y), p1xy = h2_to_f2(pos.z), p1zw = h2_to_f2(pos.w);
  const float2 n0xy = h2_to_f2(nrm.x), n0zw = h2_to_f2(nrm.y), n1xy = h2_to_f2(nrm.z), n1zw = h2_to_f2(nrm.w);
  const f2 nx = pk(n0xy.x, n1xy.x), ny = pk(n0xy.y, n1xy.y), nz = pk(n0zw.x, n1zw.x);
  const f2 dx = sub2(dup(cx), pk(p0xy.x, p1xy.x)), dy = sub2(dup(cy), pk(p0xy.y, p1xy.y)), dz = sub2(dup(cz), pk(p0zw.x, p1zw.x));
  const f2 inv = map2(fma2(dz, dz, fma2(dy, dy, mul2(dx, dx))), mufu_rsq);          // :25 normalize(cam - P)
  const f2 vx = mul2(dx, inv), vy = mul2(dy, inv), vz = mul2(dz, inv);
  const f2 ndv = max2(fma2(nz, vz, fma2(ny, vy, mul2(nx, vx))), 0.0f);              // BlinnPhong.lib.glsl:5
  f2 hx = add2(vx, vx), hy = add2(vy, vy), hz = add2(vz, vz);                        // :6 H = normalize(V + L)
  const f2 hinv = map2(fma2(hz, hz, fma2(hy, hy, mul2(hx, hx))), mufu_rsq);
  hx = mul2(hx, hinv); hy = mul2(hy, hinv); hz = mul2(hz, hinv);
  const f2 ndh = max2(fma2(nz, hz, fma2(ny, hy, mul2(nx, hx))), 0.0f);              // :7
  f2 s = mul2(ndh, ndh);                                                             // :8 pow(NdotH, 64) = 6 squarings
  s = mul2(s, s); s = mul2(s, s); s = mul2(s, s); s = mul2(s, s); s = mul2(s, s);
  const f2 k = dup(1.0f / 65535.0f), m = dup(8388608.0f);
  auto u16 = [&](uint32_t w0, uint32_t w1, uint32_t sel) {
    return mul2(sub2(pk(__uint_as_float(__byte_perm(w0, 0x4B000000u, sel)), __uint_as_float(__byte_perm(w1, 0x4B000000u, sel))), m), k);
  };
  F4x2 o;                                                                            // :9 all four channels
  o.x = fma2(u16(alb.x, alb.z, 0x7610), ndv, s);
  o.y = fma2(u16(alb.x, alb.z, 0x7632), ndv, s);
  o.z = fma2(u16(alb.y, alb.w, 0x7610), ndv, s);
  o.w = fma2(u16(alb.y, alb.w, 0x7632), ndv, s);
  return o;
}

// two RGBA16F texels {texel 0 words, texel 1 words}
__device__ __forceinline__ uint4 pack_rgba16f_pair(const F4x2& c) {
  return make_uint4(f2_to_h2(lo(c.x), lo(c.y)), f2_to_h2(lo(c.z), lo(c.w)), f2_to_h2(hi(c.x), hi(c.y)), f2_to_h2(hi(c.z), hi(c.w)));
}
__device__ __forceinline__ F4x2 unpack_rgba16f_pair(uint4 t) {
  const float2 a0 = h2_to_f2(t.x), b0 = h2_to_f2(t.y), a1 = h2_to_f2(t.z), b1 = h2_to_f2(t.w);
  return F4x2{pk(a0.x, a1.x), pk(a0.y, a1.y), pk(b0.x, b1.x), pk(b0.y, b1.y)};
}

__device__ __forceinline__ f2 rrt_odt_fit2(f2 c) {
  const f2 num = fma2(c, add2(c, dup(0.0245786f)), dup(-0.000090537f));
  const f2 den = fma2(c, fma2(dup(0.983729f), c, dup(0.4329510f)), dup(0.238081f));
  return mul2(num, map2(den, mufu_rcp));
}
__device__ __forceinline__ f2 gamma22_2(f2 x) {
  return map2(mul2(map2(max2(x, 0.0f), mufu_lg2), dup(1.0f / 2.2f)), mufu_ex2);
}
__device__ __forceinline__ f2 unorm8_bits2(f2 v) { return add2(mul2(sat2(v), dup(255.0f)), dup(12582912.0f)); }

// two RGBA8 texels
__device__ __forceinline__ uint2 tonemap_pair(const F4x2& hdr) {
  const f2 ax = fma2(dup(0.02840f), hdr.z, fma2(dup(0.07600f), hdr.y, mul2(dup(0.59719f), hdr.x)));
  const f2 ay = fma2(dup(0.13383f), hdr.z, fma2(dup(0.90834f), hdr.y, mul2(dup(0.35458f), hdr.x)));
  const f2 az = fma2(dup(0.83777f), hdr.z, fma2(dup(0.01566f), hdr.y, mul2(dup(0.04823f), hdr.x)));
  const f2 fx = rrt_odt_fit2(ax), fy = rrt_odt_fit2(ay), fz = rrt_odt_fit2(az);
  const f2 ox = fma2(dup(-0.00327f), fz, fma2(dup(-0.10208f), fy, mul2(dup(1.60475f), fx)));
  const f2 oy = fma2(dup(-0.07276f), fz, fma2(dup(1.10813f), fy, mul2(dup(-0.53108f), fx)));
  const f2 oz = fma2(dup(1.07602f), fz, fma2(dup(-0.00605f), fy, mul2(dup(-0.07367f), fx)));
  const f2 r = unorm8_bits2(gamma22_2(ox)), g = unorm8_bits2(gamma22_2(oy)), b = unorm8_bits2(gamma22_2(oz)), a = unorm8_bits2(hdr.w);
  auto pack = [](float r_, float g_, float b_, float a_) {
    const uint32_t rg = __byte_perm(__float_as_uint(r_), __float_as_uint(g_), 0x0040);
    const uint32_t ba = __byte_perm(__float_as_uint(b_), __float_as_uint(a_), 0x0040);
    return __byte_perm(rg, ba, 0x5410);
  };
  return make_uint2(pack(lo(r), lo(g), lo(b), lo(a)), pack(hi(r), hi(g), hi(b), hi(a)));
}

// One texel, CANONICAL form: the packed code above on a duplicated texel.  The packed and the scalar arithmetic agree except on exact
// rounding ties of the UNORM8 store (measured: 5 bytes of a full-HD frame, 1 LSB -- scratch/ldr_diag.py), so every path that tonemaps a
// single texel -- the tail of a row, an image that is not 16-byte aligned, K3's fused LDR store (shade_ggx.cuh) -- goes through the packed
// code too and a pixel's bits do not depend on which path shaded it.  (tonemap_texel_scalar is kept for reference and for the diagnosis.)
__device__ __forceinline__ uint32_t tonemap_texel(F4 hdr) {
  return tonemap_pair(F4x2{dup(hdr.x), dup(hdr.y), dup(hdr.z), dup(hdr.w)}).x;
}

}  // namespace pbrcu
