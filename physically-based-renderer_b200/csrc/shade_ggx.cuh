// shade_ggx.cuh -- parameters, staging helpers and the light / ambient arithmetic shared by the K3 kernels
// (shade_ggx.cu: tile lists and device-resident light lists; shade_ggx_cl.cu: lights in the kernel parameters).
#pragma once
#include <algorithm>

#include "f32x2.cuh"
#include "ibl_quads.cuh"
#include "tier_a_math.cuh"

namespace pbrcu {

constexpr int kTile = PBRCU_TILE;
constexpr int kHalfRows = kTile / 2;
constexpr int kShadeThreads = kTile * kHalfRows;  // a CTA iteration shades half a tile (16 x 8 pixels): finer register granularity
constexpr int kClusterThreads = kTile * kTile;    // K3c: one thread per pixel of a tile
constexpr int kShadeCtasPerSm = 5;  // <= 102 registers: the four IBL records stay in registers across the light loop
constexpr int kMaxLights = 1024;   // API limit for max_per_tile
constexpr int kLightChunk = 128;   // lights staged in shared memory at a time
constexpr float kInvPi = 0.31830988618379067154f;
constexpr int kClMaxLights = 32;   // shade_ggx_cl.cu: lights that fit the kernel's parameter block

struct ShadeParams {
  const uint8_t *pos, *nrm, *alb, *orm, *emi;
  uint8_t* out;
  size_t pitch8, pitch4, out_pitch;
  uint32_t x0, y0, x1, y1;        // rect in image coordinates (x1, y1 exclusive)
  uint32_t tx0, ty0;              // first tile of the rect
  uint32_t ntx, ntiles;           // tiles per row of the rect, tiles in the rect
  uint32_t ntx_magic;             // ceil(2^32 / ntx) (0 if ntx == 1): tile / ntx == __umulhi(tile, magic) while tile * ntx < 2^32
  float cx, cy, cz;
  const pbrcu_light* lights;
  uint32_t n_lights;
  const uint32_t* tile_counts;    // NULL: every pixel loops all lights
  const uint16_t* tile_indices;
  uint32_t tiles_x, max_per_tile;
  int has_ibl;
  CubeChain irr, pref;
  const float2* lut;
  uint32_t lut_res;
  const uint4 *q_pref, *q_irr, *q_lut;  // footprint records (pbrcu_ibl_accel); NULL: plain textures
  cudaTextureObject_t t_pref, t_irr;    // flagged hardware textures of the same maps (IBL_TEX)
  cudaTextureObject_t t_lut;            // the LUT as a bilinear float2 texture (shade_ggx_one_kernel<.., AMBTEX>)
  // batched multi-view (LIST_ONE_BATCHED): view v reads its planes view_stride8 / view_stride4 bytes after view
  // v - 1's, its camera from cams[3 v ..], and writes out_view_stride bytes after view v - 1's output
  uint32_t n_views;
  size_t view_stride8, view_stride4, out_view_stride;
  const float* cams;
  // pbrcu_shade_ggx_ibl_replicated: every finished texel is also stored at the same byte offset of these buffers
  // (peer-mapped memory of the other GPUs: the stores travel over NVLink while the kernel shades the next texels)
  uint32_t n_replicas;
  uint8_t* replica[PBRCU_MAX_REPLICAS];
  int out_is_multicast;  // pbrcu_shade_ggx_ibl_multicast: `out` is a multimem address
  uint32_t out_texel;    // bytes per output texel: 8 = RGBA16F HDR; 4 = RGBA8 LDR, the ACES + gamma tonemap (K2) fused into the store
                         // (peer-store instantiations only: pbrcu_shade_ggx_ibl_ldr)
  int pair_stores;       // every frame address and the pitch are 16-byte aligned: neighbouring texels may leave as one 16-byte store
};

// the finished texel: streamed to the local frame and to every replica; or ONE multimem store to a multicast address,
// which the NVSwitch delivers to the frame of every GPU bound to it (this one included)
template <bool PEERS>
__device__ __forceinline__ void store_texel(const ShadeParams& p, size_t off, uint2 v) {
  if (!PEERS) {  // the single-GPU kernels carry none of the bookkeeping below
    stg_stream8(p.out + off, v);
    return;
  }
  if (p.out_texel == 4) {
    // K2 fused into the store: tonemap the fp16-ROUNDED texel, exactly what pbrcu_tonemap_aces_gamma would read back from the HDR frame
    // (aces+gamma.glsl:30-40), so the LDR frame equals the two-pass result bit for bit; `off` was computed with 4-byte texels
    const uint32_t ldr = tonemap_texel(unpack_rgba16f(v));
    stg_stream4(p.out + off, ldr);
    for (uint32_t k = 0; k < p.n_replicas; ++k) stg_stream4(p.replica[k] + off, ldr);
    return;
  }
  if (p.out_is_multicast) {
    const unsigned long long bits = (static_cast<unsigned long long>(v.y) << 32) | v.x;
    asm volatile("multimem.st.weak.global.b64 [%0], %1;" ::"l"(p.out + off), "l"(bits) : "memory");
    return;
  }
  stg_stream8(p.out + off, v);
  for (uint32_t k = 0; k < p.n_replicas; ++k) stg_stream8(p.replica[k] + off, v);
}

// The same for the kernels whose lanes all reach the store together (shade_ggx_kernel): lanes 2k / 2k+1 hold the horizontally
// adjacent texels x (even) / x + 1 of one row, so where both exist the even lane stores BOTH as one 16-byte store -- half the
// packets on NVLink for the peer frames (8-byte stores reached ~510 GB/s of ingress at N = 8).  `have`: this lane owns a texel.
__device__ __forceinline__ void store_texel_paired(const ShadeParams& p, size_t off, uint2 v, bool have) {
  const unsigned full = 0xffffffffu;
  const uint32_t nx = __shfl_down_sync(full, v.x, 1), ny = __shfl_down_sync(full, v.y, 1);
  const bool right_has = __shfl_down_sync(full, int(have), 1) != 0, left_has = __shfl_up_sync(full, int(have), 1) != 0;
  const bool even = (threadIdx.x & 1u) == 0;
  if (p.pair_stores && even && have && right_has) {
    const uint4 w = make_uint4(v.x, v.y, nx, ny);
    if (p.out_is_multicast) {
      asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p.out + off), "f"(__uint_as_float(w.x)), "f"(__uint_as_float(w.y)),
                   "f"(__uint_as_float(w.z)), "f"(__uint_as_float(w.w))
                   : "memory");
      return;
    }
    stg_stream(p.out + off, w);
    for (uint32_t k = 0; k < p.n_replicas; ++k) stg_stream(p.replica[k] + off, w);
    return;
  }
  if (have && !(p.pair_stores && !even && left_has)) store_texel<true>(p, off, v);  // (an odd lane whose left neighbour stored for it is done)
}

// how a work item gets its light list
enum : int {
  LIST_TILED = 0,        // per-tile lists (K3c) or more than kLightChunk lights: staged per item, in chunks
  LIST_ONE = 1,          // no tile lists and at most kLightChunk lights: one list, staged once per CTA
  LIST_ONE_BATCHED = 2   // LIST_ONE over n_views G-buffers of one shape in a single launch
};

// how the ambient term samples the three IBL textures
enum : int {
  IBL_PLAIN = 0,    // four 8-byte loads per bilinear tap from the plain chains, fp32 weights
  IBL_RECORDS = 1,  // one 256-bit footprint-record load per tap, fp32 weights (bit-identical to IBL_PLAIN)
  IBL_RECORDS_AMBTEX = 3,  // IBL_RECORDS for the prefiltered cube; the two smooth maps -- irradiance (flagged copy, record
                           // fallback) and the LUT -- through the texture unit (see shade_ggx_one_kernel)
  IBL_TEX = 2       // texture unit for the prefiltered and irradiance cubes (8-bit weights), records where
                    // the filtered contrast flag is set and for the LUT
};

__device__ __forceinline__ float2 lut_sample(const float2* lut, int res, float u, float v) {
  const float fx = u * float(res) - 0.5f, fy = v * float(res) - 0.5f;
  const float x0f = floorf(fx), y0f = floorf(fy);
  const float tx = fx - x0f, ty = fy - y0f;
  const int hi = res - 1;
  const int x0 = min(max(int(x0f), 0), hi), x1 = min(max(int(x0f) + 1, 0), hi);
  const int y0 = min(max(int(y0f), 0), hi), y1 = min(max(int(y0f) + 1, 0), hi);
  const float2 t00 = __ldg(lut + y0 * res + x0), t10 = __ldg(lut + y0 * res + x1);
  const float2 t01 = __ldg(lut + y1 * res + x0), t11 = __ldg(lut + y1 * res + x1);
  return make_float2(lerpf(lerpf(t00.x, t10.x, tx), lerpf(t01.x, t11.x, tx), ty),
                     lerpf(lerpf(t00.y, t10.y, tx), lerpf(t01.y, t11.y, tx), ty));
}

__device__ __forceinline__ float sat(float x) { return __saturatef(x); }
// single MUFU ops (no denormal fix-up code: every argument here is a normal number or the result is
// multiplied by an exact 0)
__device__ __forceinline__ float rsq(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

// The four records a pixel's ambient term needs: irradiance(N), prefiltered(R) at two levels, LUT(NoV, rough).
struct IblTaps {
  int i_irr, i_p0, i_p1, i_lut;
  float irr_tx, irr_ty, p0_tx, p0_ty, p1_tx, p1_ty, lut_tx, lut_ty, lod_f;
};
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

__device__ __forceinline__ IblTaps make_ibl_taps(const ShadeParams& p, float3 N, float3 R, float NoV, float rough) {
  IblTaps t;
  int face;
  float u, v;
  cube_dir_to_face_uv(N, face, u, v);
  const int ires = int(p.irr.base_res);
  Tap a = make_tap(u, v, ires);
  t.i_irr = face * ires * ires + a.idx; t.irr_tx = a.tx; t.irr_ty = a.ty;
  cube_dir_to_face_uv(R, face, u, v);
  const float lod = fminf(rough * float(p.pref.nmips - 1), float(p.pref.nmips - 1));  // rough >= 0.045 > 0
  const float l0f = floorf(lod);
  t.lod_f = lod - l0f;
  const int l0 = int(l0f), l1 = min(l0 + 1, int(p.pref.nmips) - 1);
  const int r0 = int(p.pref.base_res >> l0), r1 = int(p.pref.base_res >> l1);
  a = make_tap(u, v, r0);
  t.i_p0 = int(p.pref.mip_offset[l0]) + face * r0 * r0 + a.idx; t.p0_tx = a.tx; t.p0_ty = a.ty;
  a = make_tap(u, v, r1);
  t.i_p1 = int(p.pref.mip_offset[l1]) + face * r1 * r1 + a.idx; t.p1_tx = a.tx; t.p1_ty = a.ty;
  a = make_tap(NoV, rough, int(p.lut_res));
  t.i_lut = a.idx; t.lut_tx = a.tx; t.lut_ty = a.ty;
  return t;
}

// one 256-bit load of a cube-map record (8 destination registers, 6 in use)
__device__ __forceinline__ Quad6 ldg_rec256(const uint4* p) {
  Quad6 q;
  uint32_t pad0, pad1;
  asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(q.w[0]), "=r"(q.w[1]), "=r"(q.w[2]), "=r"(q.w[3]), "=r"(q.w[4]), "=r"(q.w[5]), "=r"(pad0), "=r"(pad1)
               : "l"(p));
  return q;
}

// ---- cp.async staging of one pixel's five texels into this thread's shared-memory slots ----
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(uint32_t(__cvta_generic_to_shared(smem))), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(uint32_t(__cvta_generic_to_shared(smem))), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

struct TileStage {
  uint2 px[4][kShadeThreads];   // pos, nrm, alb, emi
  uint32_t orm[kShadeThreads];
};

// work item = half a tile: item >> 1 = tile of the rect (row-major), item & 1 = upper / lower 8 rows
struct ItemPos { uint32_t tx, ty, x, y, view; };
__device__ __forceinline__ ItemPos item_pos(const ShadeParams& p, uint32_t item) {
  const uint32_t tile = item >> 1;
  const uint32_t trow = p.ntx_magic ? __umulhi(tile, p.ntx_magic) : tile;
  ItemPos q;
  q.ty = p.ty0 + trow; q.tx = p.tx0 + (tile - trow * p.ntx);
  q.x = q.tx * kTile + (threadIdx.x % kTile);
  q.y = q.ty * kTile + (item & 1u) * kHalfRows + (threadIdx.x / kTile);
  q.view = 0;
  return q;
}

template <bool BATCHED>
__device__ __forceinline__ void stage_tile(const ShadeParams& p, TileStage& st, const ItemPos& q) {
  const uint32_t x = q.x, y = q.y;
  if (x >= p.x0 && x < p.x1 && y >= p.y0 && y < p.y1) {
    const size_t o8 = size_t(y) * p.pitch8 + size_t(x) * 8 + (BATCHED ? size_t(q.view) * p.view_stride8 : 0);
    cp_async8(&st.px[1][threadIdx.x], p.nrm + o8);
    cp_async8(&st.px[0][threadIdx.x], p.pos + o8);
    cp_async8(&st.px[2][threadIdx.x], p.alb + o8);
    cp_async8(&st.px[3][threadIdx.x], p.emi + o8);
    cp_async4(&st.orm[threadIdx.x], p.orm + size_t(y) * p.pitch4 + size_t(x) * 4 + (BATCHED ? size_t(q.view) * p.view_stride4 : 0));
  }
}

// ---- the light evaluation both K3 kernels share: KHR_lights_punctual point lights, Cook-Torrance GGX / Smith / Schlick,
// two lights per call in the lanes of packed f32x2 instructions ----
struct LightPixel {  // what the light loop reads of a pixel
  float3 P, N, V;
  float NoV, visV, a2, oma2, a2h, F0r, F0g, F0b, omF0r, omF0g, omF0b, cdpr, cdpg, cdpb;  // a2h = a2 / (2 pi), cdp = c_diff / pi
};
struct PairA { f2 dx, dy, dz, invd, NoL, w; };

// geometry of a staged light pair {{x0,x1,y0,y1}, {z0,z1,w0,w1}} (w = 1/range^2, 0 = unlimited): w = inverse square * window * N.L
__device__ __forceinline__ PairA light_pair_a(const LightPixel& s, const float4 l0, const float4 l1) {
  const f2 one = dup(1.0f);
  PairA a;
  a.dx = sub2(pk(l0.x, l0.y), dup(s.P.x)); a.dy = sub2(pk(l0.z, l0.w), dup(s.P.y)); a.dz = sub2(pk(l1.x, l1.y), dup(s.P.z));
  const f2 d2 = max2(fma2(a.dz, a.dz, fma2(a.dy, a.dy, mul2(a.dx, a.dx))), 1e-8f);
  a.invd = map2(d2, rsq);
  a.NoL = sat2(mul2(fma2(dup(s.N.z), a.dz, fma2(dup(s.N.y), a.dy, mul2(dup(s.N.x), a.dx))), a.invd));
  const f2 qq = mul2(d2, pk(l1.z, l1.w));               // (d/range)^2; 0 for unlimited range
  const f2 win = sat2(sub2(one, mul2(qq, qq)));         // clamp(1 - (d/range)^4, 0, 1)
  a.w = mul2(mul2(mul2(a.invd, a.invd), win), a.NoL);
  return a;
}

// radiance factor of the pair per channel: (1 - F) c_diff/pi + F D Vis = c_diff/pi + F (D Vis - c_diff/pi)
__device__ __forceinline__ void light_pair_b(const LightPixel& s, const PairA& a, f2& br, f2& bg, f2& bb) {
  const f2 one = dup(1.0f);
  const f2 hx = fma2(a.dx, a.invd, dup(s.V.x)), hy = fma2(a.dy, a.invd, dup(s.V.y)), hz = fma2(a.dz, a.invd, dup(s.V.z));  // V + L
  const f2 hh = max2(fma2(hz, hz, fma2(hy, hy, mul2(hx, hx))), 1e-12f);  // V = -L: grazing back-facing pixel
  const f2 nhinv = map2(hh, [](float v) { return -rsq(v); });
  // |V| = |L| = 1: V.h = 1 + V.L = |h|^2 / 2, so V.H = |h| / 2 (in [0, 1] by construction)
  const f2 VoH = mul2(mul2(hh, nhinv), dup(-0.5f));
  // 1 - NoH^2 without cancellation: sq = |N - H|^2 = 2 - 2 N.H  =>  u = 1 - (N.H)^2 = sq (1 - sq/4)
  // (what the oracle takes from |N x H|^2).  D's denominator NoH^2 (a2 - 1) + 1 with
  // NoH = max(N.H, 0) is then u (1 - a2) + a2, or just u where N.H < 0, i.e. sq > 2.
  const f2 ex = fma2(hx, nhinv, dup(s.N.x)), ey = fma2(hy, nhinv, dup(s.N.y)), ez = fma2(hz, nhinv, dup(s.N.z));
  const f2 sq = fma2(ez, ez, fma2(ey, ey, mul2(ex, ex)));
  const f2 u = fma2(mul2(sq, sq), dup(-0.25f), sq);
  const f2 ddn = fma2(u, dup(s.oma2), dup(s.a2));
  const f2 dd = pk(lo(sq) > 2.0f ? lo(u) : lo(ddn), hi(sq) > 2.0f ? hi(u) : hi(ddn));
  const f2 vl = fma2(mul2(a.NoL, a.NoL), dup(s.oma2), dup(s.a2));
  const f2 den = fma2(dup(s.NoV), mul2(vl, map2(vl, rsq)), mul2(a.NoL, dup(s.visV)));
  const f2 dv = mul2(dup(s.a2h), map2(mul2(mul2(dd, dd), den), rcp));  // D * Vis = a2/(pi dd^2) * 0.5/den
  const f2 t = sub2(one, VoH), t2 = mul2(t, t), fc = mul2(mul2(t2, t2), t);
  const f2 Fr = fma2(dup(s.omF0r), fc, dup(s.F0r)), Fg = fma2(dup(s.omF0g), fc, dup(s.F0g)), Fb = fma2(dup(s.omF0b), fc, dup(s.F0b));
  br = fma2(Fr, sub2(dv, dup(s.cdpr)), dup(s.cdpr));
  bg = fma2(Fg, sub2(dv, dup(s.cdpg)), dup(s.cdpg));
  bb = fma2(Fb, sub2(dv, dup(s.cdpb)), dup(s.cdpb));
}

// shade_ggx_cl.cu: K3 with the (host-resident) light list in the kernel parameters; p as filled by shade_common,
// amb = 0 no IBL, 1 irradiance + LUT through the texture unit's bilinear, 2 fp32 weights everywhere
int launch_shade_cl(const ShadeParams& p, const pbrcu_light* lights_host, uint32_t n_lights, const pbrcu_ibl_accel* accel, int amb, int lists,
                    int sm_count, cudaStream_t st);

}  // namespace pbrcu
