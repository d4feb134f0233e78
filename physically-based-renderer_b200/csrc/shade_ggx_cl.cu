// shade_ggx_cl.cu -- K3 for ONE light list that the HOST holds (at most kClMaxLights point lights; the headline 4K frame,
// the batched views, the facade's render): same formulas as shade_ggx.cu, laid out for what ncu showed to bound that
// kernel -- its instruction count (profiles/r01_k3_one_kernel_ncu.txt: 119.4 M warp instructions per 4K frame, issue
// slots 67 % busy, memory 11 % of the time):
//  * the lights live in the kernel's PARAMETER BLOCK (constant bank): every light quantity is an instruction operand,
//    there is no staging, no shared-memory load and -- for the fixed-count instantiations -- no loop; the evaluation is
//    scalar fp32 (a packed f32x2 instruction takes two issue slots on sm_100a, profiles/r01_ubench_issue.txt), the
//    lights of a pixel are independent chains the scheduler interleaves, and a light that reaches no lane of the warp
//    (N.L <= 0 or out of range) costs its 17-instruction front half only;
//  * the prefiltered cube, and in exact mode the irradiance cube and the LUT, are read with texture GATHERS (tld4 on the
//    atlases of ibl_quads.cuh): the texture unit does the addressing, the clamp and the fp16 -> fp32 unpack, the kernel
//    keeps its own fp32 weights, so the result is bit-identical to cube.cuh's four-load filter -- no footprint records
//    (1.07 GB), no 256-bit scattered loads, no HADD2.F32 unpacks, no tap-index arithmetic;
//  * integer -> float conversions (UNORM16 / UNORM8 decode, floor) run on the FMA / ALU pipes (2^23 mantissa trick)
//    instead of the quarter-rate XU pipe they would share with the light loop's MUFU ops.
// No reference counterpart; checked against oracle/pbr_oracle_ibl.c (pbro_shade_pixel).
#include <cstdlib>

#include "shade_ggx.cuh"

namespace pbrcu {

constexpr int kClCtasPerSm = 5;

struct ClLight { float x, y, z, invr2, r, g, b, pad; };  // invr2 = 1/range^2 (0 = unlimited), rgb = colour * intensity

struct ShadeClParams {
  ShadeParams p;
  cudaTextureObject_t g_pref, g_irr, g_lut;  // gather atlases (ibl_quads.cuh)
  float4 lv[17];   // per level of the prefiltered atlas: {res, ax + 2, ay + 2, res + 2}; entry [nmips] repeats the last level
  float4 irr_lv;   // the irradiance atlas' only level, same form
  float lut_res, nm1;  // float(lut_res), float(nmips - 1)
  uint32_t n_lights;
  uint32_t pad;
  ClLight lights[kClMaxLights];
};

constexpr float kMagic23 = 12582912.0f;  // 1.5 * 2^23: x + kMagic23 rounded DOWN is floor(x) + kMagic23 for |x| < 2^22

// footprint origin (as a float, in [-1, res - 1]) and fp32 weights of the bilinear tap at normalised (u, v): the same
// fx = u res - 0.5, floor, frac as cube.cuh's face_bilinear; floor() on the FMA pipe (FRND runs on the XU pipe)
struct TapF { float x0, y0, tx, ty; };
__device__ __forceinline__ TapF tap_floor(float u, float v, float resf) {
  const float fx = fmaf(u, resf, -0.5f), fy = fmaf(v, resf, -0.5f);
  TapF t;
  t.x0 = __fadd_rd(fx, kMagic23) - kMagic23;
  t.y0 = __fadd_rd(fy, kMagic23) - kMagic23;
  t.tx = fx - t.x0;
  t.ty = fy - t.y0;
  return t;
}

// cube_dir_to_face_uv with the face as a float (it only feeds the atlas row): same selects, same u, v
__device__ __forceinline__ void cube_dir_to_facef_uv(float3 d, float& facef, float& u, float& v) {
  const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
  const bool isx = ax >= ay && ax >= az;
  const bool isy = !isx && ay >= az;
  const float ma = isx ? ax : (isy ? ay : az);
  const float major = isx ? d.x : (isy ? d.y : d.z);
  const bool neg = major < 0.0f;
  facef = (isx ? 0.0f : (isy ? 2.0f : 4.0f)) + (neg ? 1.0f : 0.0f);
  const float sc = isx ? (neg ? d.z : -d.z) : ((isy || !neg) ? d.x : -d.x);
  const float tc = isy ? (neg ? -d.z : d.z) : -d.y;
  const float inv = 0.5f * rcp(ma);  // ma >= 1/sqrt(3) for a unit vector: no denormal guard needed
  u = fmaf(sc, inv, 0.5f);
  v = fmaf(tc, inv, 0.5f);
}

// tld4 returns the footprint's texels as x = (i0, j0+1), y = (i0+1, j0+1), z = (i0+1, j0), w = (i0, j0)
// lerp(lerp(t00, t10, tx), lerp(t01, t11, tx), ty) with lerp(a, b, t) = fma(t, b - a, a): cube.cuh's order
__device__ __forceinline__ float bilerp4(float4 g, float tx, float ty) {
  const float top = fmaf(tx, g.z - g.w, g.w), bot = fmaf(tx, g.y - g.x, g.x);
  return fmaf(ty, bot - top, top);
}

// UNORM decode on the FMA pipe: 0x4B000000 | v is the float 2^23 + v, so fma(that, k, -2^23 k) = v k, rounded once
__device__ __forceinline__ float unorm_bits(uint32_t v, float k) { return fmaf(__uint_as_float(0x4B000000u | v), k, -8388608.0f * k); }

// one KHR_lights_punctual point light, Cook-Torrance GGX / height-correlated Smith / Schlick: the arithmetic of
// light_pair_a / light_pair_b (shade_ggx.cuh), one light per call, the light's quantities from the constant bank
template <bool BRANCH>
__device__ __forceinline__ void cl_light(const LightPixel& s, const ClLight& L, float& dr, float& dg, float& db) {
  const float dx = L.x - s.P.x, dy = L.y - s.P.y, dz = L.z - s.P.z;
  const float d2 = fmaxf(fmaf(dz, dz, fmaf(dy, dy, dx * dx)), 1e-8f);
  const float invd = rsq(d2);
  const float NoL = sat(fmaf(s.N.z, dz, fmaf(s.N.y, dy, s.N.x * dx)) * invd);
  const float qq = d2 * L.invr2;                      // (d/range)^2; 0 for unlimited range
  const float win = sat(fmaf(-qq, qq, 1.0f));         // clamp(1 - (d/range)^4, 0, 1)
  const float w = ((invd * invd) * win) * NoL;        // inverse square * window * N.L
  if (!BRANCH || w > 0.0f) {
    const float hx = fmaf(dx, invd, s.V.x), hy = fmaf(dy, invd, s.V.y), hz = fmaf(dz, invd, s.V.z);  // V + L
    const float hh = fmaxf(fmaf(hz, hz, fmaf(hy, hy, hx * hx)), 1e-12f);  // V = -L: grazing back-facing pixel
    const float nhinv = -rsq(hh);
    // |V| = |L| = 1: V.h = 1 + V.L = |h|^2 / 2, so V.H = |h| / 2 (in [0, 1] by construction)
    const float VoH = (hh * nhinv) * -0.5f;
    // 1 - NoH^2 without cancellation: sq = |N - H|^2 = 2 - 2 N.H  =>  u = 1 - (N.H)^2 = sq (1 - sq/4)
    const float ex = fmaf(hx, nhinv, s.N.x), ey = fmaf(hy, nhinv, s.N.y), ez = fmaf(hz, nhinv, s.N.z);
    const float sq = fmaf(ez, ez, fmaf(ey, ey, ex * ex));
    const float u = fmaf(sq * sq, -0.25f, sq);
    const float dd = sq > 2.0f ? u : fmaf(u, s.oma2, s.a2);
    const float vl = fmaf(NoL * NoL, s.oma2, s.a2);
    const float den = fmaf(s.NoV, vl * rsq(vl), NoL * s.visV);
    const float dv = s.a2h * rcp((dd * dd) * den);  // D * Vis = a2/(pi dd^2) * 0.5/den
    const float t = 1.0f - VoH, t2 = t * t, fc = (t2 * t2) * t;
    const float Fr = fmaf(s.omF0r, fc, s.F0r), Fg = fmaf(s.omF0g, fc, s.F0g), Fb = fmaf(s.omF0b, fc, s.F0b);
    // (1 - F) c_diff/pi + F D Vis = c_diff/pi + F (D Vis - c_diff/pi)
    dr = fmaf(fmaf(Fr, dv - s.cdpr, s.cdpr), L.r * w, dr);
    dg = fmaf(fmaf(Fg, dv - s.cdpg, s.cdpg), L.g * w, dg);
    db = fmaf(fmaf(Fb, dv - s.cdpb, s.cdpb), L.b * w, db);
  }
}

// AMB: 0 = no IBL; 1 = prefiltered cube by gathers + fp32 weights, irradiance and LUT through the texture unit's own
// bilinear (8-bit weights; an irradiance footprint whose contrast flag is set is re-filtered by gathers); 2 = all four
// taps by gathers + fp32 weights (bit-identical to cube.cuh)
// NL: number of lights evaluated, fully unrolled (the list is padded with null lights); -1 = run-time count
template <int NL, int AMB, bool BATCHED, bool PEERS, bool BRANCH, int CTAS, bool LATE = false>
__global__ void __launch_bounds__(kShadeThreads, CTAS) shade_ggx_cl_kernel(const __grid_constant__ ShadeClParams q) {
  const ShadeParams& p = q.p;
  constexpr bool REC = AMB >= 3, AMBTEX = AMB == 1 || AMB == 3;  // 3 / 4: modes 1 / 2 with footprint records in place of gathers
  __shared__ TileStage s_tile[2];
  __shared__ float4 s_lv[17];
  __shared__ float4 s_amb[AMBTEX ? 2 : 3][kShadeThreads];  // {p0_tx, p0_ty, p1_tx, p1_ty}, {lod_f, cdr, cdg, cdb}, {irr_tx, irr_ty, lut_tx, lut_ty}
  if (AMB != 0 && !REC && threadIdx.x < 17) s_lv[threadIdx.x] = q.lv[threadIdx.x];

  const uint32_t nitems = 2 * p.ntiles;  // per view
  uint32_t tile = blockIdx.x, view = 0;
  auto carry = [&](uint32_t& t, uint32_t& v) {
    if (BATCHED)
      while (t >= nitems && v < p.n_views) { t -= nitems; ++v; }
  };
  auto live = [&](uint32_t t, uint32_t v) { return BATCHED ? v < p.n_views : t < nitems; };
  carry(tile, view);
  int stage = 0;
  ItemPos next = item_pos(p, tile);
  next.view = view;
  if (live(tile, view)) stage_tile<BATCHED>(p, s_tile[0], next);
  cp_async_commit();
  __syncthreads();

  uint32_t ntile = tile, nview = view;
  for (; live(tile, view); tile = ntile, view = nview, stage ^= 1) {
    const ItemPos ip = next;
    ntile = tile + gridDim.x; nview = view;
    carry(ntile, nview);
    if (live(ntile, nview)) {  // next item in flight
      next = item_pos(p, ntile);
      next.view = nview;
      stage_tile<BATCHED>(p, s_tile[stage ^ 1], next);
    }
    cp_async_commit();
    cp_async_wait<1>();  // this item's group has landed (each thread reads only its own slots)

    const uint32_t x = ip.x, y = ip.y;
    if (!(x >= p.x0 && x < p.x1 && y >= p.y0 && y < p.y1)) continue;
    const size_t outo = size_t(y) * p.out_pitch + size_t(x) * (PEERS ? p.out_texel : 8u) + (BATCHED ? size_t(ip.view) * p.out_view_stride : 0);
    const TileStage& st = s_tile[stage];
    const uint2 tn = st.px[1][threadIdx.x];
    const float2 nxy = h2_to_f2(tn.x), nzw = h2_to_f2(tn.y);
    if (nxy.x == 0.0f && nxy.y == 0.0f && nzw.x == 0.0f) {  // all-zero normal = cleared texel
      store_texel<PEERS>(p, outo, make_uint2(0u, 0u));
      continue;
    }
    // ---- per-pixel set-up (glTF 2.0 Appendix B material model) ----
    LightPixel s;
    const uint2 tp = st.px[0][threadIdx.x];
    const uint32_t to = st.orm[threadIdx.x];
    const float2 pxy = h2_to_f2(tp.x), pzw = h2_to_f2(tp.y);
    s.P = make_float3(pxy.x, pxy.y, pzw.x);
    const float ninv = rsq(nxy.x * nxy.x + nxy.y * nxy.y + nzw.x * nzw.x);
    s.N = make_float3(nxy.x * ninv, nxy.y * ninv, nzw.x * ninv);
    float cx = p.cx, cy = p.cy, cz = p.cz;
    if (BATCHED) { cx = __ldg(p.cams + 3 * ip.view); cy = __ldg(p.cams + 3 * ip.view + 1); cz = __ldg(p.cams + 3 * ip.view + 2); }
    const float vx = cx - s.P.x, vy = cy - s.P.y, vz = cz - s.P.z;
    const float vinv = rsq(vx * vx + vy * vy + vz * vz);
    s.V = make_float3(vx * vinv, vy * vinv, vz * vinv);
    const float NoVraw = s.N.x * s.V.x + s.N.y * s.V.y + s.N.z * s.V.z;
    s.NoV = fminf(fmaxf(NoVraw, 1e-4f), 1.0f);
    const float rough = fminf(fmaxf(unorm_bits((to >> 8) & 0xFFu, 1.0f / 255.0f), 0.045f), 1.0f);
    const float metal = unorm_bits((to >> 16) & 0xFFu, 1.0f / 255.0f), omm = 1.0f - metal;
    {
      const uint2 ta = st.px[2][threadIdx.x];
      const float base_r = unorm_bits(ta.x & 0xFFFFu, 1.0f / 65535.0f), base_g = unorm_bits(ta.x >> 16, 1.0f / 65535.0f);
      const float base_b = unorm_bits(ta.y & 0xFFFFu, 1.0f / 65535.0f);
      s.F0r = 0.04f * omm + base_r * metal; s.F0g = 0.04f * omm + base_g * metal; s.F0b = 0.04f * omm + base_b * metal;
      const float cdr = base_r * omm, cdg = base_g * omm, cdb = base_b * omm;
      s.cdpr = cdr * kInvPi; s.cdpg = cdg * kInvPi; s.cdpb = cdb * kInvPi;
      if (AMB != 0) s_amb[1][threadIdx.x] = make_float4(0.f, cdr, cdg, cdb);
    }
    const float alpha = rough * rough;
    s.a2 = alpha * alpha; s.oma2 = 1.0f - s.a2;
    const float vv = s.NoV * s.NoV * s.oma2 + s.a2;
    s.visV = vv * rsq(vv);
    s.omF0r = 1.0f - s.F0r; s.omF0g = 1.0f - s.F0g; s.omF0b = 1.0f - s.F0b;
    s.a2h = s.a2 * (0.5f * kInvPi);

    // ---- ambient set-up: every fetch is issued here and consumed after the lights, which cover its latency ----
    float4 p0r, p0g, p0b, p1r, p1g, p1b, ir, ig, ib, la, lb;
    Quad6 rq0, rq1, rqi;
    Quad8 rql;
    float4 tirr = make_float4(0.f, 0.f, 0.f, 0.f);
    float2 tlut = make_float2(0.f, 0.f);
    if (REC) {  // footprint records: one 256-bit load per bilinear tap, fp32 weights (ibl_quads.cuh)
      const float3 Rv = make_float3(2.0f * NoVraw * s.N.x - s.V.x, 2.0f * NoVraw * s.N.y - s.V.y, 2.0f * NoVraw * s.N.z - s.V.z);
      const IblTaps taps = make_ibl_taps(p, s.N, Rv, s.NoV, rough);  // (the taps a mode does not use are dead code)
      rq0 = ldg_rec256(p.q_pref + 2 * size_t(taps.i_p0));
      rq1 = ldg_rec256(p.q_pref + 2 * size_t(taps.i_p1));
      s_amb[0][threadIdx.x] = make_float4(taps.p0_tx, taps.p0_ty, taps.p1_tx, taps.p1_ty);
      reinterpret_cast<float*>(&s_amb[1][threadIdx.x])[0] = taps.lod_f;
      if (AMBTEX) {
        tirr = texCubemapLod<float4>(p.t_irr, s.N.x * 1.00000095f, s.N.y * 1.00000048f, s.N.z, 0.0f);  // (tilt: see below)
        tlut = tex2D<float2>(p.t_lut, s.NoV, rough);
      } else {
        rqi = ldg_rec256(p.q_irr + 2 * size_t(taps.i_irr));
        rql = ldg_quad(p.q_lut + 2 * size_t(taps.i_lut));
        s_amb[2][threadIdx.x] = make_float4(taps.irr_tx, taps.irr_ty, taps.lut_tx, taps.lut_ty);
      }
    } else if (AMB != 0) {
      const float3 Rv = make_float3(2.0f * NoVraw * s.N.x - s.V.x, 2.0f * NoVraw * s.N.y - s.V.y, 2.0f * NoVraw * s.N.z - s.V.z);
      float facef, u, v;
      cube_dir_to_facef_uv(Rv, facef, u, v);
      const float lod = fminf(rough * q.nm1, q.nm1);  // rough >= 0.045 > 0
      const float lb23 = __fadd_rd(lod, kMagic23), l0f = lb23 - kMagic23;
      const int l0 = __float_as_int(lb23) - 0x4B400000;
      const float4 A = s_lv[l0], B = s_lv[l0 + 1];
      const TapF t0 = tap_floor(u, v, A.x), t1 = tap_floor(u, v, B.x);
      const float c0x = t0.x0 + A.y, c0y = fmaf(facef, A.w, A.z) + t0.y0;
      const float c1x = t1.x0 + B.y, c1y = fmaf(facef, B.w, B.z) + t1.y0;
      p0r = tex2Dgather<float4>(q.g_pref, c0x, c0y, 0); p0g = tex2Dgather<float4>(q.g_pref, c0x, c0y, 1); p0b = tex2Dgather<float4>(q.g_pref, c0x, c0y, 2);
      p1r = tex2Dgather<float4>(q.g_pref, c1x, c1y, 0); p1g = tex2Dgather<float4>(q.g_pref, c1x, c1y, 1); p1b = tex2Dgather<float4>(q.g_pref, c1x, c1y, 2);
      s_amb[0][threadIdx.x] = make_float4(t0.tx, t0.ty, t1.tx, t1.ty);
      reinterpret_cast<float*>(&s_amb[1][threadIdx.x])[0] = lod - l0f;
      if (AMBTEX) {
        // the unit's own face selection breaks exact ties of the major axis (fp16 normals do produce them) its own way, and
        // without seamless filtering the two faces give different edge texels: tilt the direction by 2^-20 so that ties
        // go X, then Y, then Z like cube_dir_to_face_uv (moves the tap by 1e-6 of the face)
        tirr = texCubemapLod<float4>(p.t_irr, s.N.x * 1.00000095f, s.N.y * 1.00000048f, s.N.z, 0.0f);
        tlut = tex2D<float2>(p.t_lut, s.NoV, rough);
      } else {
        cube_dir_to_facef_uv(s.N, facef, u, v);
        const TapF ti = tap_floor(u, v, q.irr_lv.x), tl = tap_floor(s.NoV, rough, q.lut_res);
        const float cix = ti.x0 + q.irr_lv.y, ciy = fmaf(facef, q.irr_lv.w, q.irr_lv.z) + ti.y0;
        ir = tex2Dgather<float4>(q.g_irr, cix, ciy, 0); ig = tex2Dgather<float4>(q.g_irr, cix, ciy, 1); ib = tex2Dgather<float4>(q.g_irr, cix, ciy, 2);
        la = tex2Dgather<float4>(q.g_lut, tl.x0 + 2.0f, tl.y0 + 2.0f, 0); lb = tex2Dgather<float4>(q.g_lut, tl.x0 + 2.0f, tl.y0 + 2.0f, 1);
        s_amb[2][threadIdx.x] = make_float4(ti.tx, ti.ty, tl.tx, tl.ty);
      }
    }

    // ---- direct lighting ----
    float dr = 0.f, dg = 0.f, db = 0.f;
    if (NL >= 0) {
#pragma unroll
      for (int l = 0; l < NL; ++l) cl_light<BRANCH>(s, q.lights[l], dr, dg, db);
    } else {  // run-time count, padded to a multiple of four: four independent chains per iteration
      for (uint32_t l = 0; l < q.n_lights; l += 4) {
#pragma unroll
        for (int k = 0; k < 4; ++k) cl_light<BRANCH>(s, q.lights[l + k], dr, dg, db);
      }
    }

    // ---- ambient: split-sum image-based lighting (Karis 2013) ----
    float ar = 0.f, ag = 0.f, ab = 0.f;
    if (AMB != 0 && (!LATE || q.pad == 0)) {  // (LATE: an opaque, always-true test keeps ptxas from moving the filter -- and the wait for its fetches -- in front of the lights)
      const float4 w0 = s_amb[0][threadIdx.x], w1 = s_amb[1][threadIdx.x];
      RGB pre, pr1;
      if (REC) {
        pre = quad_bilinear_rgb(rq0, w0.x, w0.y);
        pr1 = quad_bilinear_rgb(rq1, w0.z, w0.w);
      } else {
        pre = RGB{bilerp4(p0r, w0.x, w0.y), bilerp4(p0g, w0.x, w0.y), bilerp4(p0b, w0.x, w0.y)};
        pr1 = RGB{bilerp4(p1r, w0.z, w0.w), bilerp4(p1g, w0.z, w0.w), bilerp4(p1b, w0.z, w0.w)};
      }
      pre.r = lerpf(pre.r, pr1.r, w1.x); pre.g = lerpf(pre.g, pr1.g, w1.x); pre.b = lerpf(pre.b, pr1.b, w1.x);
      RGB irr;
      float2 sb;
      if (AMBTEX && REC) {
        irr = RGB{tirr.x, tirr.y, tirr.z};
        sb = tlut;
        if (tirr.w > 0.0f) {  // a flagged texel took part in the fetch: redo this tap with fp32 weights from its record
          int face;
          float u, v;
          cube_dir_to_face_uv(s.N, face, u, v);
          const int ires = int(p.irr.base_res);
          const Tap a = make_tap(u, v, ires);
          irr = quad_bilinear_rgb(ldg_rec256(p.q_irr + 2 * size_t(face * ires * ires + a.idx)), a.tx, a.ty);
        }
      } else if (REC) {
        const float4 w2 = s_amb[2][threadIdx.x];
        irr = quad_bilinear_rgb(rqi, w2.x, w2.y);
        sb = quad_bilinear_lut(rql, w2.z, w2.w);
      } else if (AMBTEX) {
        irr = RGB{tirr.x, tirr.y, tirr.z};
        sb = tlut;
        if (tirr.w > 0.0f) {  // a flagged texel took part in the fetch: redo this tap with fp32 weights
          float facef, u, v;
          cube_dir_to_facef_uv(s.N, facef, u, v);
          const TapF ti = tap_floor(u, v, q.irr_lv.x);
          const float cix = ti.x0 + q.irr_lv.y, ciy = fmaf(facef, q.irr_lv.w, q.irr_lv.z) + ti.y0;
          irr.r = bilerp4(tex2Dgather<float4>(q.g_irr, cix, ciy, 0), ti.tx, ti.ty);
          irr.g = bilerp4(tex2Dgather<float4>(q.g_irr, cix, ciy, 1), ti.tx, ti.ty);
          irr.b = bilerp4(tex2Dgather<float4>(q.g_irr, cix, ciy, 2), ti.tx, ti.ty);
        }
      } else {
        const float4 w2 = s_amb[2][threadIdx.x];
        irr = RGB{bilerp4(ir, w2.x, w2.y), bilerp4(ig, w2.x, w2.y), bilerp4(ib, w2.x, w2.y)};
        sb = make_float2(bilerp4(la, w2.z, w2.w), bilerp4(lb, w2.z, w2.w));
      }
      const float occ = unorm_bits(st.orm[threadIdx.x] & 0xFFu, 1.0f / 255.0f);
      ar = (irr.r * w1.y + pre.r * (s.F0r * sb.x + sb.y)) * occ;
      ag = (irr.g * w1.z + pre.g * (s.F0g * sb.x + sb.y)) * occ;
      ab = (irr.b * w1.w + pre.b * (s.F0b * sb.x + sb.y)) * occ;
    }
    const uint2 te = st.px[3][threadIdx.x];
    const float2 exy = h2_to_f2(te.x), ezw = h2_to_f2(te.y);
    const float base_a = unorm_bits(st.px[2][threadIdx.x].y >> 16, 1.0f / 65535.0f);
    // oracle order: (direct + ambient) + emissive
    const float Lr = fminf((dr + ar) + exy.x, 65504.0f), Lg = fminf((dg + ag) + exy.y, 65504.0f), Lb = fminf((db + ab) + ezw.x, 65504.0f);
    store_texel<PEERS>(p, outo, make_uint2(f2_to_h2(Lr, Lg), f2_to_h2(Lb, base_a)));
  }
  cp_async_wait<0>();
}

// ---- dispatch ----
static void fill_level(float4& o, const AtlasLevel& a) { o = make_float4(float(a.res), float(a.ax + 2), float(a.ay + 2), float(a.res + 2)); }

template <int NL, int AMB, bool BRANCH, int CTAS>
static void launch_cl_modes(const ShadeClParams& q, int lists, bool peers, uint32_t grid, cudaStream_t st) {
  (void)lists;  // (single frames only: shade_common keeps batched calls on shade_ggx_one_kernel)
  if (peers) shade_ggx_cl_kernel<NL, AMB, false, true, BRANCH, CTAS><<<grid, kShadeThreads, 0, st>>>(q);
  else shade_ggx_cl_kernel<NL, AMB, false, false, BRANCH, CTAS><<<grid, kShadeThreads, 0, st>>>(q);
}

template <int NL, bool BRANCH>
static void launch_cl_amb2(const ShadeClParams& q, int amb, int lists, bool peers, uint32_t grid, cudaStream_t st) {
  if (amb == 0) launch_cl_modes<NL, 0, BRANCH, kClCtasPerSm>(q, lists, peers, grid, st);
  else if (amb == 1) launch_cl_modes<NL, 1, BRANCH, kClCtasPerSm>(q, lists, peers, grid, st);
  else if (amb == 2) launch_cl_modes<NL, 2, BRANCH, kClCtasPerSm>(q, lists, peers, grid, st);
  else if (amb == 3) launch_cl_modes<NL, 3, BRANCH, kClCtasPerSm>(q, lists, peers, grid, st);
  else launch_cl_modes<NL, 4, BRANCH, kClCtasPerSm>(q, lists, peers, grid, st);
}
// BRANCH: skip the back half of a light that reaches no lane of the warp.  Off for single frames (straight-line code for the four
// lights: 158.6 vs 169.0 us on the 4K frame); see launch_shade_cl for the batched launches.
template <int NL>
static void launch_cl_amb(const ShadeClParams& q, int amb, int lists, bool peers, uint32_t grid, cudaStream_t st, bool branch) {
  (void)branch;
  launch_cl_amb2<NL, false>(q, amb, lists, peers, grid, st);
}

// p: the parameters shade_common filled (rect, planes, textures, views, replicas); lights: HOST array.
// amb: 0 no IBL; 1 / 2 gathers (default / exact filtering); 3 / 4 footprint records (default / exact).
int launch_shade_cl(const ShadeParams& p, const pbrcu_light* lights_host, uint32_t n_lights, const pbrcu_ibl_accel* accel, int amb, int lists,
                    int sm_count, cudaStream_t st) {
  ShadeClParams q{};
  q.p = p;
  if (amb != 0) {
    q.g_pref = accel->g_pref.tex; q.g_irr = accel->g_irr.tex; q.g_lut = accel->g_lut.tex;
    for (uint32_t m = 0; m < 17; ++m) fill_level(q.lv[m], atlas_level(accel->pref_res, m < accel->pref_mips ? m : accel->pref_mips - 1));
    fill_level(q.irr_lv, atlas_level(accel->irr_res, 0));
    q.lut_res = float(accel->lut_res);
    q.nm1 = float(accel->pref_mips - 1);
  }
  const uint32_t n_eval = n_lights == 0 ? 0 : n_lights <= 2 ? 2 : (n_lights + 3) & ~3u;
  q.n_lights = n_eval;
  for (uint32_t l = 0; l < n_eval; ++l) {
    ClLight& c = q.lights[l];
    if (l < n_lights) {
      const pbrcu_light& s = lights_host[l];
      c.x = s.px; c.y = s.py; c.z = s.pz;
      c.invr2 = s.range > 0.0f ? 1.0f / (s.range * s.range) : 0.0f;
      c.r = s.r * s.intensity; c.g = s.g * s.intensity; c.b = s.b * s.intensity;
    } else {
      c = ClLight{0.f, 0.f, 0.f, 1e30f, 0.f, 0.f, 0.f, 0.f};  // null light: range 1e-15, its window is exactly 0 everywhere, so w = 0
    }
  }
  const bool peers = p.n_replicas != 0 || p.out_is_multicast || p.out_texel == 4;
  const uint64_t items = uint64_t(2) * p.ntiles * (lists == LIST_ONE_BATCHED ? p.n_views : 1);
  const uint32_t grid = uint32_t(std::min<uint64_t>(items, uint64_t(sm_count) * kClCtasPerSm));
  if (const char* e = getenv("PBRCU_K3_EXP")) {  // scratch/k3_cl_exp.py: A/B variants of the headline instantiation
    const int v = atoi(e);
    if (n_eval == 4 && lists == LIST_ONE && !peers && amb != 0 && v > 0) {
      const uint32_t g4 = uint32_t(std::min<uint64_t>(items, uint64_t(sm_count) * 4)), g6 = uint32_t(std::min<uint64_t>(items, uint64_t(sm_count) * 6));
#define PBRCU_EXP1(A, BR, CT, LT, G) shade_ggx_cl_kernel<4, A, false, false, BR, CT, LT><<<G, kShadeThreads, 0, st>>>(q)
#define PBRCU_EXP(BR, CT, LT, G) \
  (amb == 1 ? PBRCU_EXP1(1, BR, CT, LT, G) : amb == 2 ? PBRCU_EXP1(2, BR, CT, LT, G) : amb == 3 ? PBRCU_EXP1(3, BR, CT, LT, G) : PBRCU_EXP1(4, BR, CT, LT, G))
      if (v == 1) PBRCU_EXP(true, 5, false, grid);
      else if (v == 2) PBRCU_EXP(true, 4, false, g4);
      else if (v == 3) PBRCU_EXP(true, 6, false, g6);
      else if (v == 4) PBRCU_EXP(true, 4, true, g4);
      else if (v == 5) PBRCU_EXP(true, 5, true, grid);
      else if (v == 6) PBRCU_EXP(false, 6, false, g6);
      else if (v == 7) PBRCU_EXP(false, 5, true, grid);
      else if (v == 8) PBRCU_EXP(false, 6, true, g6);
      else if (v == 9) PBRCU_EXP(false, 4, false, g4);
      else PBRCU_EXP(true, 5, false, grid);
#undef PBRCU_EXP
#undef PBRCU_EXP1
      return launch_check("shade_ggx_cl_kernel (experiment)");
    }
  }
  const bool branch = false;
  if (n_eval == 2) launch_cl_amb<2>(q, amb, lists, peers, grid, st, branch);
  else if (n_eval == 4) launch_cl_amb<4>(q, amb, lists, peers, grid, st, branch);
  else launch_cl_amb<-1>(q, amb, lists, peers, grid, st, branch);
  return launch_check("shade_ggx_cl_kernel");
}

}  // namespace pbrcu
