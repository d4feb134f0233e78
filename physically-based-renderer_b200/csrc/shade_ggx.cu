// shade_ggx.cu -- K3 metallic-roughness Cook-Torrance shading (GGX D, height-correlated Smith V,
// Schlick F) with point lights, occlusion, emissive and split-sum IBL over the extended G-buffer, and
// K3c the per-tile light lists.  No reference counterpart; definition:
// glTF 2.0 Appendix B + KHR_lights_punctual + Karis 2013 (restated for the CPU checker as pbro_shade_pixel).
//
// The light loop evaluates TWO lights per iteration as packed f32x2 instructions (f32x2.cuh: light 2j in
// the low lane, light 2j+1 in the high lane, the pixel's own quantities as scalar-broadcast operands) --
// the kernel is bound by issue slots, not by the FMA pipe, and a packed instruction does two lights' worth
// per slot.  Persistent CTAs (3 per SM) walk the 16x16 pixel tiles of the rect.  The five G-buffer planes of the
// NEXT tile are staged into shared memory with cp.async while the current tile is shaded, so the DRAM
// latency of the streaming reads never sits on a warp's critical path; the tile's light list is staged
// in shared memory; the IBL footprint records of a pixel are prefetched into L1 before its light loop.
#include <cstdlib>
#include <mutex>

#include "shade_ggx.cuh"

namespace pbrcu {

template <int IBL, int LISTS, bool PEERS = false>
__global__ void __launch_bounds__(kShadeThreads, kShadeCtasPerSm) shade_ggx_kernel(const __grid_constant__ ShadeParams p) {
  constexpr bool ONE_LIST = LISTS != LIST_TILED, BATCHED = LISTS == LIST_ONE_BATCHED;
  // staged lights, two per entry (lane 0 = light 2j, lane 1 = light 2j+1):
  // s_lp[j] = {{x0,x1,y0,y1}, {z0,z1,w0,w1}} with w = 1/range^2 (0 = unlimited); s_lc[j] = {{r0,r1,g0,g1}, {b0,b1,-,-}}
  // with rgb = colour * intensity.  An odd count is padded with a null light (colour 0: it adds exactly 0).
  __shared__ float4 s_lp[kLightChunk];
  __shared__ float4 s_lc[kLightChunk];
  __shared__ TileStage s_tile[2];
  auto stage_lights = [&](const uint16_t* idx, uint32_t base, uint32_t cnt) {
    if (threadIdx.x < ((cnt + 1) & ~1u)) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
      if (threadIdx.x < cnt) {
        const uint32_t li = idx ? uint32_t(__ldg(idx + base + threadIdx.x)) : base + threadIdx.x;
        a = __ldg(reinterpret_cast<const float4*>(p.lights) + 2 * li);
        b = __ldg(reinterpret_cast<const float4*>(p.lights) + 2 * li + 1);
      }
      float* lp = reinterpret_cast<float*>(s_lp + (threadIdx.x & ~1u)) + (threadIdx.x & 1u);
      float* lc = reinterpret_cast<float*>(s_lc + (threadIdx.x & ~1u)) + (threadIdx.x & 1u);
      lp[0] = a.x; lp[2] = a.y; lp[4] = a.z; lp[6] = a.w > 0.0f ? __fdividef(1.0f, a.w * a.w) : 0.0f;
      lc[0] = b.x * b.w; lc[2] = b.y * b.w; lc[4] = b.z * b.w;
    }
  };
  if (ONE_LIST) stage_lights(nullptr, 0, p.n_lights);

  const uint32_t nitems = 2 * p.ntiles;  // per view
  uint32_t tile = blockIdx.x, view = 0;  // work item (half tile) of `view`
  // BATCHED: the CTAs stride over the items of all views; (view, tile) is kept normalised without a division
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
  if (ONE_LIST) __syncthreads();

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
    cp_async_wait<1>();  // this tile's group has landed (each thread reads only its own slots)

    const uint32_t tx = ip.tx, ty = ip.ty, x = ip.x, y = ip.y;
    const bool inside = x >= p.x0 && x < p.x1 && y >= p.y0 && y < p.y1;

    // ---- per-pixel set-up (glTF 2.0 Appendix B material model) ----
    float3 P = make_float3(0.f, 0.f, 0.f), N = P, V = P;
    float NoVraw = 0.f, NoV = 1.f, rough = 1.f, occ = 0.f, a2 = 1.f, oma2 = 0.f, visV = 1.f, base_a = 0.f;
    float F0r = 0.f, F0g = 0.f, F0b = 0.f, cdr = 0.f, cdg = 0.f, cdb = 0.f;
    float Lr = 0.f, Lg = 0.f, Lb = 0.f;
    bool lit = false;
    if (inside) {
      const TileStage& st = s_tile[stage];
      const uint2 tn = st.px[1][threadIdx.x];
      const float2 nxy = h2_to_f2(tn.x), nzw = h2_to_f2(tn.y);
      lit = !(nxy.x == 0.0f && nxy.y == 0.0f && nzw.x == 0.0f);  // all-zero normal = cleared texel
      if (lit) {
        const uint2 tp = st.px[0][threadIdx.x], ta = st.px[2][threadIdx.x], te = st.px[3][threadIdx.x];
        const uint32_t to = st.orm[threadIdx.x];
        const float2 pxy = h2_to_f2(tp.x), pzw = h2_to_f2(tp.y);
        P = make_float3(pxy.x, pxy.y, pzw.x);
        const float ninv = rsq(nxy.x * nxy.x + nxy.y * nxy.y + nzw.x * nzw.x);
        N = make_float3(nxy.x * ninv, nxy.y * ninv, nzw.x * ninv);
        float cx = p.cx, cy = p.cy, cz = p.cz;
        if (BATCHED) { cx = __ldg(p.cams + 3 * ip.view); cy = __ldg(p.cams + 3 * ip.view + 1); cz = __ldg(p.cams + 3 * ip.view + 2); }
        const float vx = cx - P.x, vy = cy - P.y, vz = cz - P.z;
        const float vinv = rsq(vx * vx + vy * vy + vz * vz);
        V = make_float3(vx * vinv, vy * vinv, vz * vinv);
        NoVraw = N.x * V.x + N.y * V.y + N.z * V.z;
        NoV = fminf(fmaxf(NoVraw, 1e-4f), 1.0f);
        const float base_r = unorm16_lo(ta.x), base_g = unorm16_hi(ta.x), base_b = unorm16_lo(ta.y);
        base_a = unorm16_hi(ta.y);
        occ = float(to & 0xFFu) * (1.0f / 255.0f);
        rough = fminf(fmaxf(float((to >> 8) & 0xFFu) * (1.0f / 255.0f), 0.045f), 1.0f);
        const float metal = float((to >> 16) & 0xFFu) * (1.0f / 255.0f), omm = 1.0f - metal;
        F0r = 0.04f * omm + base_r * metal; F0g = 0.04f * omm + base_g * metal; F0b = 0.04f * omm + base_b * metal;
        cdr = base_r * omm; cdg = base_g * omm; cdb = base_b * omm;
        const float alpha = rough * rough;
        a2 = alpha * alpha; oma2 = 1.0f - a2;
        const float vv = NoV * NoV * oma2 + a2;
        visV = vv * rsq(vv);
        const float2 exy = h2_to_f2(te.x), ezw = h2_to_f2(te.y);
        Lr = exy.x; Lg = exy.y; Lb = ezw.x;  // emissive
      }
    }
    const float cdpr = cdr * kInvPi, cdpg = cdg * kInvPi, cdpb = cdb * kInvPi;
    const float omF0r = 1.0f - F0r, omF0g = 1.0f - F0g, omF0b = 1.0f - F0b;
    const float a2h = a2 * (0.5f * kInvPi);

    // ---- ambient set-up: the four footprint records are LOADED here (32 registers) and consumed after
    // the light loop, whose ~400 instructions cover their L2 / DRAM latency (a prefetch into L1 does not
    // survive: 24 warps x 32 lanes x 4 incoherent 128-byte lines exceed the L1) ----
    IblTaps taps;
    Quad6 qi, q0, q1;
    Quad8 ql;
    float4 tpre = make_float4(0.f, 0.f, 0.f, 0.f), tirr = tpre;
    Tap tlut;
    float3 Rv = make_float3(0.f, 0.f, 0.f);
    if (lit && p.has_ibl) {
      Rv = make_float3(2.0f * NoVraw * N.x - V.x, 2.0f * NoVraw * N.y - V.y, 2.0f * NoVraw * N.z - V.z);
      if (IBL == IBL_RECORDS || IBL == IBL_RECORDS_AMBTEX) {
        taps = make_ibl_taps(p, N, Rv, NoV, rough);
        q0 = ldg_rec(p.q_pref + 2 * size_t(taps.i_p0));
        q1 = ldg_rec(p.q_pref + 2 * size_t(taps.i_p1));
        if (IBL == IBL_RECORDS) {
          qi = ldg_rec(p.q_irr + 2 * size_t(taps.i_irr));
          ql = ldg_quad(p.q_lut + 2 * size_t(taps.i_lut));
        } else {
          tirr = texCubemapLod<float4>(p.t_irr, N.x * 1.00000095f, N.y * 1.00000048f, N.z, 0.0f);  // tilt: see below
          const float2 l = tex2D<float2>(p.t_lut, NoV, rough);
          tpre.x = l.x; tpre.y = l.y;
        }
      }
      if (IBL == IBL_TEX) {
        // face selection, clamp-to-edge bilinear on two levels and the level blend in the texture unit
        const float lod = fminf(rough * float(p.pref.nmips - 1), float(p.pref.nmips - 1));
        // (directions tilted by 2^-20 so that exact ties of the major axis go X, Y, Z like cube_dir_to_face_uv)
        tpre = texCubemapLod<float4>(p.t_pref, Rv.x * 1.00000095f, Rv.y * 1.00000048f, Rv.z, lod);
        tirr = texCubemapLod<float4>(p.t_irr, N.x * 1.00000095f, N.y * 1.00000048f, N.z, 0.0f);
        // a level fraction the unit would round to 1 hides the finer level's flags: treat it as flagged
        if (lod - floorf(lod) > 0.995f) tirr.w = 1.0f;
        tlut = make_tap(NoV, rough, int(p.lut_res));
        ql = ldg_quad(p.q_lut + 2 * size_t(tlut.idx));
      }
    }

    // ---- direct lighting: KHR_lights_punctual point lights, Cook-Torrance GGX / Smith / Schlick ----
    uint32_t nl = p.n_lights;
    const uint16_t* idx = nullptr;
    if (!ONE_LIST && p.tile_counts != nullptr) {
      const uint32_t t = ty * p.tiles_x + tx;
      nl = min(__ldg(p.tile_counts + t), p.max_per_tile);
      idx = p.tile_indices + size_t(t) * p.max_per_tile;
    }
    f2 dr2 = dup(0.f), dg2 = dr2, db2 = dr2;  // per lane: the even / the odd lights of the list, in list order
    const LightPixel lp{P, N, V, NoV, visV, a2, oma2, a2h, F0r, F0g, F0b, omF0r, omF0g, omF0b, cdpr, cdpg, cdpb};
    for (uint32_t base = 0; base < nl; base += kLightChunk) {
      const uint32_t cnt = ONE_LIST ? nl : min(nl - base, uint32_t(kLightChunk));
      if (!ONE_LIST) {
        __syncthreads();  // everyone is done with the previous chunk / tile
        stage_lights(idx, base, cnt);
        __syncthreads();
      }
      if (!lit) continue;
      for (uint32_t j = 0; 2 * j < cnt; ++j) {
        const PairA a = light_pair_a(lp, s_lp[2 * j], s_lp[2 * j + 1]);
        if (lo(a.w) > 0.0f || hi(a.w) > 0.0f) {
          f2 br, bg, bb;
          light_pair_b(lp, a, br, bg, bb);
          const float4 c0 = s_lc[2 * j], c1 = s_lc[2 * j + 1];
          dr2 = fma2(br, mul2(pk(c0.x, c0.y), a.w), dr2);
          dg2 = fma2(bg, mul2(pk(c0.z, c0.w), a.w), dg2);
          db2 = fma2(bb, mul2(pk(c1.x, c1.y), a.w), db2);
        }
      }
    }
    const float dr = lo(dr2) + hi(dr2), dg = lo(dg2) + hi(dg2), db = lo(db2) + hi(db2);
    // every lane reaches the store at the end of the iteration together (`have`: it owns a texel of the rect; a cleared texel is
    // stored as 0), so the peer-store instantiations can pair the texels of neighbouring lanes into 16-byte stores
    const size_t outo = size_t(y) * p.out_pitch + size_t(x) * (PEERS ? p.out_texel : 8u) + (BATCHED ? size_t(ip.view) * p.out_view_stride : 0);
    const bool have = inside;
    uint2 texel = make_uint2(0u, 0u);
    if (have && lit) {
    // ---- ambient: split-sum image-based lighting (Karis 2013) ----
    float ar = 0.f, ag = 0.f, ab = 0.f;
    if (p.has_ibl) {
      RGB irr, pre;
      float2 sb;
      if (IBL == IBL_RECORDS || IBL == IBL_RECORDS_AMBTEX) {
        pre = quad_bilinear_rgb(q0, taps.p0_tx, taps.p0_ty);
        const RGB pr1 = quad_bilinear_rgb(q1, taps.p1_tx, taps.p1_ty);
        pre.r = lerpf(pre.r, pr1.r, taps.lod_f); pre.g = lerpf(pre.g, pr1.g, taps.lod_f); pre.b = lerpf(pre.b, pr1.b, taps.lod_f);
        if (IBL == IBL_RECORDS) {
          irr = quad_bilinear_rgb(qi, taps.irr_tx, taps.irr_ty);
          sb = quad_bilinear_lut(ql, taps.lut_tx, taps.lut_ty);
        } else {
          irr = RGB{tirr.x, tirr.y, tirr.z};
          sb = make_float2(tpre.x, tpre.y);
          if (tirr.w > 0.0f) {  // flagged footprint: fp32 weights from the record
            qi = ldg_rec(p.q_irr + 2 * size_t(taps.i_irr));
            irr = quad_bilinear_rgb(qi, taps.irr_tx, taps.irr_ty);
          }
        }
      } else if (IBL == IBL_TEX) {
        irr = RGB{tirr.x, tirr.y, tirr.z};
        pre = RGB{tpre.x, tpre.y, tpre.z};
        if (tpre.w > 0.0f || tirr.w > 0.0f) {
          // a flagged texel took part in the fetch: the unit's 8-bit weights are not accurate enough for
          // this footprint's contrast (around the sun), so redo both taps with fp32 weights from the records
          taps = make_ibl_taps(p, N, Rv, NoV, rough);
          q0 = ldg_rec(p.q_pref + 2 * size_t(taps.i_p0));
          q1 = ldg_rec(p.q_pref + 2 * size_t(taps.i_p1));
          qi = ldg_rec(p.q_irr + 2 * size_t(taps.i_irr));
          irr = quad_bilinear_rgb(qi, taps.irr_tx, taps.irr_ty);
          pre = quad_bilinear_rgb(q0, taps.p0_tx, taps.p0_ty);
          const RGB pr1 = quad_bilinear_rgb(q1, taps.p1_tx, taps.p1_ty);
          pre.r = lerpf(pre.r, pr1.r, taps.lod_f); pre.g = lerpf(pre.g, pr1.g, taps.lod_f); pre.b = lerpf(pre.b, pr1.b, taps.lod_f);
        }
        sb = quad_bilinear_lut(ql, tlut.tx, tlut.ty);
      } else {
        irr = cube_sample(p.irr, N, 0.0f);
        pre = cube_sample(p.pref, Rv, rough * float(p.pref.nmips - 1));
        sb = lut_sample(p.lut, int(p.lut_res), NoV, rough);
      }
      ar = (irr.r * cdr + pre.r * (F0r * sb.x + sb.y)) * occ;
      ag = (irr.g * cdg + pre.g * (F0g * sb.x + sb.y)) * occ;
      ab = (irr.b * cdb + pre.b * (F0b * sb.x + sb.y)) * occ;
    }
    // oracle order: (direct + ambient) + emissive
    Lr = fminf((dr + ar) + Lr, 65504.0f); Lg = fminf((dg + ag) + Lg, 65504.0f); Lb = fminf((db + ab) + Lb, 65504.0f);
    texel = make_uint2(f2_to_h2(Lr, Lg), f2_to_h2(Lb, base_a));
    }
    if (PEERS) store_texel_paired(p, outo, texel, have);
    else if (have) stg_stream8(p.out + outo, texel);
  }
  cp_async_wait<0>();
}

// ---- K3, single-list form (LIST_ONE / LIST_ONE_BATCHED with footprint records or no IBL) ----
// Same arithmetic as shade_ggx_kernel, laid out for what ncu showed to bound it (profiles/r01_k3_*):
//  * dependent-issue latency: the packed light chain is ONE dependency chain per thread, and 5 warps per scheduler do
//    not cover it.  Here two light pairs (four lights) are evaluated side by side, so the scheduler always has a
//    second independent chain of the same warp to issue from;
//  * every record is fetched with ONE 256-bit load (one request per lane instead of two).
// Registers for both come from deferring everything only the ambient term needs (tap weights, c_diff, occlusion,
// emissive, alpha) to shared memory: a small per-thread stash, and the staged tile itself, which stays valid for the
// whole iteration.  Unlit and outside pixels leave the iteration early instead of carrying neutral values through it.
// AMBTEX: the two SMOOTH maps of the ambient term -- irradiance (a cosine convolution) and the split-sum LUT -- come from the
// otherwise idle texture unit: one TEX each (face selection and bilinear in hardware) instead of face / tap arithmetic, a
// scattered 256-bit load and a 24-instruction filter.  The unit's 8-bit weights cost at most 2/512 of the footprint's value
// range; an irradiance footprint whose range exceeds 12 % of its minimum (the flag in the texture's alpha; error below
// 4.7e-4 relative otherwise) is re-filtered from its record.  The prefiltered cube keeps fp32 weights from the records in every mode.
template <bool BATCHED, bool AMBTEX, bool PEERS = false>
__global__ void __launch_bounds__(kShadeThreads, kShadeCtasPerSm) shade_ggx_one_kernel(const __grid_constant__ ShadeParams p) {
  __shared__ float4 s_lp[kLightChunk];
  __shared__ float4 s_lc[kLightChunk];
  __shared__ TileStage s_tile[2];
  __shared__ float4 s_amb[3][kShadeThreads];  // {irr_tx, irr_ty, p0_tx, p0_ty}, {p1_tx, p1_ty, lut_tx, lut_ty}, {lod_f, cdr, cdg, cdb}
  {
    // lights, two per entry as in shade_ggx_kernel, padded with null lights (colour 0: they add exactly 0) to a
    // multiple of FOUR so the loop below always takes two pairs
    const uint32_t cnt = p.n_lights;
    if (threadIdx.x < ((cnt + 3) & ~3u)) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
      if (threadIdx.x < cnt) {
        a = __ldg(reinterpret_cast<const float4*>(p.lights) + 2 * threadIdx.x);
        b = __ldg(reinterpret_cast<const float4*>(p.lights) + 2 * threadIdx.x + 1);
      }
      float* lp = reinterpret_cast<float*>(s_lp + (threadIdx.x & ~1u)) + (threadIdx.x & 1u);
      float* lc = reinterpret_cast<float*>(s_lc + (threadIdx.x & ~1u)) + (threadIdx.x & 1u);
      lp[0] = a.x; lp[2] = a.y; lp[4] = a.z; lp[6] = a.w > 0.0f ? __fdividef(1.0f, a.w * a.w) : 0.0f;
      lc[0] = b.x * b.w; lc[2] = b.y * b.w; lc[4] = b.z * b.w;
    }
  }
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
  const uint32_t quads = (p.n_lights + 3) >> 2;

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
    const float rough = fminf(fmaxf(float((to >> 8) & 0xFFu) * (1.0f / 255.0f), 0.045f), 1.0f);
    const float metal = float((to >> 16) & 0xFFu) * (1.0f / 255.0f), omm = 1.0f - metal;
    {
      const uint2 ta = st.px[2][threadIdx.x];
      const float base_r = unorm16_lo(ta.x), base_g = unorm16_hi(ta.x), base_b = unorm16_lo(ta.y);
      s.F0r = 0.04f * omm + base_r * metal; s.F0g = 0.04f * omm + base_g * metal; s.F0b = 0.04f * omm + base_b * metal;
      const float cdr = base_r * omm, cdg = base_g * omm, cdb = base_b * omm;
      s.cdpr = cdr * kInvPi; s.cdpg = cdg * kInvPi; s.cdpb = cdb * kInvPi;
      s_amb[2][threadIdx.x] = make_float4(0.f, cdr, cdg, cdb);
    }
    const float alpha = rough * rough;
    s.a2 = alpha * alpha; s.oma2 = 1.0f - s.a2;
    const float vv = s.NoV * s.NoV * s.oma2 + s.a2;
    s.visV = vv * rsq(vv);
    s.omF0r = 1.0f - s.F0r; s.omF0g = 1.0f - s.F0g; s.omF0b = 1.0f - s.F0b;
    s.a2h = s.a2 * (0.5f * kInvPi);

    // ---- ambient set-up: the four footprint records are loaded here and consumed after the light loop ----
    Quad6 qi, q0, q1;
    Quad8 ql;
    float4 tirr = make_float4(0.f, 0.f, 0.f, 0.f);
    float2 tlut = make_float2(0.f, 0.f);
    if (p.has_ibl) {
      const float3 Rv = make_float3(2.0f * NoVraw * s.N.x - s.V.x, 2.0f * NoVraw * s.N.y - s.V.y, 2.0f * NoVraw * s.N.z - s.V.z);
      const IblTaps taps = make_ibl_taps(p, s.N, Rv, s.NoV, rough);  // (the taps a mode does not use are dead code)
      q0 = ldg_rec256(p.q_pref + 2 * size_t(taps.i_p0));
      q1 = ldg_rec256(p.q_pref + 2 * size_t(taps.i_p1));
      if (AMBTEX) {
        // the unit's own face selection breaks exact ties of the major axis (fp16 normals do produce them) its own way, and
        // without seamless filtering the two faces give different edge texels: tilt the direction by 2^-20 so that ties
        // go X, then Y, then Z like cube_dir_to_face_uv (moves the tap by 1e-6 of the face)
        tirr = texCubemapLod<float4>(p.t_irr, s.N.x * 1.00000095f, s.N.y * 1.00000048f, s.N.z, 0.0f);
        tlut = tex2D<float2>(p.t_lut, s.NoV, rough);
        s_amb[0][threadIdx.x] = make_float4(taps.p0_tx, taps.p0_ty, taps.p1_tx, taps.p1_ty);
      } else {
        qi = ldg_rec256(p.q_irr + 2 * size_t(taps.i_irr));
        ql = ldg_quad(p.q_lut + 2 * size_t(taps.i_lut));
        s_amb[0][threadIdx.x] = make_float4(taps.irr_tx, taps.irr_ty, taps.p0_tx, taps.p0_ty);
        s_amb[1][threadIdx.x] = make_float4(taps.p1_tx, taps.p1_ty, taps.lut_tx, taps.lut_ty);
      }
      reinterpret_cast<float*>(&s_amb[2][threadIdx.x])[0] = taps.lod_f;
    }

    // ---- direct lighting: four lights per iteration as two independent packed chains ----
    f2 dr2 = dup(0.f), dg2 = dr2, db2 = dr2;
    for (uint32_t it = 0; it < quads; ++it) {
      const float4* lp = s_lp + 4 * it;
      const PairA a = light_pair_a(s, lp[0], lp[1]), b = light_pair_a(s, lp[2], lp[3]);
      if (fmaxf(fmaxf(lo(a.w), hi(a.w)), fmaxf(lo(b.w), hi(b.w))) > 0.0f) {
        f2 ar, ag, ab, br, bg, bb;
        light_pair_b(s, a, ar, ag, ab);
        light_pair_b(s, b, br, bg, bb);
        const float4* lc = s_lc + 4 * it;
        const float4 c0 = lc[0], c1 = lc[1], c2 = lc[2], c3 = lc[3];
        // the pairs are added in list order: the same sums as shade_ggx_kernel's
        dr2 = fma2(ar, mul2(pk(c0.x, c0.y), a.w), dr2); dg2 = fma2(ag, mul2(pk(c0.z, c0.w), a.w), dg2); db2 = fma2(ab, mul2(pk(c1.x, c1.y), a.w), db2);
        dr2 = fma2(br, mul2(pk(c2.x, c2.y), b.w), dr2); dg2 = fma2(bg, mul2(pk(c2.z, c2.w), b.w), dg2); db2 = fma2(bb, mul2(pk(c3.x, c3.y), b.w), db2);
      }
    }
    const float dr = lo(dr2) + hi(dr2), dg = lo(dg2) + hi(dg2), db = lo(db2) + hi(db2);

    // ---- ambient: split-sum image-based lighting (Karis 2013) ----
    float ar = 0.f, ag = 0.f, ab = 0.f;
    if (p.has_ibl) {
      const float4 w0 = s_amb[0][threadIdx.x], w2 = s_amb[2][threadIdx.x];
      RGB irr, pre, pr1;
      float2 sb;
      if (AMBTEX) {
        pre = quad_bilinear_rgb(q0, w0.x, w0.y);
        pr1 = quad_bilinear_rgb(q1, w0.z, w0.w);
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
      } else {
        const float4 w1 = s_amb[1][threadIdx.x];
        irr = quad_bilinear_rgb(qi, w0.x, w0.y);
        pre = quad_bilinear_rgb(q0, w0.z, w0.w);
        pr1 = quad_bilinear_rgb(q1, w1.x, w1.y);
        sb = quad_bilinear_lut(ql, w1.z, w1.w);
      }
      pre.r = lerpf(pre.r, pr1.r, w2.x); pre.g = lerpf(pre.g, pr1.g, w2.x); pre.b = lerpf(pre.b, pr1.b, w2.x);
      const float occ = float(st.orm[threadIdx.x] & 0xFFu) * (1.0f / 255.0f);
      ar = (irr.r * w2.y + pre.r * (s.F0r * sb.x + sb.y)) * occ;
      ag = (irr.g * w2.z + pre.g * (s.F0g * sb.x + sb.y)) * occ;
      ab = (irr.b * w2.w + pre.b * (s.F0b * sb.x + sb.y)) * occ;
    }
    const uint2 te = st.px[3][threadIdx.x];
    const float2 exy = h2_to_f2(te.x), ezw = h2_to_f2(te.y);
    const float base_a = unorm16_hi(st.px[2][threadIdx.x].y);
    // oracle order: (direct + ambient) + emissive
    const float Lr = fminf((dr + ar) + exy.x, 65504.0f), Lg = fminf((dg + ag) + exy.y, 65504.0f), Lb = fminf((db + ab) + ezw.x, 65504.0f);
    store_texel<PEERS>(p, outo, make_uint2(f2_to_h2(Lr, Lg), f2_to_h2(Lb, base_a)));
  }
  cp_async_wait<0>();
}

// ---- K3c: per 16x16 tile, world-space AABB of the tile's non-cleared positions, then every light
// whose range sphere meets it, in ascending light index (so K3 sums in the oracle's order).
// One WARP per tile (eight tiles per CTA): the 256 pixels are read as eight fully coalesced 2-row slabs, the box is
// reduced with shuffles and the lights are compacted 32 at a time with a ballot -- no block barriers at all.
__global__ void __launch_bounds__(kClusterThreads) cluster_lights_kernel(
    const uint8_t* pos, const uint8_t* nrm, size_t pitch8, uint32_t width, uint32_t height,
    const pbrcu_light* lights, uint32_t n_lights, uint32_t* counts, uint16_t* indices, uint32_t tiles_x, uint32_t n_tiles,
    uint32_t max_per_tile) {
  const uint32_t tile = (blockIdx.x * kClusterThreads + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (tile >= n_tiles) return;
  const uint32_t ty = tile / tiles_x, tx = tile - ty * tiles_x;
  const uint32_t x = tx * kTile + (lane & 15u);
  float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
#pragma unroll
  for (int k = 0; k < kTile / 2; ++k) {
    const uint32_t y = ty * kTile + 2 * k + (lane >> 4);
    if (x < width && y < height) {
      const size_t o8 = size_t(y) * pitch8 + size_t(x) * 8;
      // both planes are fetched at once (the position does not wait for the normal's test): all sixteen loads of a lane are
      // in flight together -- 165 -> 133 us on the 8K frame (scratch/k3c_time.py)
      const uint2 tn = __ldg(reinterpret_cast<const uint2*>(nrm + o8));
      const uint2 tp = __ldg(reinterpret_cast<const uint2*>(pos + o8));
      const float2 nxy = h2_to_f2(tn.x), nzw = h2_to_f2(tn.y);
      if (!(nxy.x == 0.0f && nxy.y == 0.0f && nzw.x == 0.0f)) {
        const float2 pxy = h2_to_f2(tp.x), pzw = h2_to_f2(tp.y);
        mn[0] = fminf(mn[0], pxy.x); mx[0] = fmaxf(mx[0], pxy.x);
        mn[1] = fminf(mn[1], pxy.y); mx[1] = fmaxf(mx[1], pxy.y);
        mn[2] = fminf(mn[2], pzw.x); mx[2] = fmaxf(mx[2], pzw.x);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], off));
      mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], off));
    }
  }
  const bool empty = mn[0] > mx[0];
  uint32_t total = 0;
  for (uint32_t base = 0; base < n_lights; base += 32) {
    const uint32_t i = base + lane;
    bool hit = false;
    if (i < n_lights && !empty) {
      const float4 la = __ldg(reinterpret_cast<const float4*>(lights) + 2 * i);
      if (la.w <= 0.0f) hit = true;  // unlimited range
      else {
        // fp16 positions are exact in fp32; widen the sphere by 1 ulp-ish so K3's window (which is
        // exactly 0 at d >= range) can never see a light this test dropped
        const float ddx = fmaxf(fmaxf(mn[0] - la.x, la.x - mx[0]), 0.0f);
        const float ddy = fmaxf(fmaxf(mn[1] - la.y, la.y - mx[1]), 0.0f);
        const float ddz = fmaxf(fmaxf(mn[2] - la.z, la.z - mx[2]), 0.0f);
        hit = ddx * ddx + ddy * ddy + ddz * ddz <= la.w * la.w * 1.0001f;
      }
    }
    const uint32_t ballot = __ballot_sync(0xffffffffu, hit);  // ordered compaction
    const uint32_t slot = total + __popc(ballot & ((1u << lane) - 1u));
    if (hit && slot < max_per_tile) indices[size_t(tile) * max_per_tile + slot] = uint16_t(i);
    total += __popc(ballot);
  }
  if (lane == 0) counts[tile] = min(total, max_per_tile);
}

}  // namespace pbrcu

using namespace pbrcu;

extern "C" int pbrcu_light_tiles_alloc(pbrcu_ctx* ctx, uint32_t width, uint32_t height, uint32_t max_per_tile,
                                       pbrcu_light_tiles* out) {
  PBRCU_REQUIRE(ctx && out, "NULL argument");
  PBRCU_REQUIRE(width > 0 && height > 0 && max_per_tile > 0 && max_per_tile <= uint32_t(kMaxLights), "extent or max_per_tile out of range");
  out->tiles_x = (width + kTile - 1) / kTile;
  out->tiles_y = (height + kTile - 1) / kTile;
  out->max_per_tile = max_per_tile;
  out->counts = nullptr; out->indices = nullptr;
  const size_t nt = size_t(out->tiles_x) * out->tiles_y;
  int rc;
  if ((rc = pbrcu_malloc(ctx, nt * 4, reinterpret_cast<void**>(&out->counts)))) return rc;
  if ((rc = pbrcu_malloc(ctx, nt * max_per_tile * 2, reinterpret_cast<void**>(&out->indices)))) { pbrcu_free(ctx, out->counts); out->counts = nullptr; return rc; }
  return PBRCU_OK;
}

extern "C" int pbrcu_light_tiles_free(pbrcu_ctx* ctx, pbrcu_light_tiles* t) {
  if (!t) return PBRCU_OK;
  pbrcu_free(ctx, t->counts);
  pbrcu_free(ctx, t->indices);
  t->counts = nullptr; t->indices = nullptr;
  return PBRCU_OK;
}

extern "C" int pbrcu_cluster_lights(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const pbrcu_light* lights_dev, uint32_t n_lights,
                                    const pbrcu_light_tiles* tiles, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && g && tiles && tiles->counts && tiles->indices, "NULL argument");
  PBRCU_REQUIRE(g->positions && g->normals, "gbuffer plane is NULL");
  PBRCU_REQUIRE(n_lights <= 65535 && (n_lights == 0 || lights_dev), "n_lights out of range or lights NULL");
  PBRCU_REQUIRE(tiles->tiles_x == (g->width + kTile - 1) / kTile && tiles->tiles_y == (g->height + kTile - 1) / kTile, "tile grid does not match the gbuffer");
  const uint32_t n_tiles = tiles->tiles_x * tiles->tiles_y;
  cluster_lights_kernel<<<(n_tiles + kClusterThreads / 32 - 1) / (kClusterThreads / 32), kClusterThreads, 0, static_cast<cudaStream_t>(s)>>>(
      static_cast<const uint8_t*>(g->positions), static_cast<const uint8_t*>(g->normals), g->pitch8, g->width, g->height,
      lights_dev, n_lights, tiles->counts, tiles->indices, tiles->tiles_x, n_tiles, tiles->max_per_tile);
  return launch_check("cluster_lights_kernel");
}

extern "C" int pbrcu_shade_set_texture_unit(pbrcu_ctx* ctx, int on) {
  PBRCU_REQUIRE(ctx, "ctx is NULL");
  ctx->shade_tex = on != 0;
  return PBRCU_OK;
}

template <int IBL>
static void launch_shade(const ShadeParams& p, int lists, uint32_t grid, cudaStream_t st) {
  const bool peers = p.n_replicas != 0 || p.out_is_multicast || p.out_texel == 4;  // single view only (shade_common)
  if (lists == LIST_ONE_BATCHED) shade_ggx_kernel<IBL, LIST_ONE_BATCHED><<<grid, kShadeThreads, 0, st>>>(p);
  else if (lists == LIST_ONE && peers) shade_ggx_kernel<IBL, LIST_ONE, true><<<grid, kShadeThreads, 0, st>>>(p);
  else if (lists == LIST_ONE) shade_ggx_kernel<IBL, LIST_ONE><<<grid, kShadeThreads, 0, st>>>(p);
  else if (peers) shade_ggx_kernel<IBL, LIST_TILED, true><<<grid, kShadeThreads, 0, st>>>(p);
  else shade_ggx_kernel<IBL, LIST_TILED><<<grid, kShadeThreads, 0, st>>>(p);
}

// argument checks and kernel parameters shared by the single-view and the batched entry; n_views > 1 = batched
static int shade_common(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3], const pbrcu_light* lights_dev, uint32_t n_lights,
                        const pbrcu_light_tiles* tiles, const pbrcu_ibl* ibl, pbrcu_rect rect, void* out_rgba16f, size_t out_pitch,
                        uint32_t n_views, size_t view_stride8, size_t view_stride4, size_t out_view_stride, pbrcu_stream s,
                        void* const* replicas = nullptr, uint32_t n_replicas = 0, bool multicast = false, const float* cams_table = nullptr, bool ldr = false) {
  PBRCU_REQUIRE(ctx && g && cam_pos && out_rgba16f, "NULL argument");
  PBRCU_REQUIRE(n_replicas <= PBRCU_MAX_REPLICAS && (n_replicas == 0 || replicas), "too many replicas or replicas NULL");
  for (uint32_t k = 0; k < n_replicas; ++k)
    PBRCU_REQUIRE(replicas[k] && reinterpret_cast<uintptr_t>(replicas[k]) % (ldr ? 4 : 8) == 0, "replica NULL or not aligned to the texel size");
  PBRCU_REQUIRE(g->positions && g->normals && g->albedo && g->orm && g->emissive, "gbuffer plane is NULL (K3 needs the tier-1 planes)");
  PBRCU_REQUIRE(g->pitch8 >= size_t(g->width) * 8 && g->pitch8 % 8 == 0 && g->pitch4 >= size_t(g->width) * 4 && g->pitch4 % 4 == 0, "bad gbuffer pitch");
  const uint32_t texel = ldr ? 4u : 8u;
  PBRCU_REQUIRE(out_pitch >= size_t(g->width) * texel && out_pitch % texel == 0, "out_pitch too small or not a multiple of the texel size");
  PBRCU_REQUIRE(!ldr || (n_views == 1 && !multicast), "the LDR output is for single frames without multicast");
  PBRCU_REQUIRE(n_lights <= 65535 && (n_lights == 0 || lights_dev), "n_lights out of range or lights NULL");
  if (rect.w == 0 && rect.h == 0) rect = pbrcu_rect{0, 0, g->width, g->height};
  PBRCU_REQUIRE(uint64_t(rect.x0) + rect.w <= g->width && uint64_t(rect.y0) + rect.h <= g->height, "rect outside the image");
  if (rect.w == 0 || rect.h == 0) return PBRCU_OK;
  ShadeParams p{};
  p.pos = static_cast<const uint8_t*>(g->positions); p.nrm = static_cast<const uint8_t*>(g->normals);
  p.alb = static_cast<const uint8_t*>(g->albedo); p.orm = static_cast<const uint8_t*>(g->orm);
  p.emi = static_cast<const uint8_t*>(g->emissive); p.out = static_cast<uint8_t*>(out_rgba16f);
  p.n_replicas = n_replicas;
  p.out_is_multicast = multicast ? 1 : 0;
  p.out_texel = texel;
  for (uint32_t k = 0; k < n_replicas; ++k) p.replica[k] = static_cast<uint8_t*>(replicas[k]);
  {
    bool aligned = reinterpret_cast<uintptr_t>(out_rgba16f) % 16 == 0 && out_pitch % 16 == 0;
    for (uint32_t k = 0; k < n_replicas; ++k) aligned = aligned && reinterpret_cast<uintptr_t>(replicas[k]) % 16 == 0;
    const char* e = getenv("PBRCU_PAIR_STORES");  // (0: 8-byte stores only -- the A/B switch of bench.py's fused-gather legs)
    p.pair_stores = aligned && !ldr && !(e && atoi(e) == 0) ? 1 : 0;
  }
  p.pitch8 = g->pitch8; p.pitch4 = g->pitch4; p.out_pitch = out_pitch;
  p.x0 = rect.x0; p.y0 = rect.y0; p.x1 = rect.x0 + rect.w; p.y1 = rect.y0 + rect.h;
  p.tx0 = rect.x0 / kTile; p.ty0 = rect.y0 / kTile;
  p.cx = cam_pos[0]; p.cy = cam_pos[1]; p.cz = cam_pos[2];
  p.n_views = n_views; p.view_stride8 = view_stride8; p.view_stride4 = view_stride4; p.out_view_stride = out_view_stride;
  p.cams = cams_table ? cams_table : ctx->cams_dev;
  p.lights = lights_dev; p.n_lights = n_lights;
  if (tiles) {
    PBRCU_REQUIRE(tiles->counts && tiles->indices, "light tiles not allocated");
    PBRCU_REQUIRE(tiles->tiles_x == (g->width + kTile - 1) / kTile && tiles->tiles_y == (g->height + kTile - 1) / kTile, "tile grid does not match the gbuffer");
    p.tile_counts = tiles->counts; p.tile_indices = tiles->indices; p.tiles_x = tiles->tiles_x; p.max_per_tile = tiles->max_per_tile;
  }
  if (ibl) {
    PBRCU_REQUIRE(ibl->irradiance && ibl->prefiltered && ibl->lut, "ibl texture is NULL");
    PBRCU_REQUIRE(ibl->irr_res > 0 && ibl->pref_res > 0 && ibl->pref_mips >= 1 && ibl->pref_mips <= 16 && (ibl->pref_res >> (ibl->pref_mips - 1)) >= 1 && ibl->lut_res > 0, "bad ibl shape");
    p.has_ibl = 1;
    p.irr = make_chain(ibl->irradiance, ibl->irr_res, 1);
    p.pref = make_chain(ibl->prefiltered, ibl->pref_res, ibl->pref_mips);
    p.lut = reinterpret_cast<const float2*>(ibl->lut);
    p.lut_res = ibl->lut_res;
    if (ibl->accel) {
      const pbrcu_ibl_accel* a = ibl->accel;
      PBRCU_REQUIRE(a->pref_res == ibl->pref_res && a->pref_mips == ibl->pref_mips && a->irr_res == ibl->irr_res && a->lut_res == ibl->lut_res,
                    "ibl->accel was built for textures of another shape");
      p.q_pref = a->pref; p.q_irr = a->irr; p.q_lut = a->lut;
      p.t_pref = a->pref_tex.tex; p.t_irr = a->irr_tex.tex; p.t_lut = a->lut_tex;
    }
  }
  p.ntx = (p.x1 + kTile - 1) / kTile - p.tx0;
  p.ntiles = p.ntx * ((p.y1 + kTile - 1) / kTile - p.ty0);
  PBRCU_REQUIRE(uint64_t(p.ntiles) * p.ntx < (1ull << 31), "rect too large");  // 16x16 tiles: beyond 100k x 100k pixels
  p.ntx_magic = p.ntx == 1 ? 0u : uint32_t(((1ull << 32) + p.ntx - 1) / p.ntx);  // 0: one tile per row
  cudaStream_t st = static_cast<cudaStream_t>(s);
  const bool exact = ctx->exact_ibl_filter || (ibl && (ibl->flags & PBRCU_IBL_EXACT));
  const bool tex = p.q_pref && !exact && (ctx->shade_tex || (ibl && (ibl->flags & PBRCU_IBL_TEXTURE_UNIT)));
  // ---- one light list the HOST holds (the headline frame, the batched views, the facade): lights in the kernel's
  // parameter block, IBL by texture gathers (shade_ggx_cl.cu) ----
  const bool host_lights = n_lights == 0 || is_host_pointer(lights_dev);
  // PBRCU_K3_CL=0 (test switch): never -- upload the list and run shade_ggx_one_kernel.  IBL taps: footprint records (158.6 us on the
  // bench frame), or tld4 gathers where the caller asked for them and the accel carries the atlases (PBRCU_IBL_GATHER: 218 us) --
  // profiles/r02_k3_variants_*.txt, r02_k3_cl_ncu.txt, r02_k3_cl_gather_ncu.txt
  const char* cl_env = getenv("PBRCU_K3_CL");
  const bool use_cl = !(cl_env && atoi(cl_env) == 0);
  const bool gathers = p.has_ibl && ibl->accel && ibl->accel->g_pref.tex && (ibl->flags & PBRCU_IBL_GATHER);
  // (single frames only: over a block of views -- one long launch -- shade_ggx_one_kernel measured faster, 12.1 vs 12.8 ms for 256 x 1080p,
  // scratch/batched_ab.py, so a batched call uploads its host list like any other)
  if (use_cl && n_views == 1 && !tex && p.tile_counts == nullptr && n_lights <= uint32_t(kClMaxLights) && host_lights && (!p.has_ibl || p.q_pref)) {
    int amb = !p.has_ibl ? 0 : (p.t_irr && p.t_lut && !exact) ? 1 : 2;
    if (amb != 0 && !gathers) amb += 2;  // footprint records instead of gathers
    return launch_shade_cl(p, lights_dev, n_lights, ibl ? ibl->accel : nullptr, amb, n_views > 1 ? LIST_ONE_BATCHED : LIST_ONE, ctx->sm_count, st);
  }
  if (n_lights != 0 && host_lights) {  // the kernels below read the list from device memory
    if (ctx->lights_cap < n_lights) {
      if (ctx->lights_dev) PBRCU_CUDA(cudaFree(ctx->lights_dev));
      ctx->lights_dev = nullptr; ctx->lights_cap = 0;
      PBRCU_CUDA(cudaMalloc(&ctx->lights_dev, size_t(n_lights) * sizeof(pbrcu_light)));
      ctx->lights_cap = n_lights;
    }
    PBRCU_CUDA(cudaMemcpyAsync(ctx->lights_dev, lights_dev, size_t(n_lights) * sizeof(pbrcu_light), cudaMemcpyHostToDevice, st));
    p.lights = ctx->lights_dev;
  }
  // persistent grid: kShadeCtasPerSm CTAs per SM walk the half-tile work items round-robin
  const uint32_t grid = uint32_t(std::min<uint64_t>(uint64_t(2) * p.ntiles * n_views, uint64_t(ctx->sm_count) * kShadeCtasPerSm));
  const bool one_list = p.tile_counts == nullptr && p.n_lights <= uint32_t(kLightChunk);
  PBRCU_REQUIRE(n_views == 1 || one_list, "batched shading takes no tile lists and at most 128 lights");
  const int lists = n_views > 1 ? LIST_ONE_BATCHED : one_list ? LIST_ONE : LIST_TILED;
  if (!tex && lists != LIST_TILED && (p.q_pref || !p.has_ibl)) {  // one list, footprint records or no IBL
    // irradiance and LUT through the texture unit unless the caller asked for fp32 weights everywhere
    const bool ambtex = p.has_ibl && p.t_irr && p.t_lut && !exact;
    if (lists == LIST_ONE_BATCHED) {
      if (ambtex) shade_ggx_one_kernel<true, true><<<grid, kShadeThreads, 0, st>>>(p);
      else shade_ggx_one_kernel<true, false><<<grid, kShadeThreads, 0, st>>>(p);
    } else if (p.n_replicas != 0 || p.out_is_multicast || p.out_texel == 4) {
      if (ambtex) shade_ggx_one_kernel<false, true, true><<<grid, kShadeThreads, 0, st>>>(p);
      else shade_ggx_one_kernel<false, false, true><<<grid, kShadeThreads, 0, st>>>(p);
    } else {
      if (ambtex) shade_ggx_one_kernel<false, true><<<grid, kShadeThreads, 0, st>>>(p);
      else shade_ggx_one_kernel<false, false><<<grid, kShadeThreads, 0, st>>>(p);
    }
    return launch_check("shade_ggx_one_kernel");
  }
  if (tex) launch_shade<IBL_TEX>(p, lists, grid, st);
  else if (p.q_pref && p.t_irr && p.t_lut && !exact) launch_shade<IBL_RECORDS_AMBTEX>(p, lists, grid, st);
  else if (p.q_pref) launch_shade<IBL_RECORDS>(p, lists, grid, st);
  else launch_shade<IBL_PLAIN>(p, lists, grid, st);
  return launch_check("shade_ggx_kernel");
}

extern "C" int pbrcu_shade_ggx_ibl(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3],
                                   const pbrcu_light* lights_dev, uint32_t n_lights, const pbrcu_light_tiles* tiles,
                                   const pbrcu_ibl* ibl, pbrcu_rect rect, void* out_rgba16f, size_t out_pitch, pbrcu_stream s) {
  return shade_common(ctx, g, cam_pos, lights_dev, n_lights, tiles, ibl, rect, out_rgba16f, out_pitch, 1, 0, 0, 0, s);
}

extern "C" int pbrcu_shade_ggx_ibl_replicated(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3],
                                              const pbrcu_light* lights_dev, uint32_t n_lights, const pbrcu_light_tiles* tiles,
                                              const pbrcu_ibl* ibl, pbrcu_rect rect, void* out_rgba16f, size_t out_pitch,
                                              void* const* replica_outs, uint32_t n_replicas, pbrcu_stream s) {
  return shade_common(ctx, g, cam_pos, lights_dev, n_lights, tiles, ibl, rect, out_rgba16f, out_pitch, 1, 0, 0, 0, s, replica_outs, n_replicas);
}

extern "C" int pbrcu_shade_ggx_ibl_multicast(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3],
                                             const pbrcu_light* lights_dev, uint32_t n_lights, const pbrcu_light_tiles* tiles,
                                             const pbrcu_ibl* ibl, pbrcu_rect rect, void* out_multicast, size_t out_pitch, pbrcu_stream s) {
  return shade_common(ctx, g, cam_pos, lights_dev, n_lights, tiles, ibl, rect, out_multicast, out_pitch, 1, 0, 0, 0, s, nullptr, 0, true);
}

extern "C" int pbrcu_shade_ggx_ibl_batched(pbrcu_ctx* ctx, const pbrcu_gbuffer* g0, size_t view_stride8, size_t view_stride4,
                                           uint32_t n_views, const float* cam_pos_xyz, const pbrcu_light* lights_dev, uint32_t n_lights,
                                           const pbrcu_ibl* ibl, void* out_rgba16f, size_t out_pitch, size_t out_view_stride, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && g0 && cam_pos_xyz && out_rgba16f, "NULL argument");
  PBRCU_REQUIRE(n_views >= 1 && n_views <= 65535, "n_views must be in [1, 65535]");
  PBRCU_REQUIRE(view_stride8 >= g0->pitch8 * g0->height && view_stride4 >= g0->pitch4 * g0->height && out_view_stride >= out_pitch * g0->height,
                "view stride smaller than one view");
  PBRCU_REQUIRE(view_stride8 % 8 == 0 && view_stride4 % 4 == 0 && out_view_stride % 8 == 0, "view strides must keep texels aligned");
  if (n_views == 1)
    return shade_common(ctx, g0, cam_pos_xyz, lights_dev, n_lights, nullptr, ibl, pbrcu_rect{0, 0, 0, 0}, out_rgba16f, out_pitch, 1, 0, 0, 0, s);
  if (ctx->cams_cap < n_views) {
    if (ctx->cams_dev) PBRCU_CUDA(cudaFree(ctx->cams_dev));
    ctx->cams_dev = nullptr; ctx->cams_cap = 0;
    PBRCU_CUDA(cudaMalloc(&ctx->cams_dev, size_t(n_views) * 3 * sizeof(float)));
    ctx->cams_cap = n_views;
  }
  PBRCU_CUDA(cudaMemcpyAsync(ctx->cams_dev, cam_pos_xyz, size_t(n_views) * 3 * sizeof(float), cudaMemcpyHostToDevice, static_cast<cudaStream_t>(s)));
  return shade_common(ctx, g0, cam_pos_xyz, lights_dev, n_lights, nullptr, ibl, pbrcu_rect{0, 0, 0, 0}, out_rgba16f, out_pitch, n_views, view_stride8,
                      view_stride4, out_view_stride, s);
}

// ---- pbrcu_shade_ggx_ibl_batched_static: the batched launch for callers that must not allocate ----
// The batched kernel reads one camera position per view from a table in device memory.  pbrcu_shade_ggx_ibl_batched keeps
// that table in context scratch grown with cudaMalloc; here it lives in a file-scope device array of a fixed, modest size
// (PBRCU_STATIC_MAX_VIEWS views x 3 floats = 48 KB, held by every GPU that loads the library until exit).  Every call
// shares the array and overwrites the entries it uses (nothing in it has to start at zero); two calls that overlap on
// the GPU -- different streams, or different contexts of one device -- race on it.
static __device__ __align__(256) float g_static_cams[3 * PBRCU_STATIC_MAX_VIEWS];

// In host code `g_static_cams` is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress,
// once per device, with that device current.
static int static_cams_address(pbrcu_ctx* ctx, float** out) {
  constexpr int kMaxDevices = 64;
  static std::mutex lock;
  static float* address[kMaxDevices] = {};
  PBRCU_REQUIRE(ctx->device >= 0 && ctx->device < kMaxDevices, "device ordinal out of range for the static scratch");
  std::lock_guard<std::mutex> hold(lock);
  if (!address[ctx->device]) {
    int current = -1;
    PBRCU_CUDA(cudaGetDevice(&current));
    if (current != ctx->device) PBRCU_CUDA(cudaSetDevice(ctx->device));
    float* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_static_cams);
    if (current != ctx->device) cudaSetDevice(current);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_static_cams)");
    PBRCU_REQUIRE(ptr != nullptr, "static scratch has no device address");
    address[ctx->device] = ptr;
  }
  *out = address[ctx->device];
  return PBRCU_OK;
}

extern "C" int pbrcu_shade_ggx_ibl_batched_static(pbrcu_ctx* ctx, const pbrcu_gbuffer* g0, size_t view_stride8, size_t view_stride4,
                                                  uint32_t n_views, const float* cam_pos_xyz, const pbrcu_light* lights_dev, uint32_t n_lights,
                                                  const pbrcu_ibl* ibl, void* out_rgba16f, size_t out_pitch, size_t out_view_stride, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && g0 && cam_pos_xyz && out_rgba16f, "NULL argument");
  PBRCU_REQUIRE(n_views >= 1, "n_views must be at least 1");
  if (n_views > PBRCU_STATIC_MAX_VIEWS) {
    set_error("invalid argument: n_views = %u exceeds PBRCU_STATIC_MAX_VIEWS = %u, the size of the static camera table "
              "(use pbrcu_shade_ggx_ibl_batched, or split the views over several calls)", n_views, unsigned(PBRCU_STATIC_MAX_VIEWS));
    return PBRCU_ERR_INVALID;
  }
  // a host light list would be uploaded into context scratch that may have to grow: this entry point allocates nothing
  PBRCU_REQUIRE(n_lights == 0 || (lights_dev && !is_host_pointer(lights_dev)), "lights must be a DEVICE array for the no-allocation entry point");
  PBRCU_REQUIRE(view_stride8 >= g0->pitch8 * g0->height && view_stride4 >= g0->pitch4 * g0->height && out_view_stride >= out_pitch * g0->height,
                "view stride smaller than one view");
  PBRCU_REQUIRE(view_stride8 % 8 == 0 && view_stride4 % 4 == 0 && out_view_stride % 8 == 0, "view strides must keep texels aligned");
  PBRCU_REQUIRE(n_lights <= uint32_t(kLightChunk), "batched shading takes at most 128 lights");
  float* cams = nullptr;
  if (int rc = static_cams_address(ctx, &cams)) return rc;
  if (n_views == 1)  // a single view takes its camera from the kernel parameters
    return shade_common(ctx, g0, cam_pos_xyz, lights_dev, n_lights, nullptr, ibl, pbrcu_rect{0, 0, 0, 0}, out_rgba16f, out_pitch, 1, 0, 0, 0, s);
  PBRCU_CUDA(cudaMemcpyAsync(cams, cam_pos_xyz, size_t(n_views) * 3 * sizeof(float), cudaMemcpyHostToDevice, static_cast<cudaStream_t>(s)));
  return shade_common(ctx, g0, cam_pos_xyz, lights_dev, n_lights, nullptr, ibl, pbrcu_rect{0, 0, 0, 0}, out_rgba16f, out_pitch, n_views, view_stride8,
                      view_stride4, out_view_stride, s, nullptr, 0, false, cams);
}

// K3 with the reference's tonemap pass (K2: TonemapperSystem::run, src/pbr_engine/engine/pbr/TonemapperSystem.cpp:97-134, aces+gamma.glsl:30-40)
// fused into its stores: the display-ready RGBA8 frame in one pass -- 40 B/px instead of 44 + 12 for K3 then K2 -- bit-identical to
// pbrcu_shade_ggx_ibl followed by pbrcu_tonemap_aces_gamma (the tonemapper sees the fp16-rounded value either way).  replica_outs / n_replicas
// as in pbrcu_shade_ggx_ibl_replicated (NULL, 0: none): a multi-GPU row band delivered to every peer at 4 bytes per texel.
extern "C" int pbrcu_shade_ggx_ibl_ldr(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3], const pbrcu_light* lights, uint32_t n_lights,
                                       const pbrcu_light_tiles* tiles, const pbrcu_ibl* ibl, pbrcu_rect rect, void* out_rgba8, size_t out_pitch,
                                       void* const* replica_outs, uint32_t n_replicas, pbrcu_stream s) {
  return shade_common(ctx, g, cam_pos, lights, n_lights, tiles, ibl, rect, out_rgba8, out_pitch, 1, 0, 0, 0, s, replica_outs, n_replicas, false, nullptr, true);
}
