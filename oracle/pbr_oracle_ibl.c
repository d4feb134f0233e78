/* TEST INFRASTRUCTURE ONLY -- see pbr_oracle.h.  Tier B: GGX/Smith/Schlick shading and the IBL
 * precompute.  NO REFERENCE COUNTERPART (the reference shades with Blinn-Phong and has no IBL code:
 * SURVEY.md section 0) => PARITY UNPINNED vs. the reference.  These functions restate published
 * formulas: glTF 2.0 spec Appendix B (BRDF), KHR_lights_punctual (point-light falloff), Karis 2013
 * "Real Shading in Unreal Engine 4" (PrefilterEnvMap, IntegrateBRDF), Krivanek & Colbert 2008
 * (filtered importance sampling, GPU Gems 3 ch. 20), Hammersley via 32-bit radical inverse.
 * The in-repo conventions they fix (cube face order, texel addressing, filtering) are in DESIGN.md.
 */
#include "pbr_oracle.h"
#include "pbr_detmath.h"

#include <math.h>
#include <omp.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------- */
void pbro_sincos2pi(float t, float* s, float* c) { dm_sincos2pi(t, s, c); }
uint32_t pbro_radical_inverse_bits(uint32_t i) { return dm_reverse_bits32(i); }
void pbro_hammersley(uint32_t i, uint32_t n, float out2[2]) { dm_hammersley(i, n, &out2[0], &out2[1]); }
void pbro_hammersley_table(uint32_t n, float* out_xy) {
  for (uint32_t i = 0; i < n; ++i) dm_hammersley(i, n, &out_xy[2 * i], &out_xy[2 * i + 1]);
}

/* ---------------------------------------------------------------------------------------------
 * Split-sum BRDF LUT -- Karis 2013, IntegrateBRDF.  u = NoV = (i+0.5)/res, v = roughness =
 * (j+0.5)/res, N = +Z, k = alpha/2 (the IBL form of Schlick-Smith), Fc = (1-VoH)^5.
 * Sequential sum over the Hammersley index; every operation a single fp32 op (bit-exact contract).
 * ------------------------------------------------------------------------------------------- */
void pbro_brdf_lut_texel(int res, int nsamples, int i, int j, float out2[2]) {
  float NoV = ((float)i + 0.5f) / (float)res;
  float rough = ((float)j + 0.5f) / (float)res;
  float alpha = rough * rough;
  float k = alpha * 0.5f;
  float omk = 1.0f - k;
  float Vx = sqrtf(1.0f - NoV * NoV), Vz = NoV;
  float g1v = NoV / (NoV * omk + k);
  float A = 0.0f, B = 0.0f;
  for (int s = 0; s < nsamples; ++s) {
    float x0, x1, H[3];
    dm_hammersley((uint32_t)s, (uint32_t)nsamples, &x0, &x1);
    dm_importance_sample_ggx(x0, x1, alpha, H);
    float VoH = Vx * H[0] + Vz * H[2];
    float Lz = (2.0f * VoH) * H[2] - Vz;
    float NoL = fmaxf(Lz, 0.0f), NoH = fmaxf(H[2], 0.0f), VoHc = fmaxf(VoH, 0.0f);
    if (NoL > 0.0f) {
      float g1l = NoL / (NoL * omk + k);
      float G = g1v * g1l;
      float Gvis = (G * VoHc) / (NoH * NoV);
      float t = 1.0f - VoHc, t2 = t * t;
      float Fc = (t2 * t2) * t;
      A = A + (1.0f - Fc) * Gvis;
      B = B + Fc * Gvis;
    }
  }
  float invn = 1.0f / (float)nsamples;
  out2[0] = A * invn;
  out2[1] = B * invn;
}

void pbro_brdf_lut(int res, int nsamples, float* out_rg, int nthreads) {
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads)
  for (int j = 0; j < res; ++j)
    for (int i = 0; i < res; ++i) pbro_brdf_lut_texel(res, nsamples, i, j, out_rg + 2 * ((size_t)j * res + i));
}

/* ---------------------------------------------------------------------------------------------
 * Cube maps.  Face order +X,-X,+Y,-Y,+Z,-Z with the OpenGL/Vulkan (sc,tc,ma) convention.
 * ------------------------------------------------------------------------------------------- */
size_t pbro_cube_mip_offset(int base_res, int mip) {
  size_t off = 0;
  for (int m = 0; m < mip; ++m) {
    size_t r = (size_t)(base_res >> m);
    off += 6 * r * r;
  }
  return off;
}
size_t pbro_cube_chain_texels(int base_res, int nmips) { return pbro_cube_mip_offset(base_res, nmips); }

void pbro_cube_texel_dir(int face, int x, int y, int res, float dir[3]) {
  float sc = (2.0f * (float)x + 1.0f) / (float)res - 1.0f;
  float tc = (2.0f * (float)y + 1.0f) / (float)res - 1.0f;
  switch (face) {
    case 0: dir[0] = 1.0f;  dir[1] = -tc;  dir[2] = -sc;  break;
    case 1: dir[0] = -1.0f; dir[1] = -tc;  dir[2] = sc;   break;
    case 2: dir[0] = sc;    dir[1] = 1.0f; dir[2] = tc;   break;
    case 3: dir[0] = sc;    dir[1] = -1.0f; dir[2] = -tc; break;
    case 4: dir[0] = sc;    dir[1] = -tc;  dir[2] = 1.0f; break;
    default: dir[0] = -sc;  dir[1] = -tc;  dir[2] = -1.0f; break;
  }
}

void pbro_cube_dir_to_face_uv(const float d[3], int* face, float* u, float* v) {
  float ax = fabsf(d[0]), ay = fabsf(d[1]), az = fabsf(d[2]);
  float sc, tc, ma;
  if (ax >= ay && ax >= az) {
    ma = ax;
    if (d[0] >= 0.0f) { *face = 0; sc = -d[2]; tc = -d[1]; }
    else              { *face = 1; sc = d[2];  tc = -d[1]; }
  } else if (ay >= az) {
    ma = ay;
    if (d[1] >= 0.0f) { *face = 2; sc = d[0]; tc = d[2]; }
    else              { *face = 3; sc = d[0]; tc = -d[2]; }
  } else {
    ma = az;
    if (d[2] >= 0.0f) { *face = 4; sc = d[0];  tc = -d[1]; }
    else              { *face = 5; sc = -d[0]; tc = -d[1]; }
  }
  float inv = 1.0f / ma;
  *u = 0.5f * (sc * inv + 1.0f);
  *v = 0.5f * (tc * inv + 1.0f);
}

static inline float lerpf(float a, float b, float t) { return a + t * (b - a); }

/* bilinear within one face of one level, clamp-to-edge at the face border (no cross-face filtering) */
static void face_bilinear(const uint16_t* level, int res, int face, float u, float v, float out4[4]) {
  float fx = u * (float)res - 0.5f, fy = v * (float)res - 0.5f;
  float x0f = floorf(fx), y0f = floorf(fy);
  float tx = fx - x0f, ty = fy - y0f;
  int x0 = (int)x0f, y0 = (int)y0f, x1 = x0 + 1, y1 = y0 + 1;
  int hi = res - 1;
  x0 = x0 < 0 ? 0 : (x0 > hi ? hi : x0); x1 = x1 < 0 ? 0 : (x1 > hi ? hi : x1);
  y0 = y0 < 0 ? 0 : (y0 > hi ? hi : y0); y1 = y1 < 0 ? 0 : (y1 > hi ? hi : y1);
  const uint16_t* f = level + (size_t)face * res * res * 4;
  const uint16_t* t00 = f + ((size_t)y0 * res + x0) * 4;
  const uint16_t* t10 = f + ((size_t)y0 * res + x1) * 4;
  const uint16_t* t01 = f + ((size_t)y1 * res + x0) * 4;
  const uint16_t* t11 = f + ((size_t)y1 * res + x1) * 4;
  for (int c = 0; c < 4; ++c) {
    float a = lerpf(pbro_half_to_float(t00[c]), pbro_half_to_float(t10[c]), tx);
    float b = lerpf(pbro_half_to_float(t01[c]), pbro_half_to_float(t11[c]), tx);
    out4[c] = lerpf(a, b, ty);
  }
}

/* textureLod(cube, dir, lod) with exact fp32 weights: trilinear = lerp of two per-face bilinears */
void pbro_cube_sample(const uint16_t* chain, int base_res, int nmips, const float dir[3], float lod,
                      float out4[4]) {
  int face;
  float u, v;
  pbro_cube_dir_to_face_uv(dir, &face, &u, &v);
  float top = (float)(nmips - 1);
  lod = fminf(fmaxf(lod, 0.0f), top);
  float l0f = floorf(lod);
  float f = lod - l0f;
  int l0 = (int)l0f, l1 = l0 + 1 < nmips ? l0 + 1 : nmips - 1;
  float c0[4];
  face_bilinear(chain + 4 * pbro_cube_mip_offset(base_res, l0), base_res >> l0, face, u, v, c0);
  if (f > 0.0f) {
    float c1[4];
    face_bilinear(chain + 4 * pbro_cube_mip_offset(base_res, l1), base_res >> l1, face, u, v, c1);
    for (int c = 0; c < 4; ++c) out4[c] = lerpf(c0[c], c1[c], f);
  } else {
    memcpy(out4, c0, 16);
  }
}

static inline void normalize3(const float v[3], float o[3]) {
  float inv = 1.0f / sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  o[0] = v[0] * inv; o[1] = v[1] * inv; o[2] = v[2] * inv;
}

/* K4: equirect (lat-long, RGBA32F, +Y up at v=0) -> cube base level.  u = atan2(z,x)/2pi + 0.5,
 * v = acos(y)/pi; bilinear with wrap in u and clamp in v. */
void pbro_equirect_to_cube(const float* src, int sw, int sh, int res, uint16_t* out_base, int nthreads) {
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(nthreads) collapse(2)
  for (int face = 0; face < 6; ++face)
    for (int y = 0; y < res; ++y)
      for (int x = 0; x < res; ++x) {
        float d[3], n[3];
        pbro_cube_texel_dir(face, x, y, res, d);
        normalize3(d, n);
        float u = atan2f(n[2], n[0]) * (0.5f / DM_PI) + 0.5f;
        float v = acosf(fminf(fmaxf(n[1], -1.0f), 1.0f)) * (1.0f / DM_PI);
        float fx = u * (float)sw - 0.5f, fy = v * (float)sh - 0.5f;
        float x0f = floorf(fx), y0f = floorf(fy);
        float tx = fx - x0f, ty = fy - y0f;
        int x0 = (int)x0f, y0 = (int)y0f, x1 = x0 + 1, y1 = y0 + 1;
        x0 = ((x0 % sw) + sw) % sw; x1 = ((x1 % sw) + sw) % sw;
        y0 = y0 < 0 ? 0 : (y0 > sh - 1 ? sh - 1 : y0); y1 = y1 < 0 ? 0 : (y1 > sh - 1 ? sh - 1 : y1);
        const float* t00 = src + ((size_t)y0 * sw + x0) * 4;
        const float* t10 = src + ((size_t)y0 * sw + x1) * 4;
        const float* t01 = src + ((size_t)y1 * sw + x0) * 4;
        const float* t11 = src + ((size_t)y1 * sw + x1) * 4;
        uint16_t* o = out_base + (((size_t)face * res + y) * res + x) * 4;
        for (int c = 0; c < 4; ++c) {
          float a = lerpf(t00[c], t10[c], tx), b = lerpf(t01[c], t11[c], tx);
          o[c] = pbro_float_to_half(fminf(lerpf(a, b, ty), 65504.0f));
        }
      }
}

/* K5: 2x2 box filter per face, each level from the previous STORED (fp16) level:
 * ((t00 + t10) + (t01 + t11)) * 0.25 in fp32, one RNE rounding at the store.  Bit-exact contract. */
void pbro_cube_mipgen(uint16_t* chain, int base_res, int nmips, int nthreads) {
  if (nthreads <= 0) nthreads = omp_get_max_threads();
  for (int m = 1; m < nmips; ++m) {
    int rs = base_res >> (m - 1), rd = base_res >> m;
    const uint16_t* src = chain + 4 * pbro_cube_mip_offset(base_res, m - 1);
    uint16_t* dst = chain + 4 * pbro_cube_mip_offset(base_res, m);
#pragma omp parallel for schedule(static) num_threads(nthreads) collapse(2)
    for (int face = 0; face < 6; ++face)
      for (int y = 0; y < rd; ++y)
        for (int x = 0; x < rd; ++x) {
          const uint16_t* s0 = src + (((size_t)face * rs + 2 * y) * rs + 2 * x) * 4;
          const uint16_t* s1 = s0 + (size_t)rs * 4;
          uint16_t* o = dst + (((size_t)face * rd + y) * rd + x) * 4;
          for (int c = 0; c < 4; ++c) {
            float a = pbro_half_to_float(s0[c]) + pbro_half_to_float(s0[4 + c]);
            float b = pbro_half_to_float(s1[c]) + pbro_half_to_float(s1[4 + c]);
            o[c] = pbro_float_to_half((a + b) * 0.25f);
          }
        }
  }
}

/* ---------------------------------------------------------------------------------------------
 * K7: GGX prefilter.  Under N = V = R the sample set in the local frame depends only on
 * (roughness, i): L_local = 2 (V.H) H - V with V = +Z; weight = N.L = L_local.z; source level from
 * filtered importance sampling: lod = max(0.5 log2(Ws/Wp), 0), Ws = 1/(n pdf), pdf = D NoH/(4 VoH) =
 * D/4, Wp = 4 pi / (6 base_res^2).  Samples with N.L <= 0 are dropped (Karis: `if (NoL > 0)`).
 * ------------------------------------------------------------------------------------------- */
int pbro_prefilter_table(float roughness, int nsamples, int base_res, int nmips, float* table4) {
  float alpha = roughness * roughness;
  float a2 = alpha * alpha;
  float omegaP = (4.0f * DM_PI) / (6.0f * (float)base_res * (float)base_res);
  int n = 0;
  for (int i = 0; i < nsamples; ++i) {
    float x0, x1, H[3];
    dm_hammersley((uint32_t)i, (uint32_t)nsamples, &x0, &x1);
    dm_importance_sample_ggx(x0, x1, alpha, H);
    float c = H[2];
    float Lx = (2.0f * c) * H[0], Ly = (2.0f * c) * H[1], Lz = (2.0f * c) * c - 1.0f;
    if (Lz > 0.0f) {
      float dd = (c * c) * (a2 - 1.0f) + 1.0f;
      float D = a2 / (DM_PI * dd * dd);
      float pdf = D * 0.25f;
      float omegaS = 1.0f / ((float)nsamples * pdf);
      float lod = 0.5f * log2f(omegaS / omegaP);
      lod = fminf(fmaxf(lod, 0.0f), (float)(nmips - 1));
      table4[4 * n + 0] = Lx; table4[4 * n + 1] = Ly; table4[4 * n + 2] = Lz; table4[4 * n + 3] = lod;
      ++n;
    }
  }
  return n;
}

/* tangent frame of Karis' ImportanceSampleGGX: up = |N.z| < 0.999 ? +Z : +X */
static void tangent_frame(const float N[3], float T[3], float B[3]) {
  float up[3] = {0.0f, 0.0f, 1.0f};
  if (!(fabsf(N[2]) < 0.999f)) { up[0] = 1.0f; up[2] = 0.0f; }
  float t[3] = {up[1] * N[2] - up[2] * N[1], up[2] * N[0] - up[0] * N[2], up[0] * N[1] - up[1] * N[0]};
  normalize3(t, T);
  B[0] = N[1] * T[2] - N[2] * T[1];
  B[1] = N[2] * T[0] - N[0] * T[2];
  B[2] = N[0] * T[1] - N[1] * T[0];
}

static void convolve_texel(const uint16_t* src_chain, int base_res, int nmips, const float* table4,
                           int n, int weighted, const float dirn[3], float out4[4]) {
  float T[3], B[3];
  tangent_frame(dirn, T, B);
  float acc[3] = {0.0f, 0.0f, 0.0f}, wsum = 0.0f;
  for (int s = 0; s < n; ++s) {
    const float* e = table4 + 4 * s;
    float L[3] = {T[0] * e[0] + B[0] * e[1] + dirn[0] * e[2], T[1] * e[0] + B[1] * e[1] + dirn[1] * e[2],
                  T[2] * e[0] + B[2] * e[1] + dirn[2] * e[2]};
    float c[4];
    pbro_cube_sample(src_chain, base_res, nmips, L, e[3], c);
    float w = weighted ? e[2] : 1.0f;
    acc[0] += c[0] * w; acc[1] += c[1] * w; acc[2] += c[2] * w;
    wsum += w;
  }
  float inv = 1.0f / wsum;
  out4[0] = acc[0] * inv; out4[1] = acc[1] * inv; out4[2] = acc[2] * inv; out4[3] = 1.0f;
}

void pbro_prefilter_texel(const uint16_t* src_chain, int base_res, int nmips, int mip, int nsamples,
                          int face, int x, int y, float out4[4]) {
  float* table = (float*)malloc(sizeof(float) * 4 * (size_t)nsamples);
  float rough = (float)mip / (float)(nmips - 1);
  int n = pbro_prefilter_table(rough, nsamples, base_res, nmips, table);
  float d[3], dn[3];
  pbro_cube_texel_dir(face, x, y, base_res >> mip, d);
  normalize3(d, dn);
  convolve_texel(src_chain, base_res, nmips, table, n, 1, dn, out4);
  free(table);
}

void pbro_prefilter_ggx(const uint16_t* src_chain, int base_res, int nmips, int nsamples,
                        uint16_t* dst_chain, int first_mip, int last_mip, int nthreads) {
  if (nthreads <= 0) nthreads = omp_get_max_threads();
  float* table = (float*)malloc(sizeof(float) * 4 * (size_t)nsamples);
  for (int m = first_mip; m <= last_mip; ++m) {
    int res = base_res >> m;
    size_t off = 4 * pbro_cube_mip_offset(base_res, m);
    if (m == 0) { /* roughness 0: the lobe is a delta -> exact copy of the source level */
      memcpy(dst_chain + off, src_chain + off, sizeof(uint16_t) * 4 * 6 * (size_t)res * res);
      continue;
    }
    float rough = (float)m / (float)(nmips - 1);
    int n = pbro_prefilter_table(rough, nsamples, base_res, nmips, table);
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads) collapse(2)
    for (int face = 0; face < 6; ++face)
      for (int y = 0; y < res; ++y)
        for (int x = 0; x < res; ++x) {
          float d[3], dn[3], o[4];
          pbro_cube_texel_dir(face, x, y, res, d);
          normalize3(d, dn);
          convolve_texel(src_chain, base_res, nmips, table, n, 1, dn, o);
          uint16_t* t = dst_chain + off + (((size_t)face * res + y) * res + x) * 4;
          for (int c = 0; c < 4; ++c) t[c] = pbro_float_to_half(fminf(o[c], 65504.0f));
        }
  }
  free(table);
}

/* K6: irradiance, stored as E/pi (so diffuse = irradiance * c_diff).  Cosine-weighted importance
 * sampling (pdf = cos/pi -> plain average), filtered: lod from Ws = pi/(n cos). */
static int irradiance_table(int nsamples, int base_res, int nmips, float* table4) {
  float omegaP = (4.0f * DM_PI) / (6.0f * (float)base_res * (float)base_res);
  for (int i = 0; i < nsamples; ++i) {
    float x0, x1, sp, cp;
    dm_hammersley((uint32_t)i, (uint32_t)nsamples, &x0, &x1);
    dm_sincos2pi(x0, &sp, &cp);
    float r = sqrtf(x1), cz = sqrtf(1.0f - x1);
    float omegaS = DM_PI / ((float)nsamples * cz);
    float lod = 0.5f * log2f(omegaS / omegaP);
    lod = fminf(fmaxf(lod, 0.0f), (float)(nmips - 1));
    table4[4 * i + 0] = r * cp; table4[4 * i + 1] = r * sp; table4[4 * i + 2] = cz; table4[4 * i + 3] = lod;
  }
  return nsamples;
}

void pbro_irradiance_texel(const uint16_t* src_chain, int base_res, int nmips, int res, int nsamples,
                           int face, int x, int y, float out4[4]) {
  float* table = (float*)malloc(sizeof(float) * 4 * (size_t)nsamples);
  int n = irradiance_table(nsamples, base_res, nmips, table);
  float d[3], dn[3];
  pbro_cube_texel_dir(face, x, y, res, d);
  normalize3(d, dn);
  convolve_texel(src_chain, base_res, nmips, table, n, 0, dn, out4);
  free(table);
}

void pbro_irradiance(const uint16_t* src_chain, int base_res, int nmips, int res, int nsamples,
                     uint16_t* out, int nthreads) {
  if (nthreads <= 0) nthreads = omp_get_max_threads();
  float* table = (float*)malloc(sizeof(float) * 4 * (size_t)nsamples);
  int n = irradiance_table(nsamples, base_res, nmips, table);
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads) collapse(2)
  for (int face = 0; face < 6; ++face)
    for (int y = 0; y < res; ++y)
      for (int x = 0; x < res; ++x) {
        float d[3], dn[3], o[4];
        pbro_cube_texel_dir(face, x, y, res, d);
        normalize3(d, dn);
        convolve_texel(src_chain, base_res, nmips, table, n, 0, dn, o);
        uint16_t* t = out + (((size_t)face * res + y) * res + x) * 4;
        for (int c = 0; c < 4; ++c) t[c] = pbro_float_to_half(fminf(o[c], 65504.0f));
      }
  free(table);
}

/* ---------------------------------------------------------------------------------------------
 * K3: metallic-roughness Cook-Torrance shading, glTF 2.0 Appendix B, + KHR_lights_punctual point
 * lights + split-sum IBL.  Extended G-buffer: reference planes (positions, normals, albedo=baseColor)
 * + ORM RGBA8 (R occlusion, G roughness, B metallic) + emissive RGBA16F.
 * ------------------------------------------------------------------------------------------- */
static inline float clamp01(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }

static void lut_sample(const float* lut, int res, float u, float v, float* A, float* B) {
  float fx = u * (float)res - 0.5f, fy = v * (float)res - 0.5f;
  float x0f = floorf(fx), y0f = floorf(fy);
  float tx = fx - x0f, ty = fy - y0f;
  int x0 = (int)x0f, y0 = (int)y0f, x1 = x0 + 1, y1 = y0 + 1, hi = res - 1;
  x0 = x0 < 0 ? 0 : (x0 > hi ? hi : x0); x1 = x1 < 0 ? 0 : (x1 > hi ? hi : x1);
  y0 = y0 < 0 ? 0 : (y0 > hi ? hi : y0); y1 = y1 < 0 ? 0 : (y1 > hi ? hi : y1);
  const float* t00 = lut + 2 * ((size_t)y0 * res + x0);
  const float* t10 = lut + 2 * ((size_t)y0 * res + x1);
  const float* t01 = lut + 2 * ((size_t)y1 * res + x0);
  const float* t11 = lut + 2 * ((size_t)y1 * res + x1);
  *A = lerpf(lerpf(t00[0], t10[0], tx), lerpf(t01[0], t11[0], tx), ty);
  *B = lerpf(lerpf(t00[1], t10[1], tx), lerpf(t01[1], t11[1], tx), ty);
}

void pbro_shade_pixel(const uint16_t pos_h[4], const uint16_t nrm_h[4], const uint16_t alb_u[4],
                      const uint8_t orm[4], const uint16_t emi_h[4], const float cam_pos[3],
                      const pbro_light* lights, int nlights, const pbro_ibl* ibl, float out4[4]) {
  float P[3], Nr[3], base[4], emi[3];
  for (int c = 0; c < 3; ++c) {
    P[c] = pbro_half_to_float(pos_h[c]);
    Nr[c] = pbro_half_to_float(nrm_h[c]);
    emi[c] = pbro_half_to_float(emi_h[c]);
  }
  for (int c = 0; c < 4; ++c) base[c] = (float)alb_u[c] / 65535.0f;
  if (Nr[0] == 0.0f && Nr[1] == 0.0f && Nr[2] == 0.0f) { /* cleared (background) texel */
    out4[0] = out4[1] = out4[2] = out4[3] = 0.0f;
    return;
  }
  float N[3], Vv[3] = {cam_pos[0] - P[0], cam_pos[1] - P[1], cam_pos[2] - P[2]}, V[3];
  normalize3(Nr, N);
  normalize3(Vv, V);
  float NoVraw = N[0] * V[0] + N[1] * V[1] + N[2] * V[2];
  float NoV = fminf(fmaxf(NoVraw, 1e-4f), 1.0f);
  float occ = (float)orm[0] / 255.0f;
  float rough = fminf(fmaxf((float)orm[1] / 255.0f, 0.045f), 1.0f);
  float metal = (float)orm[2] / 255.0f;
  float F0[3], cdiff[3];
  for (int c = 0; c < 3; ++c) {
    F0[c] = 0.04f * (1.0f - metal) + base[c] * metal; /* mix(0.04, baseColor, metallic) */
    cdiff[c] = base[c] * (1.0f - metal);
  }
  float alpha = rough * rough, a2 = alpha * alpha, oma2 = 1.0f - a2;
  float Lo[3] = {0.0f, 0.0f, 0.0f};
  float visV = sqrtf(NoV * NoV * oma2 + a2);
  for (int i = 0; i < nlights; ++i) {
    const pbro_light* lt = &lights[i];
    float Lv[3] = {lt->px - P[0], lt->py - P[1], lt->pz - P[2]};
    float d2 = fmaxf(Lv[0] * Lv[0] + Lv[1] * Lv[1] + Lv[2] * Lv[2], 1e-8f); /* light inside the surface point */
    float invd = 1.0f / sqrtf(d2);
    float Ld[3] = {Lv[0] * invd, Lv[1] * invd, Lv[2] * invd};
    float NoL = clamp01(N[0] * Ld[0] + N[1] * Ld[1] + N[2] * Ld[2]);
    float att = 1.0f / d2; /* KHR_lights_punctual: inverse-square falloff */
    if (lt->range > 0.0f) {
      float q = d2 / (lt->range * lt->range);
      att *= clamp01(1.0f - q * q);
    }
    if (NoL > 0.0f && att > 0.0f) {
      float Hv[3] = {V[0] + Ld[0], V[1] + Ld[1], V[2] + Ld[2]}, H[3];
      normalize3(Hv, H);
      float NoH = clamp01(N[0] * H[0] + N[1] * H[1] + N[2] * H[2]);
      float VoH = clamp01(V[0] * H[0] + V[1] * H[1] + V[2] * H[2]);
      /* GGX D with 1 - NoH^2 taken from |N x H|^2 (Lagrange identity): NoH^2 (a2 - 1) + 1 cancels
       * catastrophically in fp32 for smooth surfaces (a2 ~ 4e-6), this form does not */
      float cxh[3] = {N[1] * H[2] - N[2] * H[1], N[2] * H[0] - N[0] * H[2], N[0] * H[1] - N[1] * H[0]};
      float dd = (cxh[0] * cxh[0] + cxh[1] * cxh[1] + cxh[2] * cxh[2]) + NoH * NoH * a2;
      float D = a2 / (DM_PI * dd * dd);
      float vis = 0.5f / (NoL * visV + NoV * sqrtf(NoL * NoL * oma2 + a2));
      float t = 1.0f - VoH, t2 = t * t, fc = t2 * t2 * t;
      float s = lt->intensity * att * NoL;
      for (int c = 0; c < 3; ++c) {
        float F = F0[c] + (1.0f - F0[c]) * fc;
        float rad = (c == 0 ? lt->r : (c == 1 ? lt->g : lt->b)) * s;
        Lo[c] += ((1.0f - F) * cdiff[c] * (1.0f / DM_PI) + F * (D * vis)) * rad;
      }
    }
  }
  float amb[3] = {0.0f, 0.0f, 0.0f};
  if (ibl) {
    float R[3] = {2.0f * NoVraw * N[0] - V[0], 2.0f * NoVraw * N[1] - V[1], 2.0f * NoVraw * N[2] - V[2]};
    float irr[4], pre[4], A, B;
    pbro_cube_sample(ibl->irradiance, ibl->irr_res, 1, N, 0.0f, irr);
    pbro_cube_sample(ibl->prefiltered, ibl->pref_res, ibl->pref_mips, R,
                     rough * (float)(ibl->pref_mips - 1), pre);
    lut_sample(ibl->lut, ibl->lut_res, NoV, rough, &A, &B);
    for (int c = 0; c < 3; ++c) amb[c] = (irr[c] * cdiff[c] + pre[c] * (F0[c] * A + B)) * occ;
  }
  for (int c = 0; c < 3; ++c) out4[c] = fminf(Lo[c] + amb[c] + emi[c], 65504.0f);
  out4[3] = base[3];
}

void pbro_shade_image(const void* positions, const void* normals, const void* albedo,
                      const void* orm, const void* emissive, size_t pitch8, size_t pitch4, int x0,
                      int y0, int w, int h, const float cam_pos[3], const pbro_light* lights,
                      int nlights, const pbro_ibl* ibl, float* out_f32, void* out_f16,
                      size_t out_pitch, int nthreads) {
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads)
  for (int yy = 0; yy < h; ++yy) {
    int y = y0 + yy;
    const uint16_t* pr = (const uint16_t*)((const unsigned char*)positions + (size_t)y * pitch8);
    const uint16_t* nr = (const uint16_t*)((const unsigned char*)normals + (size_t)y * pitch8);
    const uint16_t* ar = (const uint16_t*)((const unsigned char*)albedo + (size_t)y * pitch8);
    const uint16_t* er = (const uint16_t*)((const unsigned char*)emissive + (size_t)y * pitch8);
    const uint8_t* orow = (const uint8_t*)orm + (size_t)y * pitch4;
    for (int xx = 0; xx < w; ++xx) {
      int x = x0 + xx;
      float o[4];
      pbro_shade_pixel(pr + 4 * x, nr + 4 * x, ar + 4 * x, orow + 4 * x, er + 4 * x, cam_pos, lights,
                       nlights, ibl, o);
      if (out_f32) memcpy(out_f32 + ((size_t)yy * w + xx) * 4, o, 16);
      if (out_f16) {
        uint16_t* t = (uint16_t*)((unsigned char*)out_f16 + (size_t)y * out_pitch) + 4 * x;
        for (int c = 0; c < 4; ++c) t[c] = pbro_float_to_half(o[c]);
      }
    }
  }
}
