/* TEST INFRASTRUCTURE ONLY -- see pbr_oracle.h.  Tier A: plain-C restatement of the reference GLSL.
 * PARITY PINNED against oracle/_ref (the reference's shaders compiled for the CPU) and tests/golden/.
 * Build: gcc -O3 -ffp-contract=off (no FMA contraction; every operation is a single IEEE fp32 op).
 */
#include "pbr_oracle.h"

#include <math.h>
#include <omp.h>
#include <string.h>

int pbro_max_threads(void) { return omp_get_max_threads(); }

/* ---------------------------------------------------------------------------------------------
 * fp16 <-> fp32, round-to-nearest-even.  Storage formats: GBuffer.hpp:13-16 (R16G16B16A16_SFLOAT
 * positions/normals, R16G16B16A16_UNORM albedo), PbrRenderSystem.hpp:21 (RGBA16F lighting output).
 * ------------------------------------------------------------------------------------------- */
float pbro_half_to_float(uint16_t h) {
  uint32_t s = ((uint32_t)h & 0x8000u) << 16, e = (h >> 10) & 0x1Fu, m = h & 0x3FFu, bits;
  if (e == 0) {
    if (m == 0) {
      bits = s;
    } else {
      int sh = 0;
      while (!(m & 0x400u)) { m <<= 1; ++sh; }
      m &= 0x3FFu;
      bits = s | ((uint32_t)(113 - sh) << 23) | (m << 13);
    }
  } else if (e == 31) {
    bits = s | 0x7F800000u | (m << 13);
  } else {
    bits = s | ((e + 112u) << 23) | (m << 13);
  }
  float f;
  memcpy(&f, &bits, 4);
  return f;
}

uint16_t pbro_float_to_half(float f) {
  uint32_t x;
  memcpy(&x, &f, 4);
  uint32_t s = (x >> 16) & 0x8000u;
  x &= 0x7FFFFFFFu;
  if (x > 0x7F800000u) return (uint16_t)(s | 0x7E00u | ((x >> 13) & 0x3FFu)); /* NaN (quiet) */
  if (x == 0x7F800000u) return (uint16_t)(s | 0x7C00u);
  if (x >= 0x477FF000u) return (uint16_t)(s | 0x7C00u); /* >= 65520 rounds to inf */
  if (x <= 0x33000000u) return (uint16_t)s;             /* <= 2^-25 rounds to 0 (tie -> even) */
  uint32_t e = x >> 23, m = (x & 0x7FFFFFu) | 0x800000u;
  if (e < 113) { /* result is a half subnormal */
    uint32_t shift = 126 - e, q = m >> shift, rem = m & ((1u << shift) - 1u), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1u))) ++q;
    return (uint16_t)(s | q);
  }
  uint32_t h = ((e - 112u) << 10) | ((m >> 13) & 0x3FFu), rem = m & 0x1FFFu;
  if (rem > 0x1000u || (rem == 0x1000u && (h & 1u))) ++h;
  return (uint16_t)(s | h);
}

void pbro_half_to_float_n(const uint16_t* h, float* f, size_t n) {
  for (size_t i = 0; i < n; ++i) f[i] = pbro_half_to_float(h[i]);
}
void pbro_float_to_half_n(const float* f, uint16_t* h, size_t n) {
  for (size_t i = 0; i < n; ++i) h[i] = pbro_float_to_half(f[i]);
}

/* ---------------------------------------------------------------------------------------------
 * GLSL built-ins with the semantics SURVEY.md 8(a) pins.
 * ------------------------------------------------------------------------------------------- */
static inline float g_max(float a, float b) { return fmaxf(a, b); } /* maxNum: max(NaN,0)=0 */
static inline float g_min(float a, float b) { return fminf(a, b); }
static inline float g_dot3(const float a[3], const float b[3]) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}
static inline void g_normalize3(const float v[3], float o[3]) {
  float inv = 1.0f / sqrtf(g_dot3(v, v)); /* normalize(0) -> 0 * inf = NaN, as on the GPU */
  o[0] = v[0] * inv; o[1] = v[1] * inv; o[2] = v[2] * inv;
}

/* BlinnPhong.lib.glsl:4-10 */
void pbro_blinn_phong(const float frag[4], const float N[3], const float V[3], const float L[3],
                      float out4[4]) {
  float NdotV = g_max(g_dot3(N, V), 0.0f);                /* :5 */
  float VL[3] = {V[0] + L[0], V[1] + L[1], V[2] + L[2]}, H[3];
  g_normalize3(VL, H);                                    /* :6 */
  float NdotH = g_max(g_dot3(N, H), 0.0f);                /* :7 */
  float spec = powf(NdotH, 64.0f);                        /* :8 */
  for (int c = 0; c < 4; ++c) out4[c] = frag[c] * NdotV + spec; /* :9, all four channels */
}

/* lighting.glsl:19-29 for one texel.  position/normal are RGBA16F, albedo RGBA16_UNORM. */
void pbro_lighting_pixel(const uint16_t pos_h[4], const uint16_t nrm_h[4], const uint16_t alb_u[4],
                         const float cam_pos[3], float out4[4]) {
  float P[3], N[3], A[4], d[3], V[3];
  for (int c = 0; c < 3; ++c) {
    P[c] = pbro_half_to_float(pos_h[c]);                  /* :20 */
    N[c] = pbro_half_to_float(nrm_h[c]);                  /* :21 */
  }
  for (int c = 0; c < 4; ++c) A[c] = (float)alb_u[c] / 65535.0f; /* :22, UNORM16 decode */
  /* :23 reads depth into a dead variable -- not restated */
  for (int c = 0; c < 3; ++c) d[c] = cam_pos[c] - P[c];
  g_normalize3(d, V);                                     /* :25 */
  pbro_blinn_phong(A, N, V, V, out4);                     /* :26 L = V, :28 */
}

void pbro_lighting_image(const void* positions, const void* normals, const void* albedo,
                         size_t pitch8, int x0, int y0, int w, int h, const float cam_pos[3],
                         float* out_f32, void* out_f16, size_t out_pitch, int nthreads) {
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(nthreads)
  for (int yy = 0; yy < h; ++yy) {
    int y = y0 + yy;
    const uint16_t* pr = (const uint16_t*)((const unsigned char*)positions + (size_t)y * pitch8);
    const uint16_t* nr = (const uint16_t*)((const unsigned char*)normals + (size_t)y * pitch8);
    const uint16_t* ar = (const uint16_t*)((const unsigned char*)albedo + (size_t)y * pitch8);
    for (int xx = 0; xx < w; ++xx) {
      int x = x0 + xx;
      float o[4];
      pbro_lighting_pixel(pr + 4 * x, nr + 4 * x, ar + 4 * x, cam_pos, o);
      if (out_f32) memcpy(out_f32 + ((size_t)yy * w + xx) * 4, o, 16);
      if (out_f16) {
        uint16_t* t = (uint16_t*)((unsigned char*)out_f16 + (size_t)y * out_pitch) + 4 * x;
        for (int c = 0; c < 4; ++c) t[c] = pbro_float_to_half(o[c]); /* one rounding, at the store */
      }
    }
  }
}

/* ---------------------------------------------------------------------------------------------
 * aces+gamma.glsl.  The nine literals of each mat3x3(...) fill COLUMNS (:10-19), so
 * (M*v).x = m[0]*v.x + m[3]*v.y + m[6]*v.z -- the transpose of the usual Hill fit, as written.
 * ------------------------------------------------------------------------------------------- */
static const float ACES_INPUT[9] = {0.59719f, 0.35458f, 0.04823f, 0.07600f, 0.90834f,
                                    0.01566f, 0.02840f, 0.13383f, 0.83777f};
static const float ACES_OUTPUT[9] = {1.60475f, -0.53108f, -0.07367f, -0.10208f, 1.10813f,
                                     -0.00605f, -0.00327f, -0.07276f, 1.07602f};
static inline void mat3_mul(const float m[9], const float v[3], float o[3]) {
  for (int r = 0; r < 3; ++r) o[r] = m[r] * v[0] + m[3 + r] * v[1] + m[6 + r] * v[2];
}
/* :20-24 */
static inline float rrt_odt_fit(float c) {
  float num = c * (c + 0.0245786f) - 0.000090537f;
  float den = c * (0.983729f * c + 0.4329510f) + 0.238081f;
  return num / den;
}
/* :26-28 */
void pbro_aces_tonemap(const float rgb[3], float out3[3]) {
  float a[3], f[3];
  mat3_mul(ACES_INPUT, rgb, a);
  for (int c = 0; c < 3; ++c) f[c] = rrt_odt_fit(a[c]);
  mat3_mul(ACES_OUTPUT, f, out3);
}
/* Gamma.lib.glsl:4-6.  pow(x<0, y) is undefined in GLSL and every dark pixel reaches it; pinned as
 * pow(max(x,0), 1/gamma), which equals the reference wherever the reference is defined and equals
 * the reference-after-clamp (NaN -> 0) elsewhere. */
void pbro_gamma_correct(const float rgb[3], float gamma, float out3[3]) {
  float e = 1.0f / gamma;
  for (int c = 0; c < 3; ++c) out3[c] = powf(g_max(rgb[c], 0.0f), e);
}
static inline uint8_t to_unorm8(float v) {
  float c = g_min(g_max(v, 0.0f), 1.0f); /* :38 clamp; NaN -> 0 */
  return (uint8_t)lrintf(c * 255.0f);    /* RNE */
}
/* :30-40 for one texel */
void pbro_tonemap_pixel(const float hdr[4], uint8_t out4[4]) {
  float ldr[3], g[3];
  pbro_aces_tonemap(hdr, ldr);            /* :36 */
  pbro_gamma_correct(ldr, 2.2f, g);       /* :37 */
  out4[0] = to_unorm8(g[0]); out4[1] = to_unorm8(g[1]); out4[2] = to_unorm8(g[2]);
  out4[3] = to_unorm8(hdr[3]);            /* :38 alpha passes through the clamp only */
}

void pbro_tonemap_image(const void* hdr_rgba16f, size_t hdr_pitch, int x0, int y0, int w, int h,
                        void* out_rgba8, size_t out_pitch, int nthreads) {
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(nthreads)
  for (int yy = 0; yy < h; ++yy) {
    int y = y0 + yy;
    const uint16_t* hr = (const uint16_t*)((const unsigned char*)hdr_rgba16f + (size_t)y * hdr_pitch);
    uint8_t* orow = (uint8_t*)out_rgba8 + (size_t)y * out_pitch;
    for (int xx = 0; xx < w; ++xx) {
      int x = x0 + xx;
      float c[4];
      for (int k = 0; k < 4; ++k) c[k] = pbro_half_to_float(hr[4 * x + k]); /* :35 imageLoad */
      pbro_tonemap_pixel(c, orow + 4 * x);                                   /* :39 imageStore */
    }
  }
}
