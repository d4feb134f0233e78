/* TEST INFRASTRUCTURE ONLY -- deterministic fp32 math for the bit-exact Tier-B functions (BRDF LUT,
 * Hammersley).  This is the ORACLE's own statement of the operation order; the device has its own,
 * independently written copy (physically-based-renderer_b200/csrc/detmath.cuh).  The two agree
 * bit-for-bit because every step is one correctly-rounded IEEE-754 fp32 operation (+, -, *, /, sqrt)
 * in a fixed order: this file must be compiled with -ffp-contract=off, the device copy uses
 * __fadd_rn/__fmul_rn/__fdiv_rn/__fsqrt_rn, which the compiler never fuses.  No libm sin/cos/pow.
 *
 * Spec (also in DESIGN.md "Deterministic math"):
 *   sincos2pi(t), t in [0,1):  x = t*4;  k = floor(x + 0.5);  r = x - k  (in [-0.5, 0.5]);
 *     y = r * (pi/2);  y2 = y*y;
 *     s = y * (1 + y2*(S3 + y2*(S5 + y2*(S7 + y2*S9))))          Taylor, Horner, innermost first
 *     c = 1 + y2*(C2 + y2*(C4 + y2*(C6 + y2*(C8 + y2*C10))))
 *     quadrant k&3: 0 -> (s, c); 1 -> (c, -s); 2 -> (-s, -c); 3 -> (-c, s)   as (sin, cos)
 *   radical inverse: bit-reverse the 32-bit index, convert to fp32 (RNE), multiply by 2^-32.
 */
#ifndef PBR_DETMATH_H
#define PBR_DETMATH_H
#include <math.h>
#include <stdint.h>

#define DM_PI 3.14159265358979323846f
#define DM_HALF_PI 1.57079632679489661923f

static inline void dm_sincos2pi(float t, float* s_out, float* c_out) {
  const float S3 = -1.0f / 6.0f, S5 = 1.0f / 120.0f, S7 = -1.0f / 5040.0f, S9 = 1.0f / 362880.0f;
  const float C2 = -0.5f, C4 = 1.0f / 24.0f, C6 = -1.0f / 720.0f, C8 = 1.0f / 40320.0f,
              C10 = -1.0f / 3628800.0f;
  float x = t * 4.0f;
  float kf = floorf(x + 0.5f);
  float r = x - kf;
  float y = r * DM_HALF_PI;
  float y2 = y * y;
  float ps = S7 + y2 * S9;
  ps = S5 + y2 * ps;
  ps = S3 + y2 * ps;
  ps = 1.0f + y2 * ps;
  float s = y * ps;
  float pc = C8 + y2 * C10;
  pc = C6 + y2 * pc;
  pc = C4 + y2 * pc;
  pc = C2 + y2 * pc;
  float c = 1.0f + y2 * pc;
  int k = ((int)kf) & 3;
  switch (k) {
    case 0: *s_out = s; *c_out = c; break;
    case 1: *s_out = c; *c_out = -s; break;
    case 2: *s_out = -s; *c_out = -c; break;
    default: *s_out = -c; *c_out = s; break;
  }
}

static inline uint32_t dm_reverse_bits32(uint32_t b) {
  b = (b << 16) | (b >> 16);
  b = ((b & 0x55555555u) << 1) | ((b & 0xAAAAAAAAu) >> 1);
  b = ((b & 0x33333333u) << 2) | ((b & 0xCCCCCCCCu) >> 2);
  b = ((b & 0x0F0F0F0Fu) << 4) | ((b & 0xF0F0F0F0u) >> 4);
  b = ((b & 0x00FF00FFu) << 8) | ((b & 0xFF00FF00u) >> 8);
  return b;
}

/* xi = (i/n, radicalInverse_VdC(i)) */
static inline void dm_hammersley(uint32_t i, uint32_t n, float* x0, float* x1) {
  *x0 = (float)i / (float)n;
  *x1 = (float)dm_reverse_bits32(i) * 2.3283064365386963e-10f; /* 2^-32 */
}

/* GGX importance sample in the local frame (N = +Z): H = (sinT cosP, sinT sinP, cosT) */
static inline void dm_importance_sample_ggx(float xi0, float xi1, float alpha, float H[3]) {
  float sp, cp;
  dm_sincos2pi(xi0, &sp, &cp);
  float a2m1 = alpha * alpha - 1.0f;
  float cos2 = (1.0f - xi1) / (1.0f + a2m1 * xi1);
  float cosT = sqrtf(cos2);
  float sinT = sqrtf(1.0f - cos2);
  H[0] = sinT * cp;
  H[1] = sinT * sp;
  H[2] = cosT;
}
#endif
