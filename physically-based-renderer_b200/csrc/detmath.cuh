// detmath.cuh -- deterministic fp32 device math for the bit-exact kernels (K8 BRDF LUT, K9 Hammersley).
// Every step is ONE correctly rounded IEEE-754 operation issued through an intrinsic the compiler
// never contracts into an FMA (__fmul_rn / __fadd_rn / __fsub_rn / __fdiv_rn / __fsqrt_rn), in the
// order DESIGN.md "Deterministic math" specifies, so host and device agree bit-for-bit.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pbrcu { namespace dm {

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float sqrt(float a) { return __fsqrt_rn(a); }

constexpr float kPi = 3.14159265358979323846f;
constexpr float kHalfPi = 1.57079632679489661923f;

// (sin, cos)(2 pi t): quarter-turn reduction to [-pi/4, pi/4], Taylor/Horner, quadrant swap.
__device__ __forceinline__ void sincos2pi(float t, float& s_out, float& c_out) {
  const float x = mul(t, 4.0f);
  const float kf = floorf(add(x, 0.5f));
  const float r = sub(x, kf);
  const float y = mul(r, kHalfPi);
  const float y2 = mul(y, y);
  float ps = add(-1.0f / 5040.0f, mul(y2, 1.0f / 362880.0f));
  ps = add(1.0f / 120.0f, mul(y2, ps));
  ps = add(-1.0f / 6.0f, mul(y2, ps));
  ps = add(1.0f, mul(y2, ps));
  const float s = mul(y, ps);
  float pc = add(1.0f / 40320.0f, mul(y2, -1.0f / 3628800.0f));
  pc = add(-1.0f / 720.0f, mul(y2, pc));
  pc = add(1.0f / 24.0f, mul(y2, pc));
  pc = add(-0.5f, mul(y2, pc));
  const float c = add(1.0f, mul(y2, pc));
  const int k = int(kf) & 3;
  s_out = (k == 0) ? s : (k == 1) ? c : (k == 2) ? -s : -c;
  c_out = (k == 0) ? c : (k == 1) ? -s : (k == 2) ? -c : s;
}

// Hammersley point i of n: (i/n, bit-reversed i * 2^-32)
__device__ __forceinline__ float2 hammersley(uint32_t i, uint32_t n) {
  return make_float2(div(float(i), float(n)), mul(float(__brev(i)), 2.3283064365386963e-10f));
}

// GGX half-vector in the local frame (N = +Z)
__device__ __forceinline__ float3 importance_sample_ggx(float2 xi, float alpha) {
  float sp, cp;
  sincos2pi(xi.x, sp, cp);
  const float a2m1 = sub(mul(alpha, alpha), 1.0f);
  const float cos2 = div(sub(1.0f, xi.y), add(1.0f, mul(a2m1, xi.y)));
  const float cosT = sqrt(cos2);
  const float sinT = sqrt(sub(1.0f, cos2));
  return make_float3(mul(sinT, cp), mul(sinT, sp), cosT);
}

}}  // namespace pbrcu::dm
