// f32x2.cuh -- packed pairs of fp32 values on sm_100a (FADD2 / FMUL2 / FFMA2: two IEEE fp32 operations
// per issued instruction, each lane rounded exactly like the scalar instruction).  The FMA pipe does not
// run them faster (scratch/ubench.cu: FFMA2 = 1.96 warp-inst/clk/SM vs FFMA 3.43), but a kernel that is
// bound by ISSUE SLOTS rather than by the pipe -- K3 -- halves the slots its arithmetic takes by giving
// every thread two pixels and keeping pixel A in the low and pixel B in the high lane.
#pragma once
#include <cuda_runtime.h>

namespace pbrcu {

struct f2 { unsigned long long v; };

__device__ __forceinline__ f2 pk(float lo, float hi) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ f2 dup(float a) { return pk(a, a); }
__device__ __forceinline__ float lo(f2 a) {
  float x, y;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a.v));
  return x;
}
__device__ __forceinline__ float hi(f2 a) {
  float x, y;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a.v));
  return y;
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
  f2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ f2 sub2(f2 a, f2 b) {
  f2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
  f2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
  f2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
  return r;
}
// per-lane scalar operations (MUFU, min/max, saturate and selects have no packed form)
template <typename F> __device__ __forceinline__ f2 map2(f2 a, F f) { return pk(f(lo(a)), f(hi(a))); }
__device__ __forceinline__ f2 max2(f2 a, float m) { return pk(fmaxf(lo(a), m), fmaxf(hi(a), m)); }
__device__ __forceinline__ f2 sat2(f2 a) { return pk(__saturatef(lo(a)), __saturatef(hi(a))); }

}  // namespace pbrcu
