// Issue-slot microbenchmarks for K3's instruction mix: what does a packed FFMA2 cost next to integer / MUFU / scalar FFMA work?
// nvcc -arch=sm_100a -O3 -o scratch/ubench_issue scratch/ubench_issue.cu
#include <cuda_runtime.h>
#include <cstdio>
__device__ __forceinline__ unsigned long long ffma2(unsigned long long x, unsigned long long a, unsigned long long b) {
  unsigned long long r; asm volatile("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(r) : "l"(x), "l"(a), "l"(b)); return r;
}
__device__ __forceinline__ float ffma(float x, float a, float b) { float r; asm volatile("fma.rn.f32 %0,%1,%2,%3;" : "=f"(r) : "f"(x), "f"(a), "f"(b)); return r; }
__device__ __forceinline__ unsigned lop(unsigned x, unsigned a) { unsigned r; asm volatile("shf.r.wrap.b32 %0,%1,%1,%2;" : "=r"(r) : "r"(x), "r"(a)); return r; }
__device__ __forceinline__ float mufu(float x) { float r; asm volatile("rsqrt.approx.ftz.f32 %0,%1;" : "=f"(r) : "f"(x)); return r; }
// NP packed FFMA2, NS scalar FFMA, NI integer ALU, NM MUFU per iteration, each on its own independent chain
template <int NP, int NS, int NI, int NM>
__global__ void k_mix(float* out, int iters, float a, float b, unsigned sh) {
  float2 av = make_float2(a, a), bv = make_float2(b, b);
  unsigned long long A = *reinterpret_cast<unsigned long long*>(&av), B = *reinterpret_cast<unsigned long long*>(&bv);
  unsigned long long xp[NP + 1]; float xs[NS + 1]; unsigned xi[NI + 1]; float xm[NM + 1];
  for (int j = 0; j <= NP; ++j) { float2 v = make_float2(threadIdx.x + j, threadIdx.x - j); xp[j] = *reinterpret_cast<unsigned long long*>(&v); }
  for (int j = 0; j <= NS; ++j) xs[j] = threadIdx.x + j;
  for (int j = 0; j <= NI; ++j) xi[j] = threadIdx.x * 7 + j;
  for (int j = 0; j <= NM; ++j) xm[j] = threadIdx.x + j + 1.5f;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (j < NP) xp[j] = ffma2(xp[j], A, B);
      if (j < NI) xi[j] = lop(xi[j], sh);
      if (j < NS) xs[j] = ffma(xs[j], a, b);
      if (j < NM) xm[j] = mufu(xm[j]);
    }
  }
  float s = 0;
  for (int j = 0; j < NP; ++j) { float2 v = *reinterpret_cast<float2*>(&xp[j]); s += v.x + v.y; }
  for (int j = 0; j < NS; ++j) s += xs[j];
  for (int j = 0; j < NI; ++j) s += float(xi[j]);
  for (int j = 0; j < NM; ++j) s += xm[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NP, int NS, int NI, int NM> static void run(const char* name, float* out, int sms, int khz) {
  const int blocks = sms * 8, threads = 256, iters = 4096;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k_mix<NP, NS, NI, NM><<<blocks, threads>>>(out, iters, 1.0001f, 0.5f, 7u); cudaDeviceSynchronize();
  cudaEventRecord(a); k_mix<NP, NS, NI, NM><<<blocks, threads>>>(out, iters, 1.0001f, 0.5f, 7u); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double inst = double(blocks) * threads / 32 * iters * (NP + NS + NI + NM);
  const double clk = ms * 1e-3 * sms * double(khz) * 1e3;
  printf("%-34s FFMA2 %d FFMA %d ALU %d MUFU %d: %.2f warp-inst/clk/SM total (FFMA2 %.2f FFMA %.2f ALU %.2f MUFU %.2f)\n", name, NP, NS, NI, NM, inst / clk,
         inst / clk * NP / (NP + NS + NI + NM), inst / clk * NS / (NP + NS + NI + NM), inst / clk * NI / (NP + NS + NI + NM), inst / clk * NM / (NP + NS + NI + NM));
}
int main() {
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
  const int sms = prop.multiProcessorCount; int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  float* out; cudaMalloc(&out, size_t(sms) * 8 * 256 * 4);
  run<8, 0, 0, 0>("packed only", out, sms, khz);
  run<0, 8, 0, 0>("scalar FFMA only", out, sms, khz);
  run<0, 0, 8, 0>("integer ALU only", out, sms, khz);
  run<0, 0, 0, 8>("MUFU only", out, sms, khz);
  run<4, 0, 4, 0>("packed + ALU 1:1", out, sms, khz);
  run<4, 0, 8, 0>("packed + ALU 1:2", out, sms, khz);
  run<0, 4, 4, 0>("FFMA + ALU 1:1", out, sms, khz);
  run<4, 4, 0, 0>("packed + FFMA 1:1", out, sms, khz);
  run<4, 4, 4, 0>("packed + FFMA + ALU", out, sms, khz);
  run<4, 0, 4, 1>("packed + ALU + MUFU 4:4:1", out, sms, khz);
  run<3, 3, 6, 1>("K3-like 3:3:6:1", out, sms, khz);
  run<2, 2, 3, 1>("2:2:3:1", out, sms, khz);
  run<2, 0, 6, 0>("packed + ALU 1:3", out, sms, khz);
  run<0, 2, 6, 0>("FFMA + ALU 1:3", out, sms, khz);
  run<0, 0, 8, 1>("ALU + MUFU 8:1", out, sms, khz);
  run<0, 7, 0, 1>("FFMA + MUFU 7:1", out, sms, khz);
  run<4, 0, 0, 1>("packed + MUFU 4:1", out, sms, khz);
  run<0, 6, 6, 1>("scalar 6:6:1", out, sms, khz);
  return 0;
}
