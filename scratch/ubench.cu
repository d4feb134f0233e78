// Microbenchmarks that decide the K3/K7 design: FFMA vs FFMA2 issue rate, and texture-unit rates for
// RGBA16F layered/cubemap fetches (bilinear, trilinear, tld4 gather, point).  nvcc -arch=sm_100a.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %s:%d\n",cudaGetErrorString(e),__FILE__,__LINE__); return 1;}}while(0)

__global__ void k_ffma(float* out, int iters, float a, float b) {
  float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; ++i) {
    x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
    x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}
__device__ __forceinline__ unsigned long long ffma2(unsigned long long x, unsigned long long a, unsigned long long b) {
  unsigned long long r; asm volatile("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(r) : "l"(x), "l"(a), "l"(b)); return r;
}
__global__ void k_ffma2(float* out, int iters, float a, float b) {
  float2 av = make_float2(a, a), bv = make_float2(b, b);
  unsigned long long A = *reinterpret_cast<unsigned long long*>(&av), B = *reinterpret_cast<unsigned long long*>(&bv);
  unsigned long long x[8];
  for (int j = 0; j < 8; ++j) { float2 v = make_float2(threadIdx.x + j, threadIdx.x - j); x[j] = *reinterpret_cast<unsigned long long*>(&v); }
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = ffma2(x[j], A, B);
  }
  float s = 0;
  for (int j = 0; j < 8; ++j) { float2 v = *reinterpret_cast<float2*>(&x[j]); s += v.x + v.y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// mode 0: bilinear layered level 0; 1: trilinear layered (lod .5); 2: tld4 gather (1 channel); 3: point fetch;
// 4: cubemap trilinear; 5: cubemap bilinear
template <int MODE>
__global__ void k_tex(cudaTextureObject_t t, float4* out, int iters, float res, float jitter) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  // coherent-ish coordinates: a warp covers an 8x4 texel patch that drifts
  float x = float(tid & 7) + float((blockIdx.x * 37) & 1023) + 0.3f, y = float((tid >> 3) & 3) + float((blockIdx.x * 91) & 1023) + 0.7f;
  float4 acc = make_float4(0, 0, 0, 0);
  for (int i = 0; i < iters; ++i) {
    float4 v;
    const float u = x / res, w = y / res;
    if (MODE == 0) v = tex2DLayeredLod<float4>(t, u, w, i & 3, 0.0f);
    if (MODE == 1) v = tex2DLayeredLod<float4>(t, u, w, i & 3, 0.5f);
    if (MODE == 2) {
      const int layer = i & 3;
      asm volatile("tld4.r.a2d.v4.f32.f32 {%0,%1,%2,%3}, [%4, {%5,%6,%7,%7}];"
                   : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(t), "r"(layer), "f"(u), "f"(w));
    }
    if (MODE == 3) v = tex2DLayeredLod<float4>(t, u, w, i & 3, 0.0f);
    if (MODE == 4) v = texCubemapLod<float4>(t, 1.0f, 2.0f * u - 1.0f, 2.0f * w - 1.0f, 0.5f);
    if (MODE == 5) v = texCubemapLod<float4>(t, 1.0f, 2.0f * u - 1.0f, 2.0f * w - 1.0f, 0.0f);
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    x += jitter; y += jitter * 0.5f;
  }
  out[tid] = acc;
}

template <typename F> static float time_ms(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}

static int make_tex(int res, int mips, bool cube, bool linear, cudaTextureObject_t* out) {
  cudaChannelFormatDesc d = cudaCreateChannelDescHalf4();
  cudaMipmappedArray_t arr;
  CK(cudaMallocMipmappedArray(&arr, &d, make_cudaExtent(res, res, 6), mips, cube ? cudaArrayCubemap : cudaArrayLayered));
  std::vector<__half> h(size_t(res) * res * 4 * 6);
  for (size_t i = 0; i < h.size(); ++i) h[i] = __float2half(float((i * 2654435761u) & 0xffff) / 65536.f);
  for (int m = 0; m < mips; ++m) {
    cudaArray_t lvl; CK(cudaGetMipmappedArrayLevel(&lvl, arr, m));
    int r = res >> m;
    cudaMemcpy3DParms p = {};
    p.srcPtr = make_cudaPitchedPtr(h.data(), size_t(r) * 8, r, r);
    p.dstArray = lvl; p.extent = make_cudaExtent(r, r, 6); p.kind = cudaMemcpyHostToDevice;
    CK(cudaMemcpy3D(&p));
  }
  cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeMipmappedArray; rd.res.mipmap.mipmap = arr;
  cudaTextureDesc td = {};
  td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
  td.filterMode = linear ? cudaFilterModeLinear : cudaFilterModePoint;
  td.mipmapFilterMode = linear ? cudaFilterModeLinear : cudaFilterModePoint;
  td.readMode = cudaReadModeElementType; td.normalizedCoords = 1; td.maxMipmapLevelClamp = float(mips - 1);
  CK(cudaCreateTextureObject(out, &rd, &td, nullptr));
  return 0;
}

int main() {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount; int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  printf("device %s, %d SMs, %d MHz\n", prop.name, sms, khz / 1000);
  float* out; CK(cudaMalloc(&out, size_t(sms) * 8 * 1024 * 16));
  const int blocks = sms * 8, threads = 256, iters = 4096;
  {
    float ms = time_ms([&] { k_ffma<<<blocks, threads>>>(out, iters, 1.0001f, 0.5f); });
    double inst = double(blocks) * threads / 32 * iters * 8;
    printf("FFMA : %.3f ms, %.2f warp-inst/clk/SM (at max clock), %.1f TFLOP/s\n", ms, inst / (ms * 1e-3) / (sms * double(khz) * 1e3), inst * 64 / (ms * 1e-3) / 1e12);
    ms = time_ms([&] { k_ffma2<<<blocks, threads>>>(out, iters, 1.0001f, 0.5f); });
    printf("FFMA2: %.3f ms, %.2f warp-inst/clk/SM (at max clock), %.1f TFLOP/s\n", ms, inst / (ms * 1e-3) / (sms * double(khz) * 1e3), inst * 128 / (ms * 1e-3) / 1e12);
  }
  cudaTextureObject_t lay_lin, lay_pt, cube_lin;
  if (make_tex(1024, 11, false, true, &lay_lin) || make_tex(1024, 11, false, false, &lay_pt) || make_tex(1024, 11, true, true, &cube_lin)) return 1;
  const int titers = 512;
  const char* names[] = {"layered bilinear", "layered trilinear", "layered tld4 gather", "layered point", "cubemap trilinear", "cubemap bilinear"};
  for (int jit = 0; jit < 2; ++jit) {
    const float jitter = jit ? 3.7f : 0.01f;
    for (int mode = 0; mode < 6; ++mode) {
      float ms = 0;
      float4* o4 = reinterpret_cast<float4*>(out);
      switch (mode) {
        case 0: ms = time_ms([&] { k_tex<0><<<blocks, threads>>>(lay_lin, o4, titers, 1024.f, jitter); }); break;
        case 1: ms = time_ms([&] { k_tex<1><<<blocks, threads>>>(lay_lin, o4, titers, 1024.f, jitter); }); break;
        case 2: ms = time_ms([&] { k_tex<2><<<blocks, threads>>>(lay_lin, o4, titers, 1024.f, jitter); }); break;
        case 3: ms = time_ms([&] { k_tex<3><<<blocks, threads>>>(lay_pt, o4, titers, 1024.f, jitter); }); break;
        case 4: ms = time_ms([&] { k_tex<4><<<blocks, threads>>>(cube_lin, o4, titers, 1024.f, jitter); }); break;
        case 5: ms = time_ms([&] { k_tex<5><<<blocks, threads>>>(cube_lin, o4, titers, 1024.f, jitter); }); break;
      }
      CK(cudaGetLastError());
      double fetches = double(blocks) * threads * titers;
      printf("TEX %-20s jitter %.2f: %.3f ms, %.3f fetches/clk/SM (at max clock), %.1f Gfetch/s\n", names[mode], jitter, ms,
             fetches / (ms * 1e-3) / (sms * double(khz) * 1e3), fetches / (ms * 1e-3) / 1e9);
    }
  }
  return 0;
}
