// gather_probe.cu -- prints what tld4 returns on this device: component order, footprint selection at integer coordinates
// (unnormalised), clamp behaviour at the array edge, on an RGBA16F array created with cudaArrayTextureGather.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
__global__ void probe(cudaTextureObject_t t, float x, float y, float4* out) {
  out[0] = tex2Dgather<float4>(t, x, y, 0);
  out[1] = tex2Dgather<float4>(t, x, y, 1);
  out[2] = tex2Dgather<float4>(t, x, y, 2);
}
int main() {
  const int W = 8, H = 6;
  std::vector<__half> h(W * H * 4);
  for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) for (int c = 0; c < 4; ++c) h[(y * W + x) * 4 + c] = __float2half(float(100 * c + 10 * y + x));
  cudaChannelFormatDesc fmt = cudaCreateChannelDescHalf4();
  cudaArray_t arr;
  cudaError_t e = cudaMallocArray(&arr, &fmt, W, H, cudaArrayTextureGather);
  printf("cudaMallocArray(gather): %s\n", cudaGetErrorString(e));
  cudaMemcpy2DToArray(arr, 0, 0, h.data(), W * 8, W * 8, H, cudaMemcpyHostToDevice);
  cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
  cudaTextureDesc td{}; td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp; td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
  cudaTextureObject_t t; e = cudaCreateTextureObject(&t, &rd, &td, nullptr);
  printf("cudaCreateTextureObject: %s\n", cudaGetErrorString(e));
  float4* out; cudaMallocManaged(&out, 3 * sizeof(float4));
  const float pts[][2] = {{3.0f, 2.0f}, {3.5f, 2.5f}, {2.999f, 2.0f}, {3.001f, 2.0f}, {0.0f, 0.0f}, {8.0f, 6.0f}, {1.0f, 1.0f}};
  for (auto& p : pts) {
    probe<<<1, 1>>>(t, p[0], p[1], out);
    e = cudaDeviceSynchronize();
    printf("gather(%.3f, %.3f): r = (%.0f %.0f %.0f %.0f)  g.x = %.0f  b.x = %.0f  [%s]   texel value = 100 c + 10 y + x\n", p[0], p[1], out[0].x, out[0].y, out[0].z, out[0].w,
           out[1].x, out[2].x, cudaGetErrorString(e));
  }
  cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
  printf("maxTexture2DGather = %d x %d, maxTexture2D = %d x %d\n", pr.maxTexture2DGather[0], pr.maxTexture2DGather[1], pr.maxTexture2D[0], pr.maxTexture2D[1]);
  return 0;
}
