// ibl_conv.cu -- K7 GGX-prefiltered specular chain and K6 cosine irradiance: Monte-Carlo convolution
// of a cube-map mip chain over a fixed Hammersley sample set (filtered importance sampling).
// No reference counterpart; definitions: oracle/pbr_oracle_ibl.c (Karis 2013, Krivanek &
// Colbert 2008).
//
// Under N = V = R the sample set in the local frame depends only on (roughness, i), so each block
// builds the table {L_local.xyz, source lod} once in shared memory; per sample a thread then does a
// 3x3 frame rotation, one exact-weight trilinear cube fetch and 4 FMAs.  G lanes share one texel
// (samples strided by G, warp-shuffle reduction) so the small mips still fill the machine.
#include "cube.cuh"
#include "detmath.cuh"

namespace pbrcu {

enum : int { CONV_GGX = 0, CONV_COSINE = 1 };
constexpr int kConvThreads = 256;

struct ConvParams {
  CubeChain src;
  uint2* dst_level;   // first texel of the destination level
  int res;            // destination level resolution
  int nsamples;
  float roughness;    // CONV_GGX only
  int row_begin, row_end;  // global rows (face*res + y) handled by this launch (multi-GPU shard)
  int tw, th;         // texel tile per block: tw*th*G == kConvThreads
};

template <int MODE>
__device__ __forceinline__ float4 table_entry(const ConvParams& p, uint32_t i) {
  const float2 xi = dm::hammersley(i, uint32_t(p.nsamples));
  const float omegaP = (4.0f * dm::kPi) / (6.0f * float(p.src.base_res) * float(p.src.base_res));
  float Lx, Ly, Lz, omegaS;
  if (MODE == CONV_GGX) {
    const float alpha = p.roughness * p.roughness, a2 = alpha * alpha;
    const float3 H = dm::importance_sample_ggx(xi, alpha);
    const float c = H.z;
    Lx = 2.0f * c * H.x; Ly = 2.0f * c * H.y; Lz = 2.0f * c * c - 1.0f;
    const float dd = c * c * (a2 - 1.0f) + 1.0f;
    const float pdf = a2 / (dm::kPi * dd * dd) * 0.25f;  // D NoH / (4 VoH) with NoH = VoH
    omegaS = 1.0f / (float(p.nsamples) * pdf);
  } else {
    float sp, cp;
    dm::sincos2pi(xi.x, sp, cp);
    const float r = sqrtf(xi.y);
    Lz = sqrtf(1.0f - xi.y);
    Lx = r * cp; Ly = r * sp;
    omegaS = dm::kPi / (float(p.nsamples) * Lz);  // pdf = cos / pi
  }
  float lod = 0.5f * log2f(omegaS / omegaP);
  lod = fminf(fmaxf(lod, 0.0f), float(p.src.nmips - 1));
  return make_float4(Lx, Ly, Lz, lod);
}

template <int G, int MODE>
__global__ void __launch_bounds__(kConvThreads) conv_kernel(const __grid_constant__ ConvParams p) {
  extern __shared__ float4 table[];
  for (int s = threadIdx.x; s < p.nsamples; s += kConvThreads) table[s] = table_entry<MODE>(p, uint32_t(s));
  __syncthreads();

  const int texel = threadIdx.x / G, lane = threadIdx.x % G;
  const int x = blockIdx.x * p.tw + texel % p.tw;
  const int grow = p.row_begin + blockIdx.y * p.th + texel / p.tw;
  const bool valid = x < p.res && grow < p.row_end;
  const int xc = min(x, p.res - 1), gc = min(grow, p.row_end - 1);
  const int face = gc / p.res, y = gc - face * p.res;

  // tangent frame of Karis' ImportanceSampleGGX: up = |N.z| < 0.999 ? +Z : +X
  const float3 N = normalize3(cube_texel_dir(face, xc, y, p.res));
  const bool upz = fabsf(N.z) < 0.999f;
  const float3 up = upz ? make_float3(0.f, 0.f, 1.f) : make_float3(1.f, 0.f, 0.f);
  const float3 T = normalize3(make_float3(up.y * N.z - up.z * N.y, up.z * N.x - up.x * N.z, up.x * N.y - up.y * N.x));
  const float3 B = make_float3(N.y * T.z - N.z * T.y, N.z * T.x - N.x * T.z, N.x * T.y - N.y * T.x);

  float ar = 0.f, ag = 0.f, ab = 0.f, wsum = 0.f;
  for (int s = lane; s < p.nsamples; s += G) {
    const float4 e = table[s];
    if (e.z > 0.0f) {  // Karis: `if (NoL > 0)`; G == 1: uniform across the warp
      const float3 L = make_float3(T.x * e.x + B.x * e.y + N.x * e.z, T.y * e.x + B.y * e.y + N.y * e.z,
                                   T.z * e.x + B.z * e.y + N.z * e.z);
      const RGB c = cube_sample(p.src, L, e.w);
      const float w = (MODE == CONV_GGX) ? e.z : 1.0f;
      ar = fmaf(c.r, w, ar); ag = fmaf(c.g, w, ag); ab = fmaf(c.b, w, ab);
      wsum += w;
    }
  }
#pragma unroll
  for (int off = G / 2; off > 0; off >>= 1) {  // warp-shuffle reduction over the texel's G lanes
    ar += __shfl_xor_sync(0xffffffffu, ar, off);
    ag += __shfl_xor_sync(0xffffffffu, ag, off);
    ab += __shfl_xor_sync(0xffffffffu, ab, off);
    wsum += __shfl_xor_sync(0xffffffffu, wsum, off);
  }
  if (valid && lane == 0) {
    const float inv = 1.0f / wsum;
    const size_t idx = (size_t(face) * p.res + y) * p.res + x;
    p.dst_level[idx] = make_uint2(f2_to_h2(fminf(ar * inv, 65504.f), fminf(ag * inv, 65504.f)),
                                  f2_to_h2(fminf(ab * inv, 65504.f), 1.0f));
  }
}

template <int MODE>
static int launch_conv(const pbrcu_ctx* ctx, ConvParams p, cudaStream_t s) {
  const int rows = p.row_end - p.row_begin;
  if (rows <= 0) return PBRCU_OK;
  // lanes per texel: enough threads to fill 148 SMs x 1024 threads, at most a warp
  const size_t texels = size_t(rows) * p.res;
  int G = 1;
  while (G < 32 && texels * G < size_t(ctx->sm_count) * 1024) G *= 2;
  const int per_block = kConvThreads / G;
  p.tw = p.res < 32 ? p.res : 32;
  if (p.tw > per_block) p.tw = per_block;
  p.th = per_block / p.tw;
  while (p.tw * p.th * G != kConvThreads) { --p.tw; p.th = per_block / p.tw; }  // res not a power of two
  dim3 grid((p.res + p.tw - 1) / p.tw, (rows + p.th - 1) / p.th);
  const size_t smem = size_t(p.nsamples) * sizeof(float4);
#define PBRCU_CONV_CASE(GG)                                                                          \
  case GG:                                                                                           \
    if (smem > 48 * 1024)                                                                            \
      PBRCU_CUDA(cudaFuncSetAttribute(conv_kernel<GG, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem))); \
    conv_kernel<GG, MODE><<<grid, kConvThreads, smem, s>>>(p);                                       \
    break;
  switch (G) {
    PBRCU_CONV_CASE(1) PBRCU_CONV_CASE(2) PBRCU_CONV_CASE(4) PBRCU_CONV_CASE(8) PBRCU_CONV_CASE(16) PBRCU_CONV_CASE(32)
  }
#undef PBRCU_CONV_CASE
  return launch_check("conv_kernel");
}

static void shard_rows(int total_rows, uint32_t shard, uint32_t nshards, int& b, int& e) {
  b = int(uint64_t(total_rows) * shard / nshards);
  e = int(uint64_t(total_rows) * (shard + 1) / nshards);
}

}  // namespace pbrcu

using namespace pbrcu;

extern "C" int pbrcu_ibl_prefilter_ggx(pbrcu_ctx* ctx, const void* src_chain_dev, uint32_t base_res, uint32_t nmips,
                                       uint32_t nsamples, void* dst_chain_dev, uint32_t first_mip, uint32_t last_mip,
                                       uint32_t shard, uint32_t nshards, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && src_chain_dev && dst_chain_dev, "NULL argument");
  PBRCU_REQUIRE(base_res > 0 && (base_res & (base_res - 1)) == 0, "base_res must be a power of two");
  PBRCU_REQUIRE(nmips >= 2 && nmips <= 16 && (base_res >> (nmips - 1)) >= 1, "nmips out of range for base_res");
  PBRCU_REQUIRE(first_mip <= last_mip && last_mip < nmips, "mip range out of bounds");
  PBRCU_REQUIRE(nsamples > 0 && size_t(nsamples) * 16 <= size_t(ctx->max_smem_optin), "nsamples out of range");
  PBRCU_REQUIRE(nshards >= 1 && shard < nshards, "shard out of range");
  cudaStream_t st = static_cast<cudaStream_t>(s);
  for (uint32_t m = first_mip; m <= last_mip; ++m) {
    const int res = int(base_res >> m);
    int rb, re;
    shard_rows(6 * res, shard, nshards, rb, re);
    const size_t off = cube_mip_offset(base_res, m);
    uint2* dst = static_cast<uint2*>(dst_chain_dev) + off;
    if (m == 0) {  // roughness 0: the GGX lobe is a delta -> the level is an exact copy
      const uint2* src = static_cast<const uint2*>(src_chain_dev) + off;
      PBRCU_CUDA(cudaMemcpyAsync(dst + size_t(rb) * res, src + size_t(rb) * res, size_t(re - rb) * res * 8,
                                 cudaMemcpyDeviceToDevice, st));
      continue;
    }
    ConvParams p{};
    p.src = make_chain(src_chain_dev, base_res, nmips);
    p.dst_level = dst;
    p.res = res;
    p.nsamples = int(nsamples);
    p.roughness = float(m) / float(nmips - 1);
    p.row_begin = rb;
    p.row_end = re;
    if (int rc = launch_conv<CONV_GGX>(ctx, p, st)) return rc;
  }
  return PBRCU_OK;
}

extern "C" int pbrcu_ibl_irradiance(pbrcu_ctx* ctx, const void* src_chain_dev, uint32_t base_res, uint32_t nmips,
                                    uint32_t res, uint32_t nsamples, void* out_dev, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && src_chain_dev && out_dev, "NULL argument");
  PBRCU_REQUIRE(base_res > 0 && nmips >= 1 && nmips <= 16 && (base_res >> (nmips - 1)) >= 1, "bad source chain");
  PBRCU_REQUIRE(res > 0 && nsamples > 0 && size_t(nsamples) * 16 <= size_t(ctx->max_smem_optin), "res or nsamples out of range");
  ConvParams p{};
  p.src = make_chain(src_chain_dev, base_res, nmips);
  p.dst_level = static_cast<uint2*>(out_dev);
  p.res = int(res);
  p.nsamples = int(nsamples);
  p.row_begin = 0;
  p.row_end = 6 * int(res);
  return launch_conv<CONV_COSINE>(ctx, p, static_cast<cudaStream_t>(s));
}
