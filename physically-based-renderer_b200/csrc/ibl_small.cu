// ibl_small.cu -- K9 Hammersley table, K8 split-sum BRDF LUT (both bit-exact vs the oracle),
// K4 equirect -> cube, K5 cube mip generation (bit-exact).  No reference counterpart.
#include "cube.cuh"
#include "detmath.cuh"

namespace pbrcu {

// ---- K9 ---------------------------------------------------------------------------------------
__global__ void hammersley_kernel(uint32_t n, float2* out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = dm::hammersley(i, n);
}

// ---- K8: Karis 2013 IntegrateBRDF.  One thread per texel, sequential sum over the sample index:
// the summation order IS the contract (bit-exact with oracle/pbr_oracle_ibl.c:pbro_brdf_lut_texel).
// The 1024 GGX half-vectors depend on the row's roughness only, so each block (one LUT row segment)
// first builds them in shared memory once; the per-texel loop is then ~25 dependent fp32 ops.
constexpr int kLutThreads = 128;
__global__ void __launch_bounds__(kLutThreads) brdf_lut_kernel(uint32_t res, uint32_t nsamples, float2* out) {
  extern __shared__ float smem[];  // nsamples * (Hx, Hz)
  float2* H = reinterpret_cast<float2*>(smem);
  const uint32_t j = blockIdx.y;
  const float rough = dm::div(dm::add(float(j), 0.5f), float(res));
  const float alpha = dm::mul(rough, rough);
  for (uint32_t s = threadIdx.x; s < nsamples; s += kLutThreads) {
    const float3 h = dm::importance_sample_ggx(dm::hammersley(s, nsamples), alpha);
    H[s] = make_float2(h.x, h.z);
  }
  __syncthreads();
  const uint32_t i = blockIdx.x * kLutThreads + threadIdx.x;
  if (i >= res) return;
  const float NoV = dm::div(dm::add(float(i), 0.5f), float(res));
  const float k = dm::mul(alpha, 0.5f);
  const float omk = dm::sub(1.0f, k);
  const float Vx = dm::sqrt(dm::sub(1.0f, dm::mul(NoV, NoV))), Vz = NoV;
  const float g1v = dm::div(NoV, dm::add(dm::mul(NoV, omk), k));
  float A = 0.0f, B = 0.0f;
  for (uint32_t s = 0; s < nsamples; ++s) {
    const float2 h = H[s];
    const float VoH = dm::add(dm::mul(Vx, h.x), dm::mul(Vz, h.y));
    const float Lz = dm::sub(dm::mul(dm::mul(2.0f, VoH), h.y), Vz);
    const float NoL = fmaxf(Lz, 0.0f), NoH = fmaxf(h.y, 0.0f), VoHc = fmaxf(VoH, 0.0f);
    if (NoL > 0.0f) {
      const float g1l = dm::div(NoL, dm::add(dm::mul(NoL, omk), k));
      const float G = dm::mul(g1v, g1l);
      const float Gvis = dm::div(dm::mul(G, VoHc), dm::mul(NoH, NoV));
      const float t = dm::sub(1.0f, VoHc), t2 = dm::mul(t, t);
      const float Fc = dm::mul(dm::mul(t2, t2), t);
      A = dm::add(A, dm::mul(dm::sub(1.0f, Fc), Gvis));
      B = dm::add(B, dm::mul(Fc, Gvis));
    }
  }
  const float invn = dm::div(1.0f, float(nsamples));
  out[size_t(j) * res + i] = make_float2(dm::mul(A, invn), dm::mul(B, invn));
}

// ---- K4: lat-long RGBA32F -> cube base level.  u = atan2(z,x)/2pi + 0.5, v = acos(y)/pi, bilinear,
// wrap in u, clamp in v.  Two texels per thread -> one 16-byte store.
__device__ __forceinline__ float4 equirect_fetch(const float4* src, int sw, int sh, float3 n) {
  const float u = atan2f(n.z, n.x) * (0.5f / dm::kPi) + 0.5f;
  const float v = acosf(fminf(fmaxf(n.y, -1.0f), 1.0f)) * (1.0f / dm::kPi);
  const float fx = u * float(sw) - 0.5f, fy = v * float(sh) - 0.5f;
  const float x0f = floorf(fx), y0f = floorf(fy);
  const float tx = fx - x0f, ty = fy - y0f;
  int x0 = int(x0f), y0 = int(y0f), x1 = x0 + 1, y1 = y0 + 1;
  // u is in [0, 1]: x0 in [-1, sw], x1 in [0, sw + 1] -- wrap without a division
  x0 = x0 < 0 ? x0 + sw : (x0 >= sw ? x0 - sw : x0);
  x1 = x1 >= sw ? x1 - sw : x1;
  x1 = x1 >= sw ? x1 - sw : x1;
  y0 = min(max(y0, 0), sh - 1); y1 = min(max(y1, 0), sh - 1);
  const float4 t00 = __ldg(src + size_t(y0) * sw + x0), t10 = __ldg(src + size_t(y0) * sw + x1);
  const float4 t01 = __ldg(src + size_t(y1) * sw + x0), t11 = __ldg(src + size_t(y1) * sw + x1);
  float4 o;
  o.x = fminf(lerpf(lerpf(t00.x, t10.x, tx), lerpf(t01.x, t11.x, tx), ty), 65504.0f);
  o.y = fminf(lerpf(lerpf(t00.y, t10.y, tx), lerpf(t01.y, t11.y, tx), ty), 65504.0f);
  o.z = fminf(lerpf(lerpf(t00.z, t10.z, tx), lerpf(t01.z, t11.z, tx), ty), 65504.0f);
  o.w = fminf(lerpf(lerpf(t00.w, t10.w, tx), lerpf(t01.w, t11.w, tx), ty), 65504.0f);
  return o;
}

// grid = (ceil(res / 32), ceil(res / 8), 6), block = 32 x 8: face, row and column straight from the indices
__global__ void __launch_bounds__(256) equirect_to_cube_kernel(const float4* src, int sw, int sh, int res, uint2* out) {
  const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y, face = blockIdx.z;
  if (x >= res || y >= res) return;
  const float4 c = equirect_fetch(src, sw, sh, normalize3(cube_texel_dir(face, x, y, res)));
  stg_stream8(out + (size_t(face) * res + y) * res + x, make_uint2(f2_to_h2(c.x, c.y), f2_to_h2(c.z, c.w)));
}

// ---- K5: one level from the previous STORED level: ((t00 + t10) + (t01 + t11)) * 0.25, one RNE
// rounding at the store (bit-exact contract: explicit non-contracted adds).
__global__ void __launch_bounds__(256) cube_mip_kernel(const uint2* src, uint2* dst, int rd) {
  const int rs = rd * 2;
  const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y, face = blockIdx.z;  // grid (rd/32, rd/8, 6), block 32 x 8
  if (x < rd && y < rd) {
    const size_t i = (size_t(face) * rd + y) * rd + x;
    const uint2* s0 = src + (size_t(face) * rs + 2 * y) * rs + 2 * x;
    // the two source texels of a row are 16 contiguous, 16-byte aligned bytes
    const uint4 r0 = *reinterpret_cast<const uint4*>(s0);
    const uint4 r1 = *reinterpret_cast<const uint4*>(s0 + rs);
    const float2 a0 = h2_to_f2(r0.x), a1 = h2_to_f2(r0.y), b0 = h2_to_f2(r0.z), b1 = h2_to_f2(r0.w);
    const float2 c0 = h2_to_f2(r1.x), c1 = h2_to_f2(r1.y), d0 = h2_to_f2(r1.z), d1 = h2_to_f2(r1.w);
    const float ox = __fmul_rn(__fadd_rn(__fadd_rn(a0.x, b0.x), __fadd_rn(c0.x, d0.x)), 0.25f);
    const float oy = __fmul_rn(__fadd_rn(__fadd_rn(a0.y, b0.y), __fadd_rn(c0.y, d0.y)), 0.25f);
    const float oz = __fmul_rn(__fadd_rn(__fadd_rn(a1.x, b1.x), __fadd_rn(c1.x, d1.x)), 0.25f);
    const float ow = __fmul_rn(__fadd_rn(__fadd_rn(a1.y, b1.y), __fadd_rn(c1.y, d1.y)), 0.25f);
    dst[i] = make_uint2(f2_to_h2(ox, oy), f2_to_h2(oz, ow));
  }
}


}  // namespace pbrcu

using namespace pbrcu;

extern "C" size_t pbrcu_cube_mip_offset(uint32_t base_res, uint32_t mip) { return cube_mip_offset(base_res, mip); }
extern "C" size_t pbrcu_cube_chain_texels(uint32_t base_res, uint32_t nmips) { return cube_mip_offset(base_res, nmips); }

extern "C" int pbrcu_ibl_hammersley(pbrcu_ctx* ctx, uint32_t n, float* out_xy_dev, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && out_xy_dev, "NULL argument");
  if (n == 0) return PBRCU_OK;
  hammersley_kernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(s)>>>(n, reinterpret_cast<float2*>(out_xy_dev));
  return launch_check("hammersley_kernel");
}

extern "C" int pbrcu_ibl_brdf_lut(pbrcu_ctx* ctx, uint32_t res, uint32_t nsamples, float* out_rg_dev, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && out_rg_dev, "NULL argument");
  PBRCU_REQUIRE(res > 0 && res <= 65535 && nsamples > 0, "res or nsamples out of range");
  const size_t smem = size_t(nsamples) * sizeof(float2);
  PBRCU_REQUIRE(smem <= size_t(ctx->max_smem_optin), "nsamples too large for shared memory");
  if (smem > 48 * 1024) PBRCU_CUDA(cudaFuncSetAttribute(brdf_lut_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  dim3 grid((res + kLutThreads - 1) / kLutThreads, res);
  brdf_lut_kernel<<<grid, kLutThreads, smem, static_cast<cudaStream_t>(s)>>>(res, nsamples, reinterpret_cast<float2*>(out_rg_dev));
  return launch_check("brdf_lut_kernel");
}

extern "C" int pbrcu_ibl_equirect_to_cube(pbrcu_ctx* ctx, const float* src, uint32_t sw, uint32_t sh, uint32_t res,
                                          void* out_base_dev, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && src && out_base_dev, "NULL argument");
  PBRCU_REQUIRE(sw > 0 && sh > 0 && res > 0, "empty extent");
  PBRCU_REQUIRE(res <= 65535u * 8u, "res too large");
  equirect_to_cube_kernel<<<dim3((res + 31) / 32, (res + 7) / 8, 6), dim3(32, 8), 0, static_cast<cudaStream_t>(s)>>>(
      reinterpret_cast<const float4*>(src), int(sw), int(sh), int(res), static_cast<uint2*>(out_base_dev));
  return launch_check("equirect_to_cube_kernel");
}

extern "C" int pbrcu_ibl_cube_mipgen(pbrcu_ctx* ctx, void* chain_dev, uint32_t base_res, uint32_t nmips, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && chain_dev, "NULL argument");
  PBRCU_REQUIRE(base_res > 0 && (base_res & (base_res - 1)) == 0, "base_res must be a power of two");
  PBRCU_REQUIRE(nmips >= 1 && nmips <= 16 && (base_res >> (nmips - 1)) >= 1, "nmips out of range for base_res");
  uint2* chain = static_cast<uint2*>(chain_dev);
  for (uint32_t m = 1; m < nmips; ++m) {
    const int rd = int(base_res >> m);
    cube_mip_kernel<<<dim3(unsigned((rd + 31) / 32), unsigned((rd + 7) / 8), 6), dim3(32, 8), 0, static_cast<cudaStream_t>(s)>>>(
        chain + cube_mip_offset(base_res, m - 1), chain + cube_mip_offset(base_res, m), rd);
    if (int rc = launch_check("cube_mip_kernel")) return rc;
  }
  return PBRCU_OK;
}
