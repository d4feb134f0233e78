// ibl_quads.cu -- builds the footprint-record copies of the IBL textures (see ibl_quads.cuh).
// HBM-bound gather: every record reads 4 texels (3 of them L1/L2 hits from neighbouring threads) and
// writes 32 bytes; 2048^2 x 12 mips = 33.5 M records = 1.07 GB written.
#include "ibl_quads.cuh"

#include <new>

namespace pbrcu {

constexpr float kShadeTexRelRange = 0.12f;

// faces of one level: n = 6*res*res records (n_faces*res*res in general)
template <bool CUBE>
// (grid = (ceil(res / 32), ceil(res / 8), faces), block = 32 x 8)
__global__ void __launch_bounds__(256) build_quads8_kernel(const uint2* __restrict__ src, uint4* __restrict__ dst, int res) {
  const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (x < res && y < res) {
    const size_t face = blockIdx.z;
    const size_t i = (face * res + y) * res + x;
    const int x1 = min(x + 1, res - 1), y1 = min(y + 1, res - 1);
    const uint2* f = src + face * size_t(res) * res;
    const uint2 t00 = __ldg(f + size_t(y) * res + x), t10 = __ldg(f + size_t(y) * res + x1);
    const uint2 t01 = __ldg(f + size_t(y1) * res + x), t11 = __ldg(f + size_t(y1) * res + x1);
    if (CUBE) {  // {rg00, rg10, rg01, rg11}, {b00|b10, b01|b11, 0, 0}
      dst[2 * i] = make_uint4(t00.x, t10.x, t01.x, t11.x);
      dst[2 * i + 1] = make_uint4((t00.y & 0xFFFFu) | (t10.y << 16), (t01.y & 0xFFFFu) | (t11.y << 16), 0u, 0u);
    } else {
      dst[2 * i] = make_uint4(t00.x, t00.y, t10.x, t10.y);
      dst[2 * i + 1] = make_uint4(t01.x, t01.y, t11.x, t11.y);
    }
  }
}

static int build_level(const uint2* src, uint4* dst, int res, int faces, int sms, cudaStream_t st) {
  (void)sms;
  const dim3 grid(unsigned((res + 31) / 32), unsigned((res + 7) / 8), unsigned(faces));
  if (faces == 6) build_quads8_kernel<true><<<grid, dim3(32, 8), 0, st>>>(src, dst, res);
  else build_quads8_kernel<false><<<grid, dim3(32, 8), 0, st>>>(src, dst, res);
  return launch_check("build_quads8_kernel");
}

}  // namespace pbrcu

using namespace pbrcu;

extern "C" int pbrcu_ibl_accel_free(pbrcu_ctx* ctx, pbrcu_ibl_accel* a) {
  (void)ctx;
  if (!a) return PBRCU_OK;
  if (a->pref) cudaFree(a->pref);
  if (a->irr) cudaFree(a->irr);
  if (a->lut) cudaFree(a->lut);
  cube_tex_destroy(&a->pref_tex);
  cube_tex_destroy(&a->irr_tex);
  if (a->lut_tex) cudaDestroyTextureObject(a->lut_tex);
  if (a->lut_array) cudaFreeArray(a->lut_array);
  delete a;
  return PBRCU_OK;
}

extern "C" size_t pbrcu_ibl_accel_bytes(const pbrcu_ibl_accel* a) { return a ? a->bytes : 0; }

extern "C" int pbrcu_ibl_accel_build(pbrcu_ctx* ctx, const pbrcu_ibl* ibl, pbrcu_ibl_accel** out, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && ibl && out, "NULL argument");
  PBRCU_REQUIRE(ibl->irradiance && ibl->prefiltered && ibl->lut, "ibl texture is NULL");
  PBRCU_REQUIRE(ibl->irr_res > 0 && ibl->irr_res <= 4096 && ibl->pref_res > 0 && ibl->pref_res <= 4096 && ibl->pref_mips >= 1 &&
                    ibl->pref_mips <= 16 && (ibl->pref_res >> (ibl->pref_mips - 1)) >= 1 && ibl->lut_res > 0 && ibl->lut_res <= 4096,
                "bad ibl shape");
  *out = nullptr;
  pbrcu_ibl_accel* a = new (std::nothrow) pbrcu_ibl_accel();
  if (!a) return PBRCU_ERR_NOMEM;
  a->pref_res = ibl->pref_res; a->pref_mips = ibl->pref_mips; a->irr_res = ibl->irr_res; a->lut_res = ibl->lut_res;
  const size_t n_pref = cube_mip_offset(ibl->pref_res, ibl->pref_mips);
  const size_t n_irr = size_t(6) * ibl->irr_res * ibl->irr_res, n_lut = size_t(ibl->lut_res) * ibl->lut_res;
  a->bytes = (n_pref + n_irr + n_lut) * 32;
  int rc;
  if ((rc = pbrcu_malloc(ctx, n_pref * 32, reinterpret_cast<void**>(&a->pref))) ||
      (rc = pbrcu_malloc(ctx, n_irr * 32, reinterpret_cast<void**>(&a->irr))) ||
      (rc = pbrcu_malloc(ctx, n_lut * 32, reinterpret_cast<void**>(&a->lut)))) {
    pbrcu_ibl_accel_free(ctx, a);
    return rc;
  }
  cudaStream_t st = static_cast<cudaStream_t>(s);
  for (uint32_t m = 0; m < ibl->pref_mips; ++m) {
    const size_t off = cube_mip_offset(ibl->pref_res, m);
    if ((rc = build_level(static_cast<const uint2*>(ibl->prefiltered) + off, a->pref + 2 * off, int(ibl->pref_res >> m), 6, ctx->sm_count, st))) {
      pbrcu_ibl_accel_free(ctx, a);
      return rc;
    }
  }
  // the LUT's float2 texels are 8 bytes like an RGBA16F texel: the same gather builds its records (plain copy of the four)
  if ((rc = build_level(static_cast<const uint2*>(ibl->irradiance), a->irr, int(ibl->irr_res), 6, ctx->sm_count, st)) ||
      (rc = build_level(reinterpret_cast<const uint2*>(ibl->lut), a->lut, int(ibl->lut_res), 1, ctx->sm_count, st))) {
    pbrcu_ibl_accel_free(ctx, a);
    return rc;
  }
  // hardware-texture copies with contrast flags (12 % footprint range: a flag-free sample is within
  // 3/512 * 0.12 = 7e-4 relative of the fp32-weight filter; K3 itself rules out level fractions near 1)
  if ((rc = cube_tex_upload(ctx, &a->pref_tex, ibl->prefiltered, ibl->pref_res, ibl->pref_mips, kShadeTexRelRange, false, st)) ||
      (rc = cube_tex_upload(ctx, &a->irr_tex, ibl->irradiance, ibl->irr_res, 1, kShadeTexRelRange, false, st))) {
    pbrcu_ibl_accel_free(ctx, a);
    return rc;
  }
  a->bytes += cube_tex_bytes(a->pref_tex) + cube_tex_bytes(a->irr_tex);
  {
    const cudaChannelFormatDesc fmt = cudaCreateChannelDesc<float2>();
    cudaError_t e = cudaMallocArray(&a->lut_array, &fmt, ibl->lut_res, ibl->lut_res);
    if (e == cudaSuccess)
      e = cudaMemcpy2DToArrayAsync(a->lut_array, 0, 0, ibl->lut, size_t(ibl->lut_res) * 8, size_t(ibl->lut_res) * 8, ibl->lut_res,
                                   cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess) {
      cudaResourceDesc rd{};
      rd.resType = cudaResourceTypeArray;
      rd.res.array.array = a->lut_array;
      cudaTextureDesc td{};
      td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
      td.filterMode = cudaFilterModeLinear;
      td.readMode = cudaReadModeElementType;
      td.normalizedCoords = 1;
      e = cudaCreateTextureObject(&a->lut_tex, &rd, &td, nullptr);
    }
    if (e != cudaSuccess) {
      pbrcu_ibl_accel_free(ctx, a);
      return cuda_fail(e, "LUT texture");
    }
    a->bytes += n_lut * 8;
  }
  *out = a;
  return PBRCU_OK;
}
