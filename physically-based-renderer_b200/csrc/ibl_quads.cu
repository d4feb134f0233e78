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

// ---- gather atlases (GatherTex, ibl_quads.cuh) ----
// one level: every slot texel, gutter included, copies the face texel its coordinates clamp to
// (grid = (ceil((res + 2) / 32), ceil((res + 2) / 8), faces), block = 32 x 8)
__global__ void __launch_bounds__(256) build_atlas_kernel(const uint2* __restrict__ src, uint2* __restrict__ atlas, size_t atlas_w, int res,
                                                          uint32_t ax, uint32_t ay) {
  const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (x < res + 2 && y < res + 2) {
    const size_t face = blockIdx.z;
    const int sx = min(max(x - 1, 0), res - 1), sy = min(max(y - 1, 0), res - 1);
    atlas[(size_t(ay) + face * size_t(res + 2) + y) * atlas_w + ax + x] = __ldg(src + (face * res + sy) * size_t(res) + sx);
  }
}

void gather_tex_destroy(GatherTex* t) {
  if (t->tex) cudaDestroyTextureObject(t->tex);
  if (t->array) cudaFreeArray(t->array);
  t->tex = 0; t->array = nullptr; t->width = t->height = 0;
}

// chain: `faces` faces per level, 8-byte texels (RGBA16F, or RG32F when rg32f); builds the atlas in a stream-ordered
// staging buffer and copies it into a 2-D array created for texture gathers
static int gather_tex_build(pbrcu_ctx* ctx, GatherTex* t, const void* chain, uint32_t base_res, uint32_t nmips, uint32_t faces, bool rg32f,
                            cudaStream_t st) {
  uint32_t w, h;
  atlas_extent(base_res, nmips, faces, w, h);
  cudaDeviceProp prop;
  PBRCU_CUDA(cudaGetDeviceProperties(&prop, ctx->device));
  if (int(w) > prop.maxTexture2DGather[0] || int(h) > prop.maxTexture2DGather[1]) {
    set_error("gather atlas of %u x %u texels exceeds the device's texture-gather limit (%d x %d)", w, h, prop.maxTexture2DGather[0], prop.maxTexture2DGather[1]);
    return PBRCU_ERR_INVALID;
  }
  const cudaChannelFormatDesc fmt = rg32f ? cudaCreateChannelDesc<float2>() : cudaCreateChannelDescHalf4();
  PBRCU_CUDA(cudaMallocArray(&t->array, &fmt, w, h, cudaArrayTextureGather));
  t->width = w; t->height = h;
  uint2* stage = nullptr;
  PBRCU_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&stage), size_t(w) * h * 8, st));
  PBRCU_CUDA(cudaMemsetAsync(stage, 0, size_t(w) * h * 8, st));
  const uint2* src = static_cast<const uint2*>(chain);
  size_t off = 0;
  for (uint32_t m = 0; m < nmips; ++m) {
    const AtlasLevel a = atlas_level(base_res, m);
    const dim3 grid((a.res + 2 + 31) / 32, (a.res + 2 + 7) / 8, faces);
    build_atlas_kernel<<<grid, dim3(32, 8), 0, st>>>(src + off, stage, w, int(a.res), a.ax, a.ay);
    off += size_t(faces) * a.res * a.res;
  }
  if (int rc = launch_check("build_atlas_kernel")) return rc;
  PBRCU_CUDA(cudaMemcpy2DToArrayAsync(t->array, 0, 0, stage, size_t(w) * 8, size_t(w) * 8, h, cudaMemcpyDeviceToDevice, st));
  PBRCU_CUDA(cudaFreeAsync(stage, st));
  cudaResourceDesc rd{};
  rd.resType = cudaResourceTypeArray;
  rd.res.array.array = t->array;
  cudaTextureDesc td{};
  td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
  td.filterMode = cudaFilterModeLinear;  // (a gather always takes the bilinear footprint; the kernels pass its exact centre)
  td.readMode = cudaReadModeElementType;
  td.normalizedCoords = 0;
  PBRCU_CUDA(cudaCreateTextureObject(&t->tex, &rd, &td, nullptr));
  return PBRCU_OK;
}

}  // namespace pbrcu

using namespace pbrcu;

extern "C" int pbrcu_ibl_accel_free(pbrcu_ctx* ctx, pbrcu_ibl_accel* a) {
  (void)ctx;
  if (!a) return PBRCU_OK;
  if (a->pref) cudaFree(a->pref);
  if (a->irr) cudaFree(a->irr);
  if (a->lut) cudaFree(a->lut);
  gather_tex_destroy(&a->g_pref);
  gather_tex_destroy(&a->g_irr);
  gather_tex_destroy(&a->g_lut);
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
  if (!(ibl->flags & PBRCU_IBL_GATHER)) {  // the atlases serve the opt-in gather mode only
    *out = a;
    return PBRCU_OK;
  }
  if ((rc = gather_tex_build(ctx, &a->g_pref, ibl->prefiltered, ibl->pref_res, ibl->pref_mips, 6, false, st)) ||
      (rc = gather_tex_build(ctx, &a->g_irr, ibl->irradiance, ibl->irr_res, 1, 6, false, st)) ||
      (rc = gather_tex_build(ctx, &a->g_lut, ibl->lut, ibl->lut_res, 1, 1, true, st))) {
    pbrcu_ibl_accel_free(ctx, a);
    return rc;
  }
  a->bytes += (size_t(a->g_pref.width) * a->g_pref.height + size_t(a->g_irr.width) * a->g_irr.height + size_t(a->g_lut.width) * a->g_lut.height) * 8;
  *out = a;
  return PBRCU_OK;
}
