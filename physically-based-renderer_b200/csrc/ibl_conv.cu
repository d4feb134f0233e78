// ibl_conv.cu -- K7 GGX-prefiltered specular chain and K6 cosine irradiance: Monte-Carlo convolution
// of a cube-map mip chain over a fixed Hammersley sample set (filtered importance sampling).
// No reference counterpart; definitions: oracle/pbr_oracle_ibl.c (Karis 2013, Krivanek &
// Colbert 2008).
//
// Under N = V = R the sample set in the local frame depends only on (roughness, i), so each block
// builds the table {L_local.xyz, source lod} once in shared memory; per sample a thread then does a
// 3x3 frame rotation, one exact-weight trilinear cube fetch and 4 FMAs.  G lanes share one texel
// (samples strided by G, warp-shuffle reduction) so the small mips still fill the machine.
#include <new>

#include "cube_tex.cuh"
#include "detmath.cuh"

namespace pbrcu {

enum : int { CONV_GGX = 0, CONV_COSINE = 1 };
constexpr int kConvThreads = 256;

struct ConvParams {
  CubeChain src;
  cudaTextureObject_t tex;  // TEX path: the source chain as a mipmapped cube-map texture
  uint2* dst_level;   // first texel of the destination level
  int res;            // destination level resolution
  int nsamples;
  float roughness;    // CONV_GGX only
  int row_begin, row_end;  // global rows (face*res + y) handled by this launch (multi-GPU shard)
  int tw, th;         // texel tile per block: tw*th*G == kConvThreads
  // pbrcu_ibl_prefilter_ggx_replicated: the finished texel is also stored at the same index of these levels (the
  // peers' chains mapped into this process -- the all-gather of a sharded prefilter fused into the kernel's stores)
  int n_replicas;
  uint2* replica_level[PBRCU_MAX_REPLICAS];
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

// TEX = true: one hardware trilinear cube fetch per sample (face selection, clamp-to-edge bilinear on two
// levels and the level blend all happen in the texture unit: ~20 instructions per sample instead of
// ~150, bound by the unit's RGBA16F rate of ~2 fetches/clk/SM).  The unit filters with 8-bit
// fixed-point weights, i.e. a sample is off by up to 3/512 of the value range of its 8-texel footprint.
// The texture copy therefore carries a flag in alpha wherever that range exceeds 25 % of the footprint's
// minimum; flagged samples fall back to the fp32-weight software filter, so every sample -- and hence
// the weighted mean -- is within 1.5e-3 relative of the exact filter (tolerance 2^-9).
// TEX = false: fp32 weights everywhere.
template <int G, int MODE, bool TEX>
__global__ void __launch_bounds__(kConvThreads) conv_kernel(const __grid_constant__ ConvParams p) {
  extern __shared__ float4 table[];
  for (int s = threadIdx.x; s < p.nsamples; s += kConvThreads) table[s] = table_entry<MODE>(p, uint32_t(s));
  __syncthreads();

  // A warp holds TPW = 32 / G texels, each with G lanes that take every G-th sample.  The texels of a warp are a compact
  // patch walked in Morton order (8x4 for G = 1, 4x4, 4x2, 2x2, 2x1, 1x1), lane = sample group * TPW + texel: the four lanes
  // the texture unit filters together (4k .. 4k+3) are then a 2x2 block of texels taking the SAME sample, so their
  // footprints overlap in the source level, instead of a 4x1 strip (G = 1: 25.7 -> 18.7 ms for the 2048^2 chain) or four
  // unrelated samples of one texel (G >= 4).
  constexpr int TPW = 32 / G;
  constexpr int PW = TPW >= 32 ? 8 : TPW >= 8 ? 4 : TPW >= 2 ? 2 : 1, PH = TPW / PW;
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, t = l % TPW, lane = l / TPW;
  const int mx = (t & 1) | ((t >> 1) & 2) | ((t >> 2) & 4), my = ((t >> 1) & 1) | ((t >> 2) & 2);
  const int ppr = p.tw / PW;  // patches per row of the block's tile
  const int tx_in = (w % ppr) * PW + mx, ty_in = (w / ppr) * PH + my;
  const int x = blockIdx.x * p.tw + tx_in;
  const int grow = p.row_begin + blockIdx.y * p.th + ty_in;
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
  auto dir_of = [&](const float4 e) {
    return make_float3(T.x * e.x + B.x * e.y + N.x * e.z, T.y * e.x + B.y * e.y + N.y * e.z, T.z * e.x + B.z * e.y + N.z * e.z);
  };
  auto add = [&](const float4 e, const RGB c) {
    const float w = (MODE == CONV_GGX) ? e.z : 1.0f;
    ar = fmaf(c.r, w, ar); ag = fmaf(c.g, w, ag); ab = fmaf(c.b, w, ab);
    wsum += w;
  };
  int s = lane;
  if (TEX) {
    // four samples per iteration: their fetches are issued back to back, BEFORE any of the flag tests, so a thread has four
    // fetches in flight instead of one (the small levels run a single partial wave and were bound by the fetch latency:
    // level 4 of the 2048^2 chain 0.70 -> see DESIGN 3.2).  The sums are taken in sample order, as before.
    for (; s + 3 * G < p.nsamples; s += 4 * G) {
      float4 e[4], t[4];
      float3 L[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        e[k] = table[s + k * G];
        L[k] = dir_of(e[k]);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k)  // Karis: `if (NoL > 0)`; a sample below the horizon is fetched and ignored
        t[k] = texCubemapLod<float4>(p.tex, L[k].x, L[k].y, L[k].z, e[k].w);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (e[k].z > 0.0f) add(e[k], t[k].w > 0.0f ? cube_sample(p.src, L[k], e[k].w) : RGB{t[k].x, t[k].y, t[k].z});
    }
  }
  for (; s < p.nsamples; s += G) {
    const float4 e = table[s];
    if (e.z > 0.0f) {  // Karis: `if (NoL > 0)`
      const float3 L = dir_of(e);
      RGB c;
      if (TEX) {
        // alpha of the texture copy = contrast flag of the footprint (see contrast_flags_kernel): where the
        // 8 texels differ by more than 25 % the unit's 8-bit weights are not accurate enough, so that
        // sample is filtered in software from the plain chain instead
        const float4 t = texCubemapLod<float4>(p.tex, L.x, L.y, L.z, e.w);
        c = t.w > 0.0f ? cube_sample(p.src, L, e.w) : RGB{t.x, t.y, t.z};
      } else {
        c = cube_sample(p.src, L, e.w);
      }
      add(e, c);
    }
  }
#pragma unroll
  for (int off = TPW; off < 32; off <<= 1) {  // warp-shuffle reduction over the texel's G lanes (TPW apart)
    ar += __shfl_xor_sync(0xffffffffu, ar, off);
    ag += __shfl_xor_sync(0xffffffffu, ag, off);
    ab += __shfl_xor_sync(0xffffffffu, ab, off);
    wsum += __shfl_xor_sync(0xffffffffu, wsum, off);
  }
  if (valid && lane == 0) {
    const float inv = 1.0f / wsum;
    const size_t idx = (size_t(face) * p.res + y) * p.res + x;
    const uint2 texel = make_uint2(f2_to_h2(fminf(ar * inv, 65504.f), fminf(ag * inv, 65504.f)),
                                   f2_to_h2(fminf(ab * inv, 65504.f), 1.0f));
    p.dst_level[idx] = texel;
    for (int k = 0; k < p.n_replicas; ++k) p.replica_level[k][idx] = texel;
  }
}

template <int MODE, bool TEX>
static int launch_conv(const pbrcu_ctx* ctx, ConvParams p, cudaStream_t s) {
  const int rows = p.row_end - p.row_begin;
  if (rows <= 0) return PBRCU_OK;
  // lanes per texel: enough threads to fill 148 SMs x 1024 threads, at most a warp.  Chosen from the WHOLE level,
  // not from this launch's rows: G fixes the order in which a texel's samples are summed, and a shard of the
  // level must produce the same bits as the unsharded run
  const size_t texels = size_t(6) * p.res * p.res;
  int G = 1;
  while (G < 32 && texels * G < size_t(ctx->sm_count) * 1024) G *= 2;
  // block tile = the 8 warps' patches, 4 x 2 of them where the level is wide enough
  const int tpw = 32 / G, pw = tpw >= 32 ? 8 : tpw >= 8 ? 4 : tpw >= 2 ? 2 : 1, ph = tpw / pw;
  int ppr = 4;
  while (ppr > 1 && ppr * pw > p.res) ppr /= 2;
  p.tw = ppr * pw;
  p.th = (kConvThreads / 32 / ppr) * ph;
  dim3 grid((p.res + p.tw - 1) / p.tw, (rows + p.th - 1) / p.th);
  const size_t smem = size_t(p.nsamples) * sizeof(float4);
#define PBRCU_CONV_CASE(GG)                                                                          \
  case GG:                                                                                           \
    if (smem > 48 * 1024)                                                                            \
      PBRCU_CUDA(cudaFuncSetAttribute(conv_kernel<GG, MODE, TEX>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem))); \
    conv_kernel<GG, MODE, TEX><<<grid, kConvThreads, smem, s>>>(p);                                  \
    break;
  switch (G) {
    PBRCU_CONV_CASE(1) PBRCU_CONV_CASE(2) PBRCU_CONV_CASE(4) PBRCU_CONV_CASE(8) PBRCU_CONV_CASE(16) PBRCU_CONV_CASE(32)
  }
#undef PBRCU_CONV_CASE
  return launch_check("conv_kernel");
}

// ---- contrast flags for the texture-unit path ----
// pass A: per-channel min / max over the 3x3 neighbourhood (clamped to the face) of every texel
// (grid = (ceil(res / 32), ceil(res / 8), 6), block = 32 x 8: face, row and column straight from the indices)
__global__ void __launch_bounds__(256) minmax3_kernel(const uint2* __restrict__ level, uint4* __restrict__ mm, int res) {
  const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (x < res && y < res) {
    const size_t i = (size_t(blockIdx.z) * res + y) * res + x;
    const uint2* f = level + size_t(blockIdx.z) * res * res;
    float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx) {
        const RGB t = texel_rgb(f + size_t(min(max(y + dy, 0), res - 1)) * res + min(max(x + dx, 0), res - 1));
        mn[0] = fminf(mn[0], t.r); mn[1] = fminf(mn[1], t.g); mn[2] = fminf(mn[2], t.b);
        mx[0] = fmaxf(mx[0], t.r); mx[1] = fmaxf(mx[1], t.g); mx[2] = fmaxf(mx[2], t.b);
      }
    mm[i] = make_uint4(f2_to_h2(mn[0], mn[1]), f2_to_h2(mn[2], 0.f), f2_to_h2(mx[0], mx[1]), f2_to_h2(mx[2], 0.f));  // exact: the texels are fp16
  }
}
struct MinMax { float mn[3], mx[3]; };
__device__ __forceinline__ void mm_merge(MinMax& a, const uint4* p) {
  const uint4 v = __ldg(p);
  const float2 a0 = h2_to_f2(v.x), a1 = h2_to_f2(v.y), b0 = h2_to_f2(v.z), b1 = h2_to_f2(v.w);
  a.mn[0] = fminf(a.mn[0], a0.x); a.mn[1] = fminf(a.mn[1], a0.y); a.mn[2] = fminf(a.mn[2], a1.x);
  a.mx[0] = fmaxf(a.mx[0], b0.x); a.mx[1] = fmaxf(a.mx[1], b0.y); a.mx[2] = fmaxf(a.mx[2], b1.x);
}
// pass B: a texel is flagged when the value range over every footprint it can share a trilinear sample
// with exceeds rel_range x the minimum.  As the finer level of a sample its own 3x3 holds the 2x2 footprint
// and its parent's 3x3 holds the coarser footprint; as the coarser level its own 3x3 holds its footprint
// and the 3x3s of the 4x4 block of texels around its children hold the finer one.
__global__ void __launch_bounds__(256) contrast_flags_kernel(const uint2* __restrict__ level, const uint4* __restrict__ mm,
                                                             const uint4* __restrict__ mm_up, const uint4* __restrict__ mm_down,
                                                             uint2* __restrict__ out, int res, float rel_range) {
  const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (x < res && y < res) {
    const size_t face = blockIdx.z;
    const size_t i = (face * res + y) * res + x;
    MinMax a{{3.0e38f, 3.0e38f, 3.0e38f}, {-3.0e38f, -3.0e38f, -3.0e38f}};
    mm_merge(a, mm + i);
    if (mm_up != nullptr) {
      const int r1 = res >> 1;
      mm_merge(a, mm_up + (face * r1 + (y >> 1)) * r1 + (x >> 1));
    }
    if (mm_down != nullptr) {
      const int r0 = res * 2;
      // the 3x3 windows around the 4x4 block of children [2x-1, 2x+2]^2 cover the texels [2x-2, 2x+3]^2; min / max are idempotent, so the
      // FOUR windows centred on the block's corners cover the same texels ([2x-2, 2x] and [2x+1, 2x+3] per axis, clamped alike): 4 loads, not 16
      for (int cy = 2 * y - 1; cy <= 2 * y + 2; cy += 3)
        for (int cx = 2 * x - 1; cx <= 2 * x + 2; cx += 3)
          mm_merge(a, mm_down + (face * r0 + min(max(cy, 0), r0 - 1)) * r0 + min(max(cx, 0), r0 - 1));
    }
    bool flag = false;
#pragma unroll
    for (int c = 0; c < 3; ++c) flag = flag || !(a.mx[c] <= fmaf(rel_range, a.mn[c], a.mn[c]) + 1.0e-4f);  // NaN -> flagged
    const uint2 t = __ldg(level + i);
    out[i] = make_uint2(t.x, (t.y & 0xFFFFu) | (flag ? 0x3C000000u : 0u));  // alpha = 1.0h or 0
  }
}

static int grow(void** buf, size_t* have, size_t want) {
  if (*have >= want) return PBRCU_OK;
  if (*buf) PBRCU_CUDA(cudaFree(*buf));
  *buf = nullptr; *have = 0;
  PBRCU_CUDA(cudaMalloc(buf, want));
  *have = want;
  return PBRCU_OK;
}

}  // namespace pbrcu

void pbrcu::cube_tex_destroy(CubeTex* t) {
  if (t->tex) cudaDestroyTextureObject(t->tex);
  if (t->array) cudaFreeMipmappedArray(t->array);
  t->tex = 0; t->array = nullptr; t->res = t->mips = 0;
}

int pbrcu::cube_tex_upload(pbrcu_ctx* ctx, CubeTex* t, const void* chain, uint32_t base_res, uint32_t nmips, float rel_range,
                           bool guard_coarser, cudaStream_t st) {
  if (!t->tex || t->res != base_res || t->mips != nmips) {  // array allocation costs milliseconds: keep it while the shape holds
    cube_tex_destroy(t);
    const cudaChannelFormatDesc fmt = cudaCreateChannelDescHalf4();
    PBRCU_CUDA(cudaMallocMipmappedArray(&t->array, &fmt, make_cudaExtent(base_res, base_res, 6), nmips, cudaArrayCubemap));
    cudaResourceDesc rd{};
    rd.resType = cudaResourceTypeMipmappedArray;
    rd.res.mipmap.mipmap = t->array;
    cudaTextureDesc td{};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;  // per face: no seamless filtering
    td.filterMode = cudaFilterModeLinear;
    td.mipmapFilterMode = cudaFilterModeLinear;
    td.readMode = cudaReadModeElementType;
    td.normalizedCoords = 1;
    td.maxMipmapLevelClamp = float(nmips - 1);
    PBRCU_CUDA(cudaCreateTextureObject(&t->tex, &rd, &td, nullptr));
    t->res = base_res;
    t->mips = nmips;
  }
  const size_t total = cube_mip_offset(base_res, nmips);
  if (int rc = grow(&ctx->cube_stage, &ctx->cube_stage_bytes, total * 8)) return rc;
  if (int rc = grow(&ctx->cube_mm, &ctx->cube_mm_bytes, total * 16)) return rc;
  const uint2* src = static_cast<const uint2*>(chain);
  uint4* mm = static_cast<uint4*>(ctx->cube_mm);
  uint2* stage = static_cast<uint2*>(ctx->cube_stage);
  auto grid_of = [](size_t r) { return dim3(unsigned((r + 31) / 32), unsigned((r + 7) / 8), 6); };
  for (uint32_t m = 0; m < nmips; ++m) {
    const size_t off = cube_mip_offset(base_res, m), r = base_res >> m;
    minmax3_kernel<<<grid_of(r), dim3(32, 8), 0, st>>>(src + off, mm + off, int(r));
  }
  if (int rc = launch_check("minmax3_kernel")) return rc;
  for (uint32_t m = 0; m < nmips; ++m) {
    const size_t off = cube_mip_offset(base_res, m), r = base_res >> m;
    contrast_flags_kernel<<<grid_of(r), dim3(32, 8), 0, st>>>(src + off, mm + off, m + 1 < nmips ? mm + cube_mip_offset(base_res, m + 1) : nullptr,
                                                             m > 0 && guard_coarser ? mm + cube_mip_offset(base_res, m - 1) : nullptr, stage + off, int(r), rel_range);
  }
  if (int rc = launch_check("contrast_flags_kernel")) return rc;
  for (uint32_t m = 0; m < nmips; ++m) {
    cudaArray_t level;
    PBRCU_CUDA(cudaGetMipmappedArrayLevel(&level, t->array, m));
    const size_t r = base_res >> m;
    cudaMemcpy3DParms cp{};
    cp.srcPtr = make_cudaPitchedPtr(stage + cube_mip_offset(base_res, m), r * 8, r, r);
    cp.dstArray = level;
    cp.extent = make_cudaExtent(r, r, 6);
    cp.kind = cudaMemcpyDeviceToDevice;
    PBRCU_CUDA(cudaMemcpy3DAsync(&cp, st));
  }
  return PBRCU_OK;
}

namespace pbrcu {

static void shard_rows(int total_rows, uint32_t shard, uint32_t nshards, int& b, int& e) {
  b = int(uint64_t(total_rows) * shard / nshards);
  e = int(uint64_t(total_rows) * (shard + 1) / nshards);
}

}  // namespace pbrcu

using namespace pbrcu;

extern "C" int pbrcu_ibl_set_exact_filtering(pbrcu_ctx* ctx, int exact) {
  PBRCU_REQUIRE(ctx, "ctx is NULL");
  ctx->exact_ibl_filter = exact != 0;
  return PBRCU_OK;
}

static int prefilter_common(pbrcu_ctx* ctx, const void* src_chain_dev, uint32_t base_res, uint32_t nmips,
                            uint32_t nsamples, void* dst_chain_dev, uint32_t first_mip, uint32_t last_mip,
                            uint32_t shard, uint32_t nshards, void* const* replicas, uint32_t n_replicas, pbrcu_stream s,
                            const CubeTex* prepared = nullptr) {
  PBRCU_REQUIRE(ctx && src_chain_dev && dst_chain_dev, "NULL argument");
  PBRCU_REQUIRE(n_replicas <= PBRCU_MAX_REPLICAS && (n_replicas == 0 || replicas), "too many replicas or replicas NULL");
  for (uint32_t k = 0; k < n_replicas; ++k) PBRCU_REQUIRE(replicas[k] && reinterpret_cast<uintptr_t>(replicas[k]) % 8 == 0, "replica NULL or not 8-byte aligned");
  PBRCU_REQUIRE(base_res > 0 && (base_res & (base_res - 1)) == 0, "base_res must be a power of two");
  PBRCU_REQUIRE(nmips >= 2 && nmips <= 16 && (base_res >> (nmips - 1)) >= 1, "nmips out of range for base_res");
  PBRCU_REQUIRE(first_mip <= last_mip && last_mip < nmips, "mip range out of bounds");
  PBRCU_REQUIRE(nsamples > 0 && size_t(nsamples) * 16 <= size_t(ctx->max_smem_optin), "nsamples out of range");
  PBRCU_REQUIRE(nshards >= 1 && shard < nshards, "shard out of range");
  cudaStream_t st = static_cast<cudaStream_t>(s);
  const bool tex = !ctx->exact_ibl_filter;
  if (tex && last_mip >= 1 && !prepared)  // (a prepared source, pbrcu_ibl_source, carries its own flagged texture copy)
    if (int rc = cube_tex_upload(ctx, &ctx->cube, src_chain_dev, base_res, nmips, 0.25f, true, st)) return rc;
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
      for (uint32_t k = 0; k < n_replicas; ++k)
        PBRCU_CUDA(cudaMemcpyAsync(static_cast<uint2*>(replicas[k]) + off + size_t(rb) * res, src + size_t(rb) * res, size_t(re - rb) * res * 8,
                                   cudaMemcpyDefault, st));
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
    p.tex = prepared ? prepared->tex : ctx->cube.tex;
    p.n_replicas = int(n_replicas);
    for (uint32_t k = 0; k < n_replicas; ++k) p.replica_level[k] = static_cast<uint2*>(replicas[k]) + off;
    if (int rc = tex ? launch_conv<CONV_GGX, true>(ctx, p, st) : launch_conv<CONV_GGX, false>(ctx, p, st)) return rc;
  }
  return PBRCU_OK;
}

extern "C" int pbrcu_ibl_prefilter_ggx(pbrcu_ctx* ctx, const void* src_chain_dev, uint32_t base_res, uint32_t nmips,
                                       uint32_t nsamples, void* dst_chain_dev, uint32_t first_mip, uint32_t last_mip,
                                       uint32_t shard, uint32_t nshards, pbrcu_stream s) {
  return prefilter_common(ctx, src_chain_dev, base_res, nmips, nsamples, dst_chain_dev, first_mip, last_mip, shard, nshards, nullptr, 0, s);
}

extern "C" int pbrcu_ibl_prefilter_ggx_replicated(pbrcu_ctx* ctx, const void* src_chain_dev, uint32_t base_res, uint32_t nmips,
                                                  uint32_t nsamples, void* dst_chain_dev, uint32_t first_mip, uint32_t last_mip,
                                                  uint32_t shard, uint32_t nshards, void* const* replica_chains, uint32_t n_replicas,
                                                  pbrcu_stream s) {
  return prefilter_common(ctx, src_chain_dev, base_res, nmips, nsamples, dst_chain_dev, first_mip, last_mip, shard, nshards,
                          replica_chains, n_replicas, s);
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
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (ctx->exact_ibl_filter) return launch_conv<CONV_COSINE, false>(ctx, p, st);
  if (int rc = cube_tex_upload(ctx, &ctx->cube, src_chain_dev, base_res, nmips, 0.25f, true, st)) return rc;
  p.tex = ctx->cube.tex;
  return launch_conv<CONV_COSINE, true>(ctx, p, st);
}

// ---- a prepared source: the flagged texture copy K6 / K7 sample, built ONCE for all the calls that convolve one environment ----
struct pbrcu_ibl_source {
  pbrcu::CubeTex tex;
  const void* chain;  // the plain chain (caller-owned): flagged samples are re-filtered from it with fp32 weights
  uint32_t base_res, nmips;
};

extern "C" int pbrcu_ibl_source_create(pbrcu_ctx* ctx, const void* src_chain_dev, uint32_t base_res, uint32_t nmips, pbrcu_stream s,
                                       pbrcu_ibl_source** out) {
  PBRCU_REQUIRE(ctx && src_chain_dev && out, "NULL argument");
  PBRCU_REQUIRE(base_res > 0 && (base_res & (base_res - 1)) == 0, "base_res must be a power of two");
  PBRCU_REQUIRE(nmips >= 1 && nmips <= 16 && (base_res >> (nmips - 1)) >= 1, "nmips out of range for base_res");
  *out = nullptr;
  pbrcu_ibl_source* src = new (std::nothrow) pbrcu_ibl_source();
  if (!src) return PBRCU_ERR_NOMEM;
  src->tex = CubeTex{};
  src->chain = src_chain_dev; src->base_res = base_res; src->nmips = nmips;
  if (int rc = cube_tex_upload(ctx, &src->tex, src_chain_dev, base_res, nmips, 0.25f, true, static_cast<cudaStream_t>(s))) {
    cube_tex_destroy(&src->tex);
    delete src;
    return rc;
  }
  *out = src;
  return PBRCU_OK;
}

extern "C" int pbrcu_ibl_source_free(pbrcu_ctx* ctx, pbrcu_ibl_source* src) {
  (void)ctx;
  if (!src) return PBRCU_OK;
  cube_tex_destroy(&src->tex);
  delete src;
  return PBRCU_OK;
}

extern "C" int pbrcu_ibl_prefilter_ggx_prepared(pbrcu_ctx* ctx, const pbrcu_ibl_source* src, uint32_t nsamples, void* dst_chain_dev,
                                                uint32_t first_mip, uint32_t last_mip, uint32_t shard, uint32_t nshards,
                                                void* const* replica_chains, uint32_t n_replicas, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && src, "NULL argument");
  PBRCU_REQUIRE(src->nmips >= 2, "the source chain needs at least two levels");
  return prefilter_common(ctx, src->chain, src->base_res, src->nmips, nsamples, dst_chain_dev, first_mip, last_mip, shard, nshards, replica_chains,
                          n_replicas, s, &src->tex);
}

extern "C" int pbrcu_ibl_irradiance_prepared(pbrcu_ctx* ctx, const pbrcu_ibl_source* src, uint32_t res, uint32_t nsamples, void* out_dev,
                                             pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && src && out_dev, "NULL argument");
  PBRCU_REQUIRE(res > 0 && nsamples > 0 && size_t(nsamples) * 16 <= size_t(ctx->max_smem_optin), "res or nsamples out of range");
  ConvParams p{};
  p.src = make_chain(src->chain, src->base_res, src->nmips);
  p.dst_level = static_cast<uint2*>(out_dev);
  p.res = int(res);
  p.nsamples = int(nsamples);
  p.row_begin = 0;
  p.row_end = 6 * int(res);
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (ctx->exact_ibl_filter) return launch_conv<CONV_COSINE, false>(ctx, p, st);
  p.tex = src->tex.tex;
  return launch_conv<CONV_COSINE, true>(ctx, p, st);
}
