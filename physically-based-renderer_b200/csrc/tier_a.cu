// tier_a.cu -- K1 deferred Blinn-Phong lighting, K2 ACES+gamma tonemap, and the fused K1+K2 pass.
// These replace PbrRenderSystem::recordLightingPass (PbrRenderSystem.cpp:466-501) and
// TonemapperSystem::run (TonemapperSystem.cpp:97-134).  HBM-bound streaming kernels: 128-bit
// coalesced loads of two 8-byte texels per plane per thread, two independent texel pairs in flight
// per thread, persistent grid sized to the SM count, evict-first/streaming cache policy.
#include "tier_a_math.cuh"

namespace pbrcu {

enum : int { MODE_LIGHT = 0, MODE_TONEMAP = 1, MODE_FUSED_HDR_LDR = 2, MODE_FUSED_LDR = 3 };

struct TierAParams {
  const uint8_t* pos;
  const uint8_t* nrm;
  const uint8_t* alb;
  const uint8_t* hdr_in;  // MODE_TONEMAP input
  uint8_t* hdr;           // RGBA16F output
  uint8_t* ldr;           // RGBA8 output
  size_t pitch8, hdr_pitch, ldr_pitch;
  uint32_t w, h;          // rect extent in texels (plane pointers already offset to the rect origin)
  float cx, cy, cz;
  // batched multi-view: blockIdx.y selects the view
  const float* cams;      // device, n_views*3 (NULL: use cx,cy,cz)
  size_t view_stride8, hdr_view_stride, ldr_view_stride;
};

constexpr int kThreads = 256;
constexpr int kUnroll = 2;  // independent texel PAIRS per thread per iteration

template <int MODE>
__device__ __forceinline__ void process_pair(uint8_t* hdr, uint8_t* ldr, uint4 a, uint4 b, uint4 c,
                                             float cx, float cy, float cz) {
  F4x2 cc;
  if (MODE == MODE_TONEMAP) {
    cc = unpack_rgba16f_pair(a);
  } else {
    cc = blinnphong_pair(a, b, c, cx, cy, cz);
    const uint4 h = pack_rgba16f_pair(cc);
    if (MODE != MODE_FUSED_LDR) stg_stream(hdr, h);
    if (MODE != MODE_LIGHT) cc = unpack_rgba16f_pair(h);  // the tonemapper reads the fp16-rounded attachment value
  }
  if (MODE != MODE_LIGHT) stg_stream8(ldr, tonemap_pair(cc));
}

template <int MODE>
__device__ __forceinline__ void process_single(const uint8_t* in0, const uint8_t* in1, const uint8_t* in2,
                                               uint8_t* hdr, uint8_t* ldr, float cx, float cy, float cz) {
  F4 c0;
  if (MODE == MODE_TONEMAP) {
    c0 = unpack_rgba16f(ldg_stream8(in0));
  } else {
    c0 = blinnphong_texel(ldg_stream8(in0), ldg_stream8(in1), ldg_stream8(in2), cx, cy, cz);
  }
  if (MODE != MODE_TONEMAP) {
    uint2 h0 = pack_rgba16f(c0);
    if (MODE != MODE_FUSED_LDR) stg_stream8(hdr, h0);
    if (MODE != MODE_LIGHT) c0 = unpack_rgba16f(h0);
  }
  if (MODE != MODE_LIGHT) stg_stream4(ldr, tonemap_texel(c0));
}

// VEC = 2: every row start and pitch is 16-byte aligned -> 128-bit accesses of texel pairs.
// VEC = 1: arbitrary alignment (odd x0, odd pitch): 64-bit accesses, one texel per item.
// FLAT (VEC = 2 only): the image is one contiguous run of texel pairs (w = width * height, h = 1) below 2^31 texels:
// 32-bit item arithmetic, one multiply-add per address.
template <int MODE, int VEC, bool FLAT>
__global__ void __launch_bounds__(kThreads) tier_a_kernel(const __grid_constant__ TierAParams p) {
  float cx = p.cx, cy = p.cy, cz = p.cz;
  const uint8_t *pos = p.pos, *nrm = p.nrm, *alb = p.alb, *hdr_in = p.hdr_in;
  uint8_t *hdr = p.hdr, *ldr = p.ldr;
  if (p.cams != nullptr) {  // batched multi-view: blockIdx.y selects the view
    const uint32_t v = blockIdx.y;
    cx = __ldg(p.cams + 3 * v);
    cy = __ldg(p.cams + 3 * v + 1);
    cz = __ldg(p.cams + 3 * v + 2);
    const size_t o8 = size_t(v) * p.view_stride8;
    pos += o8; nrm += o8; alb += o8;
    if (hdr) hdr += size_t(v) * p.hdr_view_stride;
    if (ldr) ldr += size_t(v) * p.ldr_view_stride;
  }
  if (FLAT) {
    const uint32_t npairs = p.w / 2, step = gridDim.x * kThreads * kUnroll;
    for (uint32_t base = blockIdx.x * kThreads * kUnroll + threadIdx.x; base < npairs; base += step) {
      uint4 a[kUnroll], b[kUnroll], c[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {  // issue every load of the iteration before any math
        const uint32_t item = base + u * kThreads;
        if (item < npairs) {
          if (MODE == MODE_TONEMAP) {
            a[u] = ldg_stream(hdr_in + size_t(item) * 16);
          } else {
            a[u] = ldg_stream(pos + size_t(item) * 16);
            b[u] = ldg_stream(nrm + size_t(item) * 16);
            c[u] = ldg_stream(alb + size_t(item) * 16);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const uint32_t item = base + u * kThreads;
        if (item < npairs) process_pair<MODE>(hdr + size_t(item) * 16, ldr + size_t(item) * 8, a[u], b[u], c[u], cx, cy, cz);
      }
    }
    return;
  }
  const uint32_t items_per_row = (VEC == 2) ? (p.w + 1) / 2 : p.w;
  const uint64_t total = uint64_t(items_per_row) * p.h;
  const uint64_t stride = uint64_t(gridDim.x) * kThreads * kUnroll;
  for (uint64_t base = uint64_t(blockIdx.x) * kThreads * kUnroll + threadIdx.x; base < total;
       base += stride) {
    size_t off8[kUnroll], offh[kUnroll], offl[kUnroll];
    bool valid[kUnroll], pair[kUnroll];
    uint4 a[kUnroll], b[kUnroll], c[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      uint64_t item = base + uint64_t(u) * kThreads;
      valid[u] = item < total;
      uint32_t row = 0, col = uint32_t(item);
      if (p.h > 1) {
        row = uint32_t(item / items_per_row);
        col = uint32_t(item - uint64_t(row) * items_per_row);
      }
      uint32_t x = col * VEC;
      pair[u] = (VEC == 2) && (x + 1 < p.w);
      off8[u] = size_t(row) * p.pitch8 + size_t(x) * 8;
      offh[u] = size_t(row) * p.hdr_pitch + size_t(x) * 8;
      offl[u] = size_t(row) * p.ldr_pitch + size_t(x) * 4;
      if (valid[u] && pair[u]) {  // issue every load of the iteration before any math
        if (MODE == MODE_TONEMAP) {
          a[u] = ldg_stream(hdr_in + offh[u]);
        } else {
          a[u] = ldg_stream(pos + off8[u]);
          b[u] = ldg_stream(nrm + off8[u]);
          c[u] = ldg_stream(alb + off8[u]);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (!valid[u]) continue;
      if (pair[u]) process_pair<MODE>(hdr + offh[u], ldr + offl[u], a[u], b[u], c[u], cx, cy, cz);
      else if (MODE == MODE_TONEMAP) process_single<MODE>(hdr_in + offh[u], nullptr, nullptr, hdr + offh[u], ldr + offl[u], cx, cy, cz);
      else process_single<MODE>(pos + off8[u], nrm + off8[u], alb + off8[u], hdr + offh[u], ldr + offl[u], cx, cy, cz);
    }
  }
}

// CTAs of tier_a_kernel<MODE, VEC> one SM holds (asked once per instantiation)
template <int MODE>
static int resident_ctas(bool vec) {
  static int cached[2] = {0, 0};
  int& n = cached[vec ? 1 : 0];
  if (n == 0) {
    int occ = 0;
    const cudaError_t e = vec ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, tier_a_kernel<MODE, 2, true>, kThreads, 0)
                              : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, tier_a_kernel<MODE, 1, false>, kThreads, 0);
    n = (e == cudaSuccess && occ > 0) ? occ : 4;
  }
  return n;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

template <int MODE>
static int launch_tier_a(const pbrcu_ctx* ctx, TierAParams p, uint32_t n_views, cudaStream_t s) {
  if (p.w == 0 || p.h == 0 || n_views == 0) return PBRCU_OK;
  bool vec = true;
  if (MODE == MODE_TONEMAP) {
    vec = aligned16(p.hdr_in) && p.hdr_pitch % 16 == 0;
  } else {
    vec = aligned16(p.pos) && aligned16(p.nrm) && aligned16(p.alb) && p.pitch8 % 16 == 0 &&
          p.view_stride8 % 16 == 0;
    if (MODE != MODE_FUSED_LDR) vec = vec && aligned16(p.hdr) && p.hdr_pitch % 16 == 0 && p.hdr_view_stride % 16 == 0;
  }
  if (MODE != MODE_LIGHT) vec = vec && (reinterpret_cast<uintptr_t>(p.ldr) & 7u) == 0 && p.ldr_pitch % 8 == 0 && p.ldr_view_stride % 8 == 0;
  // Collapse a fully contiguous image into one long row: no per-item division in the kernel.
  bool contiguous = (p.w % 2 == 0);
  if (MODE == MODE_TONEMAP) contiguous = contiguous && p.hdr_pitch == size_t(p.w) * 8;
  else contiguous = contiguous && p.pitch8 == size_t(p.w) * 8 && (MODE == MODE_FUSED_LDR || p.hdr_pitch == size_t(p.w) * 8);
  if (MODE != MODE_LIGHT) contiguous = contiguous && p.ldr_pitch == size_t(p.w) * 4;
  bool flat = false;
  if (contiguous && uint64_t(p.w) * p.h < (1ull << 31)) {
    p.w = p.w * p.h;
    p.h = 1;
    flat = vec;
  }
  const uint32_t vecw = vec ? 2 : 1;
  const uint64_t items = uint64_t((p.w + vecw - 1) / vecw) * p.h;
  const uint64_t per_block = uint64_t(kThreads) * kUnroll;
  uint64_t blocks = (items + per_block - 1) / per_block;
  // persistent grid: exactly the CTAs that are resident at once (32 to 63 registers per thread depending on the
  // mode: 4 to 8 CTAs of 256 threads per SM), so the grid is one full wave with no partial second one
  const uint64_t cap = uint64_t(ctx->sm_count) * resident_ctas<MODE>(vec);
  uint64_t per_view_cap = n_views > 1 ? (cap + n_views - 1) / n_views : cap;
  if (per_view_cap < 1) per_view_cap = 1;
  if (blocks > per_view_cap) blocks = per_view_cap;
  dim3 grid(uint32_t(blocks), n_views, 1);
  if (flat) tier_a_kernel<MODE, 2, true><<<grid, kThreads, 0, s>>>(p);
  else if (vec) tier_a_kernel<MODE, 2, false><<<grid, kThreads, 0, s>>>(p);
  else tier_a_kernel<MODE, 1, false><<<grid, kThreads, 0, s>>>(p);
  return launch_check("tier_a_kernel");
}

static int resolve_rect(uint32_t width, uint32_t height, pbrcu_rect& r) {
  if (r.w == 0 && r.h == 0) { r.x0 = 0; r.y0 = 0; r.w = width; r.h = height; }
  if (uint64_t(r.x0) + r.w > width || uint64_t(r.y0) + r.h > height) {
    set_error("invalid argument: rect (%u,%u,%u,%u) outside %ux%u image", r.x0, r.y0, r.w, r.h, width, height);
    return PBRCU_ERR_INVALID;
  }
  return PBRCU_OK;
}

static int fill_gbuffer_params(const pbrcu_gbuffer* g, const pbrcu_rect& r, TierAParams& p) {
  PBRCU_REQUIRE(g != nullptr, "gbuffer is NULL");
  PBRCU_REQUIRE(g->positions && g->normals && g->albedo, "gbuffer plane is NULL");
  PBRCU_REQUIRE(g->pitch8 >= size_t(g->width) * 8 && g->pitch8 % 8 == 0, "gbuffer pitch8 too small or not a multiple of 8");
  size_t off = size_t(r.y0) * g->pitch8 + size_t(r.x0) * 8;
  p.pos = static_cast<const uint8_t*>(g->positions) + off;
  p.nrm = static_cast<const uint8_t*>(g->normals) + off;
  p.alb = static_cast<const uint8_t*>(g->albedo) + off;
  p.pitch8 = g->pitch8;
  p.w = r.w;
  p.h = r.h;
  return PBRCU_OK;
}

}  // namespace pbrcu

using namespace pbrcu;

extern "C" int pbrcu_lighting_blinnphong(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3],
                                         pbrcu_rect rect, void* out_rgba16f, size_t out_pitch,
                                         pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && g && cam_pos && out_rgba16f, "NULL argument");
  if (int rc = resolve_rect(g->width, g->height, rect)) return rc;
  PBRCU_REQUIRE(out_pitch >= size_t(g->width) * 8 && out_pitch % 8 == 0, "out_pitch too small or not a multiple of 8");
  TierAParams p{};
  if (int rc = fill_gbuffer_params(g, rect, p)) return rc;
  p.hdr = static_cast<uint8_t*>(out_rgba16f) + size_t(rect.y0) * out_pitch + size_t(rect.x0) * 8;
  p.hdr_pitch = out_pitch;
  p.cx = cam_pos[0]; p.cy = cam_pos[1]; p.cz = cam_pos[2];
  return launch_tier_a<MODE_LIGHT>(ctx, p, 1, static_cast<cudaStream_t>(s));
}

extern "C" int pbrcu_tonemap_aces_gamma(pbrcu_ctx* ctx, const void* hdr_rgba16f, size_t hdr_pitch,
                                        uint32_t width, uint32_t height, pbrcu_rect rect,
                                        void* out_rgba8, size_t out_pitch, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && hdr_rgba16f && out_rgba8, "NULL argument");
  if (int rc = resolve_rect(width, height, rect)) return rc;
  PBRCU_REQUIRE(hdr_pitch >= size_t(width) * 8 && hdr_pitch % 8 == 0, "hdr_pitch too small or not a multiple of 8");
  PBRCU_REQUIRE(out_pitch >= size_t(width) * 4 && out_pitch % 4 == 0, "out_pitch too small or not a multiple of 4");
  TierAParams p{};
  p.hdr_in = static_cast<const uint8_t*>(hdr_rgba16f) + size_t(rect.y0) * hdr_pitch + size_t(rect.x0) * 8;
  p.hdr_pitch = hdr_pitch;
  p.ldr = static_cast<uint8_t*>(out_rgba8) + size_t(rect.y0) * out_pitch + size_t(rect.x0) * 4;
  p.ldr_pitch = out_pitch;
  p.w = rect.w; p.h = rect.h;
  return launch_tier_a<MODE_TONEMAP>(ctx, p, 1, static_cast<cudaStream_t>(s));
}

extern "C" int pbrcu_lighting_tonemap_fused(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3],
                                            pbrcu_rect rect, void* out_rgba16f, size_t hdr_pitch,
                                            void* out_rgba8, size_t ldr_pitch, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && g && cam_pos && out_rgba8, "NULL argument");
  if (int rc = resolve_rect(g->width, g->height, rect)) return rc;
  PBRCU_REQUIRE(ldr_pitch >= size_t(g->width) * 4 && ldr_pitch % 4 == 0, "ldr_pitch too small or not a multiple of 4");
  TierAParams p{};
  if (int rc = fill_gbuffer_params(g, rect, p)) return rc;
  p.ldr = static_cast<uint8_t*>(out_rgba8) + size_t(rect.y0) * ldr_pitch + size_t(rect.x0) * 4;
  p.ldr_pitch = ldr_pitch;
  p.cx = cam_pos[0]; p.cy = cam_pos[1]; p.cz = cam_pos[2];
  if (out_rgba16f) {
    PBRCU_REQUIRE(hdr_pitch >= size_t(g->width) * 8 && hdr_pitch % 8 == 0, "hdr_pitch too small or not a multiple of 8");
    p.hdr = static_cast<uint8_t*>(out_rgba16f) + size_t(rect.y0) * hdr_pitch + size_t(rect.x0) * 8;
    p.hdr_pitch = hdr_pitch;
    return launch_tier_a<MODE_FUSED_HDR_LDR>(ctx, p, 1, static_cast<cudaStream_t>(s));
  }
  return launch_tier_a<MODE_FUSED_LDR>(ctx, p, 1, static_cast<cudaStream_t>(s));
}

extern "C" int pbrcu_lighting_blinnphong_batched(pbrcu_ctx* ctx, const pbrcu_gbuffer* g0,
                                                 size_t view_stride8, size_t view_stride4,
                                                 uint32_t n_views, const float* cam_pos_xyz,
                                                 void* out_rgba16f, size_t out_pitch,
                                                 size_t out_view_stride, pbrcu_stream s) {
  (void)view_stride4;
  PBRCU_REQUIRE(ctx && g0 && cam_pos_xyz && out_rgba16f, "NULL argument");
  PBRCU_REQUIRE(n_views >= 1 && n_views <= 65535, "n_views must be in [1, 65535]");
  PBRCU_REQUIRE(out_pitch >= size_t(g0->width) * 8 && out_pitch % 8 == 0, "out_pitch too small or not a multiple of 8");
  PBRCU_REQUIRE(view_stride8 >= g0->pitch8 * g0->height && out_view_stride >= out_pitch * g0->height, "view stride smaller than one view");
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (ctx->cams_cap < n_views) {
    if (ctx->cams_dev) PBRCU_CUDA(cudaFree(ctx->cams_dev));
    ctx->cams_dev = nullptr; ctx->cams_cap = 0;
    PBRCU_CUDA(cudaMalloc(&ctx->cams_dev, size_t(n_views) * 3 * sizeof(float)));
    ctx->cams_cap = n_views;
  }
  PBRCU_CUDA(cudaMemcpyAsync(ctx->cams_dev, cam_pos_xyz, size_t(n_views) * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
  pbrcu_rect rect{0, 0, g0->width, g0->height};
  TierAParams p{};
  if (int rc = fill_gbuffer_params(g0, rect, p)) return rc;
  p.hdr = static_cast<uint8_t*>(out_rgba16f);
  p.hdr_pitch = out_pitch;
  p.cams = ctx->cams_dev;
  p.view_stride8 = view_stride8;
  p.hdr_view_stride = out_view_stride;
  return launch_tier_a<MODE_LIGHT>(ctx, p, n_views, st);
}

// The reference's per-view frame is lighting THEN tonemap (src/gltf_viewer/App.cpp:387-392); over a block of views both passes
// run as ONE launch: blockIdx.y selects the view, the tonemapper consumes the fp16-rounded lighting value, so every view equals the
// two-pass result bit for bit.  out_rgba16f may be NULL (LDR only: 28 B/px instead of 36).
extern "C" int pbrcu_lighting_tonemap_fused_batched(pbrcu_ctx* ctx, const pbrcu_gbuffer* g0, size_t view_stride8, uint32_t n_views,
                                                    const float* cam_pos_xyz, void* out_rgba16f, size_t hdr_pitch, size_t hdr_view_stride,
                                                    void* out_rgba8, size_t ldr_pitch, size_t ldr_view_stride, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && g0 && cam_pos_xyz && out_rgba8, "NULL argument");
  PBRCU_REQUIRE(n_views >= 1 && n_views <= 65535, "n_views must be in [1, 65535]");
  PBRCU_REQUIRE(ldr_pitch >= size_t(g0->width) * 4 && ldr_pitch % 4 == 0, "ldr_pitch too small or not a multiple of 4");
  PBRCU_REQUIRE(view_stride8 >= g0->pitch8 * g0->height && view_stride8 % 8 == 0 && ldr_view_stride >= ldr_pitch * g0->height && ldr_view_stride % 4 == 0,
                "view stride smaller than one view or misaligned");
  if (out_rgba16f)
    PBRCU_REQUIRE(hdr_pitch >= size_t(g0->width) * 8 && hdr_pitch % 8 == 0 && hdr_view_stride >= hdr_pitch * g0->height && hdr_view_stride % 8 == 0,
                  "hdr pitch / view stride too small or misaligned");
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (ctx->cams_cap < n_views) {
    if (ctx->cams_dev) PBRCU_CUDA(cudaFree(ctx->cams_dev));
    ctx->cams_dev = nullptr; ctx->cams_cap = 0;
    PBRCU_CUDA(cudaMalloc(&ctx->cams_dev, size_t(n_views) * 3 * sizeof(float)));
    ctx->cams_cap = n_views;
  }
  PBRCU_CUDA(cudaMemcpyAsync(ctx->cams_dev, cam_pos_xyz, size_t(n_views) * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
  pbrcu_rect rect{0, 0, g0->width, g0->height};
  TierAParams p{};
  if (int rc = fill_gbuffer_params(g0, rect, p)) return rc;
  p.ldr = static_cast<uint8_t*>(out_rgba8);
  p.ldr_pitch = ldr_pitch;
  p.ldr_view_stride = ldr_view_stride;
  p.cams = ctx->cams_dev;
  p.view_stride8 = view_stride8;
  if (out_rgba16f) {
    p.hdr = static_cast<uint8_t*>(out_rgba16f);
    p.hdr_pitch = hdr_pitch;
    p.hdr_view_stride = hdr_view_stride;
    return launch_tier_a<MODE_FUSED_HDR_LDR>(ctx, p, n_views, st);
  }
  return launch_tier_a<MODE_FUSED_LDR>(ctx, p, n_views, st);
}
