// capi.cu -- context, memory, streams, timers, G-buffer / image allocation and the host-buffer
// frame entry point of the C-ABI (include/pbrcu.h).  No CPU fallback anywhere: every compute entry
// point needs a CUDA device and fails with PBRCU_ERR_CUDA otherwise.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>

#include "cube_tex.cuh"

namespace pbrcu {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) in %s: %s", int(e), cudaGetErrorName(e), what, cudaGetErrorString(e));
  return PBRCU_ERR_CUDA;
}

bool is_host_pointer(const void* p) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();  // (pre-CUDA-11 behaviour for unregistered memory)
    return true;
  }
  return a.type == cudaMemoryTypeUnregistered || a.type == cudaMemoryTypeHost;
}

}  // namespace pbrcu

using namespace pbrcu;

extern "C" int pbrcu_version(void) { return PBRCU_VERSION; }
extern "C" const char* pbrcu_last_error(void) { return g_error; }

extern "C" int pbrcu_device_count(int* out) {
  PBRCU_REQUIRE(out, "out is NULL");
  *out = 0;
  PBRCU_CUDA(cudaGetDeviceCount(out));
  return PBRCU_OK;
}

extern "C" int pbrcu_ctx_create(int device, pbrcu_ctx** out) {
  PBRCU_REQUIRE(out, "out is NULL");
  *out = nullptr;
  int n = 0;
  PBRCU_CUDA(cudaGetDeviceCount(&n));
  PBRCU_REQUIRE(device >= 0 && device < n, "device ordinal out of range");
  PBRCU_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  PBRCU_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) {
    set_error("device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major, prop.minor);
    return PBRCU_ERR_CUDA;
  }
  pbrcu_ctx* c = new (std::nothrow) pbrcu_ctx();
  if (!c) return PBRCU_ERR_NOMEM;
  memset(c, 0, sizeof(*c));
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  c->max_smem_optin = int(prop.sharedMemPerBlockOptin);
  *out = c;
  return PBRCU_OK;
}

extern "C" void pbrcu_ctx_destroy(pbrcu_ctx* ctx) {
  if (!ctx) return;
  for (int i = 0; i < 8; ++i)
    if (ctx->stage[i]) cudaFree(ctx->stage[i]);
  if (ctx->cams_dev) cudaFree(ctx->cams_dev);
  if (ctx->lights_dev) cudaFree(ctx->lights_dev);
  for (int i = 0; i < 2; ++i) {
    if (ctx->pipe[i]) cudaStreamDestroy(ctx->pipe[i]);
    if (ctx->pipe_done[i]) cudaEventDestroy(ctx->pipe_done[i]);
  }
  if (ctx->pipe_start) cudaEventDestroy(ctx->pipe_start);
  for (auto& e : ctx->band_ready)
    if (e) cudaEventDestroy(e);
  pbrcu::cube_tex_destroy(&ctx->cube);
  if (ctx->raster_scratch) cudaFree(ctx->raster_scratch);
  for (int k = 0; k < pbrcu::kRasterSlots; ++k) {
    if (ctx->raster_slot_stream[k]) { cudaStreamSynchronize(ctx->raster_slot_stream[k]); cudaStreamDestroy(ctx->raster_slot_stream[k]); }
    if (ctx->raster_slot_done[k]) cudaEventDestroy(ctx->raster_slot_done[k]);
    if (ctx->raster_slot_scratch[k]) cudaFree(ctx->raster_slot_scratch[k]);
  }
  if (ctx->raster_start) cudaEventDestroy(ctx->raster_start);
  if (ctx->cube_stage) cudaFree(ctx->cube_stage);
  if (ctx->cube_mm) cudaFree(ctx->cube_mm);
  delete ctx;
}

extern "C" int pbrcu_ctx_sm_count(const pbrcu_ctx* ctx) { return ctx ? ctx->sm_count : 0; }

extern "C" int pbrcu_malloc(pbrcu_ctx* ctx, size_t bytes, void** out) {
  PBRCU_REQUIRE(ctx && out, "NULL argument");
  *out = nullptr;
  if (bytes == 0) return PBRCU_OK;
  cudaError_t e = cudaMalloc(out, bytes);
  if (e == cudaErrorMemoryAllocation) { cuda_fail(e, "cudaMalloc"); return PBRCU_ERR_NOMEM; }
  PBRCU_CUDA(e);
  return PBRCU_OK;
}
extern "C" int pbrcu_free(pbrcu_ctx* ctx, void* ptr) {
  (void)ctx;
  if (ptr) PBRCU_CUDA(cudaFree(ptr));
  return PBRCU_OK;
}
extern "C" int pbrcu_host_alloc(size_t bytes, void** out) {
  PBRCU_REQUIRE(out, "out is NULL");
  *out = nullptr;
  PBRCU_CUDA(cudaMallocHost(out, bytes ? bytes : 1));
  return PBRCU_OK;
}
extern "C" int pbrcu_host_free(void* ptr) {
  if (ptr) PBRCU_CUDA(cudaFreeHost(ptr));
  return PBRCU_OK;
}
extern "C" int pbrcu_memcpy_h2d(void* dst, const void* src, size_t bytes, pbrcu_stream s) {
  PBRCU_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(s)));
  return PBRCU_OK;
}
extern "C" int pbrcu_memcpy_d2h(void* dst, const void* src, size_t bytes, pbrcu_stream s) {
  PBRCU_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(s)));
  return PBRCU_OK;
}
extern "C" int pbrcu_memset(void* dst, int value, size_t bytes, pbrcu_stream s) {
  PBRCU_CUDA(cudaMemsetAsync(dst, value, bytes, static_cast<cudaStream_t>(s)));
  return PBRCU_OK;
}
extern "C" int pbrcu_memset_2d(void* dst, size_t dst_pitch, int value, size_t width_bytes, size_t height, pbrcu_stream s) {
  PBRCU_REQUIRE(dst_pitch >= width_bytes, "pitch smaller than a row");
  if (width_bytes == 0 || height == 0) return PBRCU_OK;
  PBRCU_CUDA(cudaMemset2DAsync(dst, dst_pitch, value, width_bytes, height, static_cast<cudaStream_t>(s)));
  return PBRCU_OK;
}
extern "C" int pbrcu_memcpy_h2d_2d(void* dst, size_t dst_pitch, const void* src, size_t src_pitch, size_t width_bytes, size_t height,
                                   pbrcu_stream s) {
  PBRCU_REQUIRE(dst_pitch >= width_bytes && src_pitch >= width_bytes, "pitch smaller than a row");
  if (width_bytes == 0 || height == 0) return PBRCU_OK;
  if (dst_pitch == width_bytes && src_pitch == width_bytes)  // tightly packed on both sides: one linear copy
    PBRCU_CUDA(cudaMemcpyAsync(dst, src, width_bytes * height, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(s)));
  else
    PBRCU_CUDA(cudaMemcpy2DAsync(dst, dst_pitch, src, src_pitch, width_bytes, height, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(s)));
  return PBRCU_OK;
}
extern "C" int pbrcu_stream_create(pbrcu_stream* out) {
  PBRCU_REQUIRE(out, "out is NULL");
  cudaStream_t st;
  PBRCU_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  *out = st;
  return PBRCU_OK;
}
extern "C" int pbrcu_stream_destroy(pbrcu_stream s) {
  if (s) PBRCU_CUDA(cudaStreamDestroy(static_cast<cudaStream_t>(s)));
  return PBRCU_OK;
}
extern "C" int pbrcu_stream_synchronize(pbrcu_stream s) {
  PBRCU_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(s)));
  return PBRCU_OK;
}

struct pbrcu_timer { cudaEvent_t a, b; };
extern "C" int pbrcu_timer_create(void** out) {
  PBRCU_REQUIRE(out, "out is NULL");
  pbrcu_timer* t = new (std::nothrow) pbrcu_timer();
  if (!t) return PBRCU_ERR_NOMEM;
  PBRCU_CUDA(cudaEventCreate(&t->a));
  PBRCU_CUDA(cudaEventCreate(&t->b));
  *out = t;
  return PBRCU_OK;
}
extern "C" int pbrcu_timer_destroy(void* tp) {
  pbrcu_timer* t = static_cast<pbrcu_timer*>(tp);
  if (!t) return PBRCU_OK;
  cudaEventDestroy(t->a);
  cudaEventDestroy(t->b);
  delete t;
  return PBRCU_OK;
}
extern "C" int pbrcu_timer_start(void* tp, pbrcu_stream s) {
  PBRCU_REQUIRE(tp, "timer is NULL");
  PBRCU_CUDA(cudaEventRecord(static_cast<pbrcu_timer*>(tp)->a, static_cast<cudaStream_t>(s)));
  return PBRCU_OK;
}
extern "C" int pbrcu_timer_stop(void* tp, pbrcu_stream s) {
  PBRCU_REQUIRE(tp, "timer is NULL");
  PBRCU_CUDA(cudaEventRecord(static_cast<pbrcu_timer*>(tp)->b, static_cast<cudaStream_t>(s)));
  return PBRCU_OK;
}
extern "C" int pbrcu_timer_elapsed_ms(void* tp, float* ms) {
  PBRCU_REQUIRE(tp && ms, "NULL argument");
  pbrcu_timer* t = static_cast<pbrcu_timer*>(tp);
  PBRCU_CUDA(cudaEventSynchronize(t->b));
  PBRCU_CUDA(cudaEventElapsedTime(ms, t->a, t->b));
  return PBRCU_OK;
}

// ---- G-buffer / images -------------------------------------------------------------------------
static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

__global__ void fill_f32_kernel(float* p, size_t n, float v) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) p[i] = v;
}

extern "C" int pbrcu_gbuffer_clear(const pbrcu_gbuffer* g, pbrcu_stream s) {
  PBRCU_REQUIRE(g && g->positions && g->normals && g->albedo, "gbuffer plane is NULL");
  cudaStream_t st = static_cast<cudaStream_t>(s);
  size_t b8 = g->pitch8 * g->height, b4 = g->pitch4 * g->height;
  // PbrRenderSystem.cpp:370-415: colour targets cleared to 0, depth to 1.0
  PBRCU_CUDA(cudaMemsetAsync(g->positions, 0, b8, st));
  PBRCU_CUDA(cudaMemsetAsync(g->normals, 0, b8, st));
  PBRCU_CUDA(cudaMemsetAsync(g->albedo, 0, b8, st));
  if (g->orm) PBRCU_CUDA(cudaMemsetAsync(g->orm, 0, b4, st));
  if (g->emissive) PBRCU_CUDA(cudaMemsetAsync(g->emissive, 0, b8, st));
  if (g->depth) {
    fill_f32_kernel<<<1184, 256, 0, st>>>(g->depth, b4 / 4, 1.0f);
    if (int rc = launch_check("fill_f32_kernel")) return rc;
  }
  return PBRCU_OK;
}

extern "C" int pbrcu_gbuffer_alloc(pbrcu_ctx* ctx, uint32_t w, uint32_t h, int tier, pbrcu_gbuffer* out) {
  PBRCU_REQUIRE(ctx && out, "NULL argument");
  PBRCU_REQUIRE(w > 0 && h > 0, "empty extent");
  PBRCU_REQUIRE(tier == 0 || tier == 1, "tier must be 0 or 1");
  memset(out, 0, sizeof(*out));
  out->width = w;
  out->height = h;
  // rows are 128-byte aligned (a whole number of 128-B lines per row, 16-B vector access always legal);
  // at every named width (1024, 1920, 3840, 7680) this equals the tight pitch.
  out->pitch8 = align_up(size_t(w) * 8, 128);
  out->pitch4 = align_up(size_t(w) * 4, 128);
  size_t b8 = out->pitch8 * h, b4 = out->pitch4 * h;
  int rc;
  if ((rc = pbrcu_malloc(ctx, b8, &out->positions)) || (rc = pbrcu_malloc(ctx, b8, &out->normals)) ||
      (rc = pbrcu_malloc(ctx, b8, &out->albedo)) || (rc = pbrcu_malloc(ctx, b4, reinterpret_cast<void**>(&out->depth))) ||
      (tier == 1 && ((rc = pbrcu_malloc(ctx, b4, &out->orm)) || (rc = pbrcu_malloc(ctx, b8, &out->emissive))))) {
    pbrcu_gbuffer_free(ctx, out);
    return rc;
  }
  // the clear must have landed before the call returns: streams from pbrcu_stream_create are non-blocking and do not
  // order themselves against the legacy stream, so a clear still in flight could wipe the caller's first upload or pass
  if ((rc = pbrcu_gbuffer_clear(out, nullptr))) return rc;
  PBRCU_CUDA(cudaStreamSynchronize(nullptr));
  return PBRCU_OK;
}

extern "C" int pbrcu_gbuffer_free(pbrcu_ctx* ctx, pbrcu_gbuffer* g) {
  (void)ctx;
  if (!g) return PBRCU_OK;
  void* planes[6] = {g->positions, g->normals, g->albedo, g->depth, g->orm, g->emissive};
  for (void* p : planes)
    if (p) cudaFree(p);
  memset(g, 0, sizeof(*g));
  return PBRCU_OK;
}

extern "C" int pbrcu_image_alloc(pbrcu_ctx* ctx, uint32_t w, uint32_t h, uint32_t texel_bytes, void** out,
                                 size_t* pitch) {
  PBRCU_REQUIRE(ctx && out && pitch, "NULL argument");
  PBRCU_REQUIRE(w > 0 && h > 0, "empty extent");
  PBRCU_REQUIRE(texel_bytes == 4 || texel_bytes == 8 || texel_bytes == 16, "texel_bytes must be 4, 8 or 16");
  *pitch = align_up(size_t(w) * texel_bytes, 128);
  int rc = pbrcu_malloc(ctx, *pitch * h, out);
  if (rc) return rc;
  PBRCU_CUDA(cudaMemsetAsync(*out, 0, *pitch * h, nullptr));  // render targets start cleared to 0 ...
  PBRCU_CUDA(cudaStreamSynchronize(nullptr));                 // ... before the caller's (non-blocking) streams touch them
  return PBRCU_OK;
}

// ---- one frame through host buffers (end-to-end path) ------------------------------------------
static int ensure_stage(pbrcu_ctx* ctx, int i, size_t bytes) {
  if (ctx->stage_bytes[i] >= bytes) return PBRCU_OK;
  if (ctx->stage[i]) PBRCU_CUDA(cudaFree(ctx->stage[i]));
  ctx->stage[i] = nullptr;
  ctx->stage_bytes[i] = 0;
  PBRCU_CUDA(cudaMalloc(&ctx->stage[i], bytes));
  ctx->stage_bytes[i] = bytes;
  return PBRCU_OK;
}

// ---- one GGX+IBL frame through host buffers, pipelined in row bands ---------------------------
static int ensure_pipe(pbrcu_ctx* ctx) {
  if (ctx->pipe[0]) return PBRCU_OK;
  for (int i = 0; i < 2; ++i) {
    PBRCU_CUDA(cudaStreamCreateWithFlags(&ctx->pipe[i], cudaStreamNonBlocking));
    PBRCU_CUDA(cudaEventCreateWithFlags(&ctx->pipe_done[i], cudaEventDisableTiming));
  }
  PBRCU_CUDA(cudaEventCreateWithFlags(&ctx->pipe_start, cudaEventDisableTiming));
  for (auto& e : ctx->band_ready) PBRCU_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  return PBRCU_OK;
}

// First row of band b of n: equal bands, except that the last three shrink by halves (8, .., 8, 4, 2, 1 units).  What is left
// to do after the frame's last byte has arrived -- shading and reading back the final band -- is then a forty-seventh of the
// frame at n = 8, not an eighth, for the same number of copies.
static uint32_t host_band_row(uint32_t height, uint32_t b, uint32_t n) {
  if (n < 4) return uint32_t(uint64_t(height) * b / n);
  const uint32_t body = 8 * (n - 3), units = body + 7;
  const uint32_t u = b <= n - 3 ? 8 * b : b == n - 2 ? body + 4 : b == n - 1 ? body + 6 : units;
  return uint32_t(uint64_t(height) * u / units);
}

// The reference's frame (lighting.glsl then aces+gamma.glsl) through host buffers.  Frames of half a megapixel or more are cut
// into 4 tapered row bands and pipelined like pbrcu_shade_frame_host: copies-in back to back on one stream, the fused
// K1+K2 launch and the read-back of each band on the other.  Smaller frames (the viewer's) go in one piece on the caller's stream.
extern "C" int pbrcu_frame_host(pbrcu_ctx* ctx, const void* positions_host, const void* normals_host,
                                const void* albedo_host, uint32_t width, uint32_t height,
                                const float cam_pos[3], void* hdr_out_host, void* ldr_out_host,
                                pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && positions_host && normals_host && albedo_host && cam_pos, "NULL argument");
  PBRCU_REQUIRE(width > 0 && height > 0, "empty extent");
  PBRCU_REQUIRE(hdr_out_host || ldr_out_host, "no output requested");
  cudaStream_t st = static_cast<cudaStream_t>(s);
  const size_t p8 = size_t(width) * 8, p4 = size_t(width) * 4;
  int rc;
  for (int i = 0; i < 3; ++i)
    if ((rc = ensure_stage(ctx, i, p8 * height))) return rc;
  if (hdr_out_host && (rc = ensure_stage(ctx, 3, p8 * height))) return rc;
  if (ldr_out_host && (rc = ensure_stage(ctx, 4, p4 * height))) return rc;
  const bool piped = uint64_t(width) * height >= (1u << 19) && height >= 4;
  const uint32_t n_bands = piped ? 4 : 1;
  cudaStream_t in = st, work = st;
  if (piped) {
    if ((rc = ensure_pipe(ctx))) return rc;
    in = ctx->pipe[0]; work = ctx->pipe[1];
    PBRCU_CUDA(cudaEventRecord(ctx->pipe_start, st));
    for (int i = 0; i < 2; ++i) PBRCU_CUDA(cudaStreamWaitEvent(ctx->pipe[i], ctx->pipe_start, 0));
  }
  pbrcu_gbuffer g{};
  g.positions = ctx->stage[0]; g.normals = ctx->stage[1]; g.albedo = ctx->stage[2];
  g.width = width; g.height = height; g.pitch8 = p8; g.pitch4 = p4;
  const void* src[3] = {positions_host, normals_host, albedo_host};
  for (uint32_t b = 0; b < n_bands; ++b) {
    const uint32_t y0 = host_band_row(height, b, n_bands), y1 = host_band_row(height, b + 1, n_bands);
    if (y1 == y0) continue;
    for (int i = 0; i < 3; ++i)
      PBRCU_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(ctx->stage[i]) + y0 * p8, static_cast<const uint8_t*>(src[i]) + y0 * p8,
                                 size_t(y1 - y0) * p8, cudaMemcpyHostToDevice, in));
    if (piped) {
      PBRCU_CUDA(cudaEventRecord(ctx->band_ready[b], in));
      PBRCU_CUDA(cudaStreamWaitEvent(work, ctx->band_ready[b], 0));
    }
    const pbrcu_rect band{0, y0, width, y1 - y0};
    if (ldr_out_host)
      rc = pbrcu_lighting_tonemap_fused(ctx, &g, cam_pos, band, hdr_out_host ? ctx->stage[3] : nullptr, p8, ctx->stage[4], p4, work);
    else
      rc = pbrcu_lighting_blinnphong(ctx, &g, cam_pos, band, ctx->stage[3], p8, work);
    if (rc) return rc;
    if (hdr_out_host)
      PBRCU_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(hdr_out_host) + y0 * p8, static_cast<uint8_t*>(ctx->stage[3]) + y0 * p8,
                                 size_t(y1 - y0) * p8, cudaMemcpyDeviceToHost, work));
    if (ldr_out_host)
      PBRCU_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(ldr_out_host) + y0 * p4, static_cast<uint8_t*>(ctx->stage[4]) + y0 * p4,
                                 size_t(y1 - y0) * p4, cudaMemcpyDeviceToHost, work));
  }
  if (piped)
    for (int i = 0; i < 2; ++i) {
      PBRCU_CUDA(cudaEventRecord(ctx->pipe_done[i], ctx->pipe[i]));
      PBRCU_CUDA(cudaStreamWaitEvent(st, ctx->pipe_done[i], 0));
    }
  PBRCU_CUDA(cudaStreamSynchronize(st));
  return PBRCU_OK;
}

extern "C" int pbrcu_shade_frame_host(pbrcu_ctx* ctx, const void* positions_host, const void* normals_host,
                                      const void* albedo_host, const void* orm_host, const void* emissive_host,
                                      uint32_t width, uint32_t height, const float cam_pos[3],
                                      const pbrcu_light* lights_host, uint32_t n_lights, const pbrcu_ibl* ibl,
                                      void* hdr_out_host, uint32_t n_bands, pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && positions_host && normals_host && albedo_host && orm_host && emissive_host && cam_pos && hdr_out_host, "NULL argument");
  PBRCU_REQUIRE(width > 0 && height > 0, "empty extent");
  PBRCU_REQUIRE(n_lights == 0 || lights_host, "lights_host is NULL");
  if (n_bands == 0) n_bands = 4;
  if (n_bands > 64) n_bands = 64;
  if (n_bands > height) n_bands = height;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  const size_t p8 = size_t(width) * 8, p4 = size_t(width) * 4;
  int rc;
  const size_t need[6] = {p8 * height, p8 * height, p8 * height, p4 * height, p8 * height, p8 * height};
  for (int i = 0; i < 6; ++i)
    if ((rc = ensure_stage(ctx, i, need[i]))) return rc;
  if ((rc = ensure_pipe(ctx))) return rc;
  PBRCU_CUDA(cudaEventRecord(ctx->pipe_start, st));
  for (int i = 0; i < 2; ++i) PBRCU_CUDA(cudaStreamWaitEvent(ctx->pipe[i], ctx->pipe_start, 0));
  pbrcu_gbuffer g{};
  g.positions = ctx->stage[0]; g.normals = ctx->stage[1]; g.albedo = ctx->stage[2];
  g.orm = ctx->stage[3]; g.emissive = ctx->stage[4];
  g.width = width; g.height = height; g.pitch8 = p8; g.pitch4 = p4;
  const void* src[5] = {positions_host, normals_host, albedo_host, orm_host, emissive_host};
  // pipe[0] carries every host-to-device copy, band after band; pipe[1] shades a band as soon as its planes are in and
  // sends the result back.  Copies-in thus run back to back at the full PCIe rate (issued on two streams they would
  // share it and every band would arrive late), copies-out overlap them on the other direction of the link.
  cudaStream_t in = ctx->pipe[0], work = ctx->pipe[1];
  for (uint32_t b = 0; b < n_bands; ++b) {
    const uint32_t y0 = host_band_row(height, b, n_bands), y1 = host_band_row(height, b + 1, n_bands);
    if (y1 == y0) continue;
    for (int i = 0; i < 5; ++i) {
      const size_t pitch = (i == 3) ? p4 : p8;
      PBRCU_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(ctx->stage[i]) + y0 * pitch, static_cast<const uint8_t*>(src[i]) + y0 * pitch,
                                 size_t(y1 - y0) * pitch, cudaMemcpyHostToDevice, in));
    }
    PBRCU_CUDA(cudaEventRecord(ctx->band_ready[b], in));
    PBRCU_CUDA(cudaStreamWaitEvent(work, ctx->band_ready[b], 0));
    if ((rc = pbrcu_shade_ggx_ibl(ctx, &g, cam_pos, lights_host, n_lights, nullptr, ibl, pbrcu_rect{0, y0, width, y1 - y0},
                                  ctx->stage[5], p8, work)))
      return rc;
    PBRCU_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(hdr_out_host) + y0 * p8, static_cast<uint8_t*>(ctx->stage[5]) + y0 * p8,
                               size_t(y1 - y0) * p8, cudaMemcpyDeviceToHost, work));
  }
  for (int i = 0; i < 2; ++i) {
    PBRCU_CUDA(cudaEventRecord(ctx->pipe_done[i], ctx->pipe[i]));
    PBRCU_CUDA(cudaStreamWaitEvent(st, ctx->pipe_done[i], 0));
  }
  PBRCU_CUDA(cudaStreamSynchronize(st));
  return PBRCU_OK;
}
