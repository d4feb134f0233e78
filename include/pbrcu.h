/* pbrcu.h -- the C-ABI boundary of the B200-native shading hot path (libpbrcu.so).
 *
 * The reference (WeaponizedSchizophrenia/physically-based-renderer) has no FFI layer: the path sits
 * behind two C++ classes, pbr::PbrRenderSystem and pbr::TonemapperSystem, that RECORD Vulkan commands
 * which run GLSL shaders.  Each entry point below replaces one recorded pass; the C++ facade in
 * include/pbr/ (same class / method names as the reference) is a thin wrapper over this header.
 *
 *   pbrcu_lighting_blinnphong   <- PbrRenderSystem::recordLightingPass  (PbrRenderSystem.cpp:466-501)
 *                                  running assets/shaders/pbr/lighting.glsl:19-29
 *   pbrcu_tonemap_aces_gamma    <- TonemapperSystem::run                (TonemapperSystem.cpp:97-134)
 *                                  running assets/shaders/tonemappers/aces+gamma.glsl:30-40
 *   pbrcu_lighting_tonemap_fused<- the two above back to back (App.cpp:387-392) without the HDR trip
 *   pbrcu_gbuffer_alloc/free    <- PbrRenderSystem::allocateGBuffer     (PbrRenderSystem.cpp:342-355,
 *                                  GBuffer.cpp:41-95; formats GBuffer.hpp:13-16)
 *   pbrcu_image_alloc/free      <- TonemapperSystem::allocateHdrImage   (TonemapperSystem.cpp:87-95)
 *   pbrcu_frame_host            <- one App::renderAndPresent iteration (App.cpp:444-483) with host
 *                                  buffers: upload, shade, tonemap, read back
 * Tier B (north-star extension; the reference has no counterpart -- SURVEY.md section 0):
 *   pbrcu_shade_ggx_ibl, pbrcu_cluster_lights, pbrcu_ibl_* (K3..K9)
 *
 * Conventions: every function is extern "C", returns 0 on success or a pbrcu_status (never throws);
 * pbrcu_last_error() gives a thread-local message.  Device work is ENQUEUED on the given stream
 * (a cudaStream_t passed as void*; NULL = the legacy default stream) and is asynchronous.  Buffers
 * are caller-owned device memory unless the name ends in _host.  There is no CPU fallback: without
 * a CUDA device every compute entry point fails with PBRCU_ERR_CUDA.
 *
 * Ordering rules.  (1) The allocation calls (pbrcu_gbuffer_alloc, pbrcu_image_alloc) return with their
 * clears COMPLETE, so the memory may be used on any stream at once.  (2) One stream at a time per
 * pbrcu_ctx for the entry points that use scratch memory owned by the context: pbrcu_geometry_pass and
 * pbrcu_geometry_pass_batched (visibility buffer -- each pass leaves it empty for the next one, which relies on
 * that --, vertex cache, draw tables), pbrcu_ibl_prefilter_ggx / pbrcu_ibl_irradiance (the texture copy of
 * the source chain), the *_batched entries (camera table) and the *_host entries (staging planes).  Calls
 * of these on DIFFERENT streams of one context must be ordered by the caller (an event), or use one
 * context per stream -- contexts are cheap (no device memory until a scratch is first needed).
 */
#ifndef PBRCU_H
#define PBRCU_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built with -fvisibility=hidden */
#endif

#define PBRCU_VERSION 100

typedef enum {
  PBRCU_OK = 0,
  PBRCU_ERR_INVALID = 1, /* bad argument (NULL plane, misaligned pitch, empty rect outside image...) */
  PBRCU_ERR_CUDA = 2,    /* a CUDA runtime call failed; see pbrcu_last_error() */
  PBRCU_ERR_NOMEM = 3
} pbrcu_status;

typedef struct pbrcu_ctx pbrcu_ctx; /* device ordinal + SM count + scratch */
typedef void* pbrcu_stream;         /* cudaStream_t */

/* The reference G-buffer (GBuffer.hpp:13-16) as row-major device planes of 8-byte texels, plus the
 * two Tier-B planes.  pitch8 = bytes per row of the 8-byte planes, pitch4 of the 4-byte planes. */
typedef struct {
  void* positions; /* R16G16B16A16_SFLOAT  world xyz, w = 0 */
  void* normals;   /* R16G16B16A16_SFLOAT  unit world normal, w = 0; all-zero = cleared texel */
  void* albedo;    /* R16G16B16A16_UNORM   base colour rgba */
  float* depth;    /* D32_SFLOAT, optional (the lighting shader reads it into a dead variable) */
  void* orm;       /* Tier B: R8G8B8A8_UNORM  R occlusion, G roughness, B metallic */
  void* emissive;  /* Tier B: R16G16B16A16_SFLOAT */
  uint32_t width, height;
  size_t pitch8, pitch4;
} pbrcu_gbuffer;

typedef struct { uint32_t x0, y0, w, h; } pbrcu_rect; /* tile to shade; w = h = 0 means whole image */

/* ---- context, memory, streams -------------------------------------------------------------- */
int pbrcu_version(void);
const char* pbrcu_last_error(void);
int pbrcu_device_count(int* out);
int pbrcu_ctx_create(int device, pbrcu_ctx** out);
void pbrcu_ctx_destroy(pbrcu_ctx* ctx);
int pbrcu_ctx_sm_count(const pbrcu_ctx* ctx);
int pbrcu_malloc(pbrcu_ctx* ctx, size_t bytes, void** out);
int pbrcu_free(pbrcu_ctx* ctx, void* ptr);
int pbrcu_host_alloc(size_t bytes, void** out); /* pinned */
int pbrcu_host_free(void* ptr);
int pbrcu_memcpy_h2d(void* dst, const void* src, size_t bytes, pbrcu_stream s);
int pbrcu_memcpy_d2h(void* dst, const void* src, size_t bytes, pbrcu_stream s);
int pbrcu_memset(void* dst, int value, size_t bytes, pbrcu_stream s);
/* pitched forms: `height` rows of `width_bytes` bytes, rows dst_pitch (src_pitch) bytes apart -- one call per image */
int pbrcu_memset_2d(void* dst, size_t dst_pitch, int value, size_t width_bytes, size_t height, pbrcu_stream s);
int pbrcu_memcpy_h2d_2d(void* dst, size_t dst_pitch, const void* src, size_t src_pitch, size_t width_bytes,
                        size_t height, pbrcu_stream s);
int pbrcu_stream_create(pbrcu_stream* out);
int pbrcu_stream_destroy(pbrcu_stream s);
int pbrcu_stream_synchronize(pbrcu_stream s);
/* CUDA-event timing on stream s around whatever the caller enqueues between start and stop */
int pbrcu_timer_create(void** out);
int pbrcu_timer_destroy(void* t);
int pbrcu_timer_start(void* t, pbrcu_stream s);
int pbrcu_timer_stop(void* t, pbrcu_stream s);
int pbrcu_timer_elapsed_ms(void* t, float* ms); /* synchronises on the stop event */

/* tier 0: reference planes only; tier 1: + orm + emissive.  Cleared like the reference
 * (PbrRenderSystem.cpp:370-415): colour planes 0, depth 1.0. */
int pbrcu_gbuffer_alloc(pbrcu_ctx* ctx, uint32_t w, uint32_t h, int tier, pbrcu_gbuffer* out);
int pbrcu_gbuffer_free(pbrcu_ctx* ctx, pbrcu_gbuffer* g);
int pbrcu_gbuffer_clear(const pbrcu_gbuffer* g, pbrcu_stream s);
/* a pitched image of `texel_bytes`-byte texels (8: RGBA16F HDR target, 4: RGBA8 LDR output) */
int pbrcu_image_alloc(pbrcu_ctx* ctx, uint32_t w, uint32_t h, uint32_t texel_bytes, void** out,
                      size_t* pitch);

/* ---- Tier A: the reference's passes -------------------------------------------------------- */
/* K1: out = albedo * max(N.V,0) + max(N.H,0)^64 with L = V = normalize(cam - P), RGBA16F out. */
int pbrcu_lighting_blinnphong(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3],
                              pbrcu_rect rect, void* out_rgba16f, size_t out_pitch, pbrcu_stream s);
/* K2: RGBA16F -> ACES fit (matrices column-major as written) -> gamma 2.2 -> clamp -> RGBA8. */
int pbrcu_tonemap_aces_gamma(pbrcu_ctx* ctx, const void* hdr_rgba16f, size_t hdr_pitch,
                             uint32_t width, uint32_t height, pbrcu_rect rect, void* out_rgba8,
                             size_t out_pitch, pbrcu_stream s);
/* K1+K2 in one pass; the tonemapper sees the fp16-ROUNDED lighting value exactly as it would after
 * the HDR round trip.  out_rgba16f may be NULL (LDR only: 28 B/px instead of 36). */
int pbrcu_lighting_tonemap_fused(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3],
                                 pbrcu_rect rect, void* out_rgba16f, size_t hdr_pitch,
                                 void* out_rgba8, size_t ldr_pitch, pbrcu_stream s);
/* Batched multi-view (BASELINE config 5): n_views G-buffers of identical shape laid out
 * `view_stride8` / `view_stride4` bytes apart starting at g[0]'s planes; one camera per view. */
int pbrcu_lighting_blinnphong_batched(pbrcu_ctx* ctx, const pbrcu_gbuffer* g0, size_t view_stride8,
                                      size_t view_stride4, uint32_t n_views,
                                      const float* cam_pos_xyz /* host, n_views*3 */,
                                      void* out_rgba16f, size_t out_pitch, size_t out_view_stride,
                                      pbrcu_stream s);
/* The batched form of pbrcu_lighting_tonemap_fused: the reference's per-view frame is lighting then tonemap
 * (src/gltf_viewer/App.cpp:387-392), here one launch for the whole block of views.  View v reads its planes view_stride8
 * bytes after view v-1's and writes its LDR frame ldr_view_stride bytes (and, when out_rgba16f is not NULL, its HDR frame
 * hdr_view_stride bytes) after view v-1's.  Every view equals the two-pass result bit for bit. */
int pbrcu_lighting_tonemap_fused_batched(pbrcu_ctx* ctx, const pbrcu_gbuffer* g0, size_t view_stride8, uint32_t n_views,
                                         const float* cam_pos_xyz /* host, n_views*3 */, void* out_rgba16f, size_t hdr_pitch,
                                         size_t hdr_view_stride, void* out_rgba8, size_t ldr_pitch, size_t ldr_view_stride,
                                         pbrcu_stream s);
/* One frame through HOST buffers (the end-to-end path): H2D of the three planes, K1 (+K2), D2H of
 * the HDR and/or LDR result.  Host pointers should be pinned (pbrcu_host_alloc) for full PCIe rate.
 * hdr_out_host / ldr_out_host may each be NULL.  Synchronises the stream before returning. */
int pbrcu_frame_host(pbrcu_ctx* ctx, const void* positions_host, const void* normals_host,
                     const void* albedo_host, uint32_t width, uint32_t height,
                     const float cam_pos[3], void* hdr_out_host, void* ldr_out_host, pbrcu_stream s);

/* ---- Tier B: GGX + point lights + IBL (no reference counterpart) ---------------------------- */
typedef struct {
  float px, py, pz, range; /* range <= 0: unlimited */
  float r, g, b, intensity;
} pbrcu_light;

/* cube chain layout: for mip: for face (+X,-X,+Y,-Y,+Z,-Z): row-major (res>>mip)^2 RGBA16F */
typedef struct pbrcu_ibl_accel pbrcu_ibl_accel; /* footprint-record copies of the three textures */
typedef struct {
  const void* irradiance; uint32_t irr_res;                       /* 6*irr_res^2 RGBA16F (E/pi) */
  const void* prefiltered; uint32_t pref_res; uint32_t pref_mips; /* GGX-prefiltered chain */
  const float* lut; uint32_t lut_res;                             /* split-sum RG32F */
  /* optional (NULL = sample the plain textures): built once by pbrcu_ibl_accel_build.  Every texel is
   * stored with its right / lower / diagonal neighbour (clamped at the face edge) as one 32-byte
   * record, so a bilinear tap is ONE 256-bit load instead of four 8-byte loads -- 4x the texture
   * memory (1.07 GB for a 2048^2 12-mip chain: sized for 180 GB of HBM) for a quarter of the L1
   * wavefronts and no address clamping in the shading kernel.  Results are bit-identical. */
  const pbrcu_ibl_accel* accel;
  /* per-call filtering of the ambient term (K3), OR-ed with the context's defaults (pbrcu_ibl_set_exact_filtering /
   * pbrcu_shade_set_texture_unit): 0 = default (prefiltered cube with fp32 weights, irradiance + LUT through the texture unit) */
  uint32_t flags;
} pbrcu_ibl;
#define PBRCU_IBL_EXACT 1u        /* fp32 weights for all four taps (bit-identical to the plain four-load filter) */
#define PBRCU_IBL_TEXTURE_UNIT 2u /* prefiltered + irradiance cubes through the texture unit, records where the contrast flag is set */
#define PBRCU_IBL_GATHER 4u       /* pbrcu_ibl_accel_build: also build the texture-gather atlases; K3 (host light list): read the bilinear
                                   * footprints with tld4 gathers instead of record loads -- same fp32 weights, bit-identical; measured
                                   * slower on incoherent materials (218 vs 159 us per 4K frame), for content with coherent ones */

/* tile light lists built by pbrcu_cluster_lights: tiles of PBRCU_TILE x PBRCU_TILE pixels */
#define PBRCU_TILE 16
typedef struct {
  uint32_t* counts;  /* tiles_x*tiles_y */
  uint16_t* indices; /* tiles_x*tiles_y*max_per_tile, ascending light index */
  uint32_t tiles_x, tiles_y, max_per_tile;
} pbrcu_light_tiles;

size_t pbrcu_cube_chain_texels(uint32_t base_res, uint32_t nmips);
size_t pbrcu_cube_mip_offset(uint32_t base_res, uint32_t mip); /* texels */

/* K3: lights = array of n_lights pbrcu_light in DEVICE or HOST memory; tiles may be NULL (every pixel loops all
 * lights); ibl may be NULL (direct light + emissive only).  A HOST list of at most 32 lights without tile lists -- the
 * usual frame: a handful of scene lights -- is read during the call and travels in the kernel's parameter block (no
 * upload, no staging: the fastest path, csrc/shade_ggx_cl.cu); any other host list is uploaded into context scratch
 * first.  The same holds for the _replicated, _multicast and _batched forms. */
int pbrcu_shade_ggx_ibl(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3],
                        const pbrcu_light* lights_dev, uint32_t n_lights,
                        const pbrcu_light_tiles* tiles, const pbrcu_ibl* ibl, pbrcu_rect rect,
                        void* out_rgba16f, size_t out_pitch, pbrcu_stream s);
/* K3 with the gather fused into its stores (multi-GPU row bands, north_star: "NCCL over NVLink is used only to
 * gather the finished tiles" -- here the finished texels leave the shading kernel already gathered).  Same
 * arguments and results as pbrcu_shade_ggx_ibl; in addition every texel of rect is stored at the same byte
 * offset (y * out_pitch + x * 8) of each of the n_replicas buffers in replica_outs (a HOST array of device
 * pointers, 8-byte aligned).  A replica is the same-shaped frame of another GPU mapped into this process
 * (cudaIpcOpenMemHandle, CUDA VMM / torch symmetric memory, or peer access in a one-process setup): its stores
 * travel over NVLink while the kernel shades the following texels, so a rank that shades rows [y0, y1) into its own
 * frame and into its peers' frames leaves the whole frame on every GPU with no separate exchange pass.  The caller
 * orders frames: a barrier across the ranks after the launches complete and before anyone reads (or overwrites)
 * a frame.  The reference renders on one GPU (`core/GpuHandle.cpp:77-81`) and has no counterpart. */
#define PBRCU_MAX_REPLICAS 15
int pbrcu_shade_ggx_ibl_replicated(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3],
                                   const pbrcu_light* lights_dev, uint32_t n_lights,
                                   const pbrcu_light_tiles* tiles, const pbrcu_ibl* ibl, pbrcu_rect rect,
                                   void* out_rgba16f, size_t out_pitch, void* const* replica_outs,
                                   uint32_t n_replicas, pbrcu_stream s);
/* The same gather with ONE store per texel: out_multicast is a multicast (multimem) address -- a CUDA VMM multicast
 * object (cuMulticastCreate / cuMulticastBindMem; torch symmetric memory's multicast_ptr) bound to the same-shaped
 * frame on every GPU of the team, this one included.  K3 stores each finished texel with `multimem.st` and the
 * NVSwitch writes it into every bound frame, so a rank's NVLink egress is its own band once, whatever the team size.
 * Needs NVSwitch multicast (NVLS); results are read through the ordinary per-GPU address of the frame after the
 * cross-rank barrier.  Same arguments otherwise as pbrcu_shade_ggx_ibl. */
int pbrcu_shade_ggx_ibl_multicast(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3],
                                  const pbrcu_light* lights_dev, uint32_t n_lights,
                                  const pbrcu_light_tiles* tiles, const pbrcu_ibl* ibl, pbrcu_rect rect,
                                  void* out_multicast, size_t out_pitch, pbrcu_stream s);
/* K3 with the reference's tonemap pass fused into its stores (K2: TonemapperSystem::run, TonemapperSystem.cpp:97-134 running
 * assets/shaders/tonemappers/aces+gamma.glsl:30-40): the display-ready RGBA8 frame in ONE pass, bit-identical to pbrcu_shade_ggx_ibl followed by
 * pbrcu_tonemap_aces_gamma (the tonemapper sees the fp16-rounded lighting value either way), 40 B/px of traffic instead of 44 + 12.
 * out_pitch counts bytes of 4-byte texels.  replica_outs / n_replicas as in pbrcu_shade_ggx_ibl_replicated (NULL, 0 = none): a row band
 * delivered to every peer's LDR frame at half the bytes of the HDR gather. */
int pbrcu_shade_ggx_ibl_ldr(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float cam_pos[3], const pbrcu_light* lights,
                            uint32_t n_lights, const pbrcu_light_tiles* tiles, const pbrcu_ibl* ibl, pbrcu_rect rect,
                            void* out_rgba8, size_t out_pitch, void* const* replica_outs, uint32_t n_replicas, pbrcu_stream s);
/* One GGX+IBL frame through HOST buffers (the end-to-end path of the headline metric): the frame is
 * cut into n_bands row bands (0 = 4; the last three shrink by halves so little is left to do after
 * the last byte has arrived) pipelined over two internal streams, so the H2D copy of band i+1 overlaps the
 * shading and D2H read-back of band i.  lights_host is a host array; ibl textures
 * are device-resident (precomputed once).  Host planes should be pinned.  Synchronises before return. */
int pbrcu_shade_frame_host(pbrcu_ctx* ctx, const void* positions_host, const void* normals_host,
                           const void* albedo_host, const void* orm_host, const void* emissive_host,
                           uint32_t width, uint32_t height, const float cam_pos[3],
                           const pbrcu_light* lights_host, uint32_t n_lights, const pbrcu_ibl* ibl,
                           void* hdr_out_host, uint32_t n_bands, pbrcu_stream s);
/* Builds / frees the footprint-record copies of ibl's textures (ibl->accel is ignored on input).  The
 * records are a snapshot: rebuild after the textures change. */
int pbrcu_ibl_accel_build(pbrcu_ctx* ctx, const pbrcu_ibl* ibl, pbrcu_ibl_accel** out, pbrcu_stream s);
int pbrcu_ibl_accel_free(pbrcu_ctx* ctx, pbrcu_ibl_accel* accel);
size_t pbrcu_ibl_accel_bytes(const pbrcu_ibl_accel* accel);
/* K3 batched multi-view (BASELINE config 5): n_views extended G-buffers of identical shape laid out view_stride8 /
 * view_stride4 bytes apart starting at g0's planes, one camera per view (host array, n_views*3), the same lights
 * (at most 128, no tile lists) and IBL for all; view v's frame lands out_view_stride bytes after view v-1's.
 * One launch: the persistent CTAs stride over the work items of all views. */
int pbrcu_shade_ggx_ibl_batched(pbrcu_ctx* ctx, const pbrcu_gbuffer* g0, size_t view_stride8, size_t view_stride4,
                                uint32_t n_views, const float* cam_pos_xyz, const pbrcu_light* lights_dev,
                                uint32_t n_lights, const pbrcu_ibl* ibl, void* out_rgba16f, size_t out_pitch,
                                size_t out_view_stride, pbrcu_stream s);
/* pbrcu_shade_ggx_ibl_batched for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments and
 * bit-identical results; the per-view camera table lives in a file-scope device array instead of context scratch, so the
 * call makes no cudaMalloc / cudaFree and calls no allocating library routine.  Limits: n_views <= PBRCU_STATIC_MAX_VIEWS
 * (beyond it: PBRCU_ERR_INVALID with a message), lights_dev must be DEVICE memory (a host list would need an upload
 * buffer), at most 128 lights, no tile lists.  The array (48 KB) is held by every GPU that loads the library until exit and
 * is shared by ALL calls on a device: each call overwrites the entries it uses, and calls that overlap on the GPU
 * (different streams or contexts of one device) race on it -- order them with events or use the allocating entry point. */
#define PBRCU_STATIC_MAX_VIEWS 4096
int pbrcu_shade_ggx_ibl_batched_static(pbrcu_ctx* ctx, const pbrcu_gbuffer* g0, size_t view_stride8, size_t view_stride4,
                                       uint32_t n_views, const float* cam_pos_xyz, const pbrcu_light* lights_dev,
                                       uint32_t n_lights, const pbrcu_ibl* ibl, void* out_rgba16f, size_t out_pitch,
                                       size_t out_view_stride, pbrcu_stream s);
/* K3c: per 16x16 tile, the lights whose range sphere meets the tile's world-space position AABB */
int pbrcu_light_tiles_alloc(pbrcu_ctx* ctx, uint32_t width, uint32_t height, uint32_t max_per_tile,
                            pbrcu_light_tiles* out);
int pbrcu_light_tiles_free(pbrcu_ctx* ctx, pbrcu_light_tiles* t);
int pbrcu_cluster_lights(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const pbrcu_light* lights_dev,
                         uint32_t n_lights, const pbrcu_light_tiles* tiles, pbrcu_stream s);

/* K9: xi_i = (i/n, radical_inverse(i)) as float2, bit-exact vs the oracle */
int pbrcu_ibl_hammersley(pbrcu_ctx* ctx, uint32_t n, float* out_xy_dev, pbrcu_stream s);
/* K8: split-sum LUT, res*res RG32F, bit-exact vs the oracle */
int pbrcu_ibl_brdf_lut(pbrcu_ctx* ctx, uint32_t res, uint32_t nsamples, float* out_rg_dev,
                       pbrcu_stream s);
/* K4: lat-long RGBA32F -> cube base level (6*res^2 RGBA16F) */
int pbrcu_ibl_equirect_to_cube(pbrcu_ctx* ctx, const float* src_rgba32f_dev, uint32_t sw, uint32_t sh,
                               uint32_t res, void* out_base_dev, pbrcu_stream s);
/* K5: 2x2 box mip chain in place (levels 1..nmips-1 from level 0), bit-exact vs the oracle.
 * face_row0/face_rows select a band of base-level rows per face for sharding (0,0 = everything). */
int pbrcu_ibl_cube_mipgen(pbrcu_ctx* ctx, void* chain_dev, uint32_t base_res, uint32_t nmips,
                          pbrcu_stream s);
/* K6/K7 read the source chain through the texture unit: ONE hardware trilinear cube fetch per sample.
 * The unit's filter weights are 8-bit fixed point (error <= 3/512 of the value range of the sample's
 * 8-texel footprint), so the texture copy carries a contrast flag and every sample whose footprint
 * spans more than 25 % falls back to fp32-weight software filtering: the result is within 1.5e-3
 * relative of the exact filter.  exact != 0 turns the texture unit off altogether (fp32 weights for
 * every sample, about 4x slower on a smooth environment).
 * The switch also covers K3's default path (ibl->accel set, one light list): there the two SMOOTH maps
 * of the ambient term -- irradiance and the split-sum LUT -- are fetched through the texture unit (the
 * irradiance copy carries the same kind of flag at 12 % footprint range with the record as fallback;
 * within 5e-4 relative of the fp32-weight filter), the prefiltered cube always from the fp32-weight
 * footprint records; exact != 0 takes all four taps from the records. */
int pbrcu_ibl_set_exact_filtering(pbrcu_ctx* ctx, int exact); /* context default; per call for K3: pbrcu_ibl.flags */
/* K3 (when ibl->accel is set): on != 0 fetches the prefiltered and irradiance cubes through the texture
 * unit (one TEX each; flagged copies, 12 % footprint range -> within 7e-4 relative of the fp32-weight
 * filter) and falls back to the fp32-weight footprint records where the filtered flag is set.  Default
 * off: with per-pixel random roughness almost every warp holds a flagged lane (coarse levels differ by
 * more than 12 % texel to texel), so the fallback runs anyway; measured 229 us vs 219 us per 4K frame.
 * Worth it only for content whose materials are coherent over a warp (179 us with no fallback). */
int pbrcu_shade_set_texture_unit(pbrcu_ctx* ctx, int on); /* context default; per call: pbrcu_ibl.flags */
/* K7: GGX prefilter of levels [first_mip,last_mip] (level 0 = copy), nsamples per texel.
 * shard/nshards split each level's texels into contiguous blocks for multi-GPU (0,1 = all). */
int pbrcu_ibl_prefilter_ggx(pbrcu_ctx* ctx, const void* src_chain_dev, uint32_t base_res,
                            uint32_t nmips, uint32_t nsamples, void* dst_chain_dev,
                            uint32_t first_mip, uint32_t last_mip, uint32_t shard, uint32_t nshards,
                            pbrcu_stream s);
/* K7 with the all-gather of a sharded prefilter fused into its stores: as pbrcu_ibl_prefilter_ggx, and every texel
 * this shard finishes is also stored at the same index of each replica chain (HOST array of n_replicas <=
 * PBRCU_MAX_REPLICAS device pointers: the peers' destination chains mapped into this process, see
 * pbrcu_shade_ggx_ibl_replicated).  When every rank has run its shard and the ranks have passed a barrier, every
 * GPU holds the whole chain -- no exchange pass. */
int pbrcu_ibl_prefilter_ggx_replicated(pbrcu_ctx* ctx, const void* src_chain_dev, uint32_t base_res,
                                       uint32_t nmips, uint32_t nsamples, void* dst_chain_dev,
                                       uint32_t first_mip, uint32_t last_mip, uint32_t shard, uint32_t nshards,
                                       void* const* replica_chains, uint32_t n_replicas, pbrcu_stream s);
/* A PREPARED source for K6 / K7: the texture copy (with contrast flags) the two convolutions sample, built once for every
 * call that convolves one environment -- the plain entry points above rebuild it inside each call (about 1.3 ms at 2048^2:
 * 3x3 min / max, flags, copy into the mip-mapped cube array), which is what stops a sharded prefilter from scaling and is
 * paid twice by a precompute that runs K6 and K7 from the same chain.  The object is a SNAPSHOT of the chain's texels
 * (rebuild it after they change) and keeps the chain's pointer: the chain must stay alive and unchanged while the object is
 * used (flagged samples are re-filtered from it with fp32 weights).  Results are bit-identical to the plain entry points.
 * pbrcu_ibl_prefilter_ggx_prepared takes the replica arguments of pbrcu_ibl_prefilter_ggx_replicated (NULL, 0: none). */
typedef struct pbrcu_ibl_source pbrcu_ibl_source;
int pbrcu_ibl_source_create(pbrcu_ctx* ctx, const void* src_chain_dev, uint32_t base_res, uint32_t nmips, pbrcu_stream s,
                            pbrcu_ibl_source** out);
int pbrcu_ibl_source_free(pbrcu_ctx* ctx, pbrcu_ibl_source* src);
int pbrcu_ibl_prefilter_ggx_prepared(pbrcu_ctx* ctx, const pbrcu_ibl_source* src, uint32_t nsamples, void* dst_chain_dev,
                                     uint32_t first_mip, uint32_t last_mip, uint32_t shard, uint32_t nshards,
                                     void* const* replica_chains, uint32_t n_replicas, pbrcu_stream s);
int pbrcu_ibl_irradiance_prepared(pbrcu_ctx* ctx, const pbrcu_ibl_source* src, uint32_t res, uint32_t nsamples, void* out_dev,
                                  pbrcu_stream s);
/* K6: cosine-weighted irradiance (E/pi), 6*res^2 RGBA16F */
int pbrcu_ibl_irradiance(pbrcu_ctx* ctx, const void* src_chain_dev, uint32_t base_res, uint32_t nmips,
                         uint32_t res, uint32_t nsamples, void* out_dev, pbrcu_stream s);

/* ---- K10: the geometry pass (fills the G-buffer) -------------------------------------------------
 * Replaces PbrRenderSystem::recordGeometryPass (src/pbr_engine/engine/pbr/PbrRenderSystem.cpp:367-465):
 * clears the planes (colour 0, depth 1.0), then for every draw runs vertex.glsl:25-37, the fixed-function
 * rasteriser (triangle list, uint16 indices + vertexOffset, back-face culling with front = CLOCKWISE, depth test
 * LESS with depth write, viewport = G-buffer extent; PbrRenderSystem.cpp:176-187, core/PipelineBuilder.hpp:92-107)
 * and fragment.glsl:19-31 (nearest + repeat sampling of R8G8B8A8_UNORM textures, gltf/Asset.cpp:116-117,
 * image/LoadImage.cpp:27).  Coverage, depth and the three colour planes are bit-exact vs the CPU checker. */
typedef struct {
  const void* texels; /* device, R8G8B8A8_UNORM, row-major */
  uint32_t width, height;
  size_t pitch; /* bytes */
} pbrcu_texture;
typedef struct {
  float color[4];                      /* MaterialData.color (Material.hpp:14-16) */
  pbrcu_texture color_tex, normal_tex; /* both required, like the reference's loader (gltf/Asset.cpp:142-157) */
  /* Tier B (no reference counterpart): constants written to the ORM / emissive planes when the G-buffer has them */
  float occlusion, roughness, metallic, emissive[3];
} pbrcu_material;
typedef struct {
  const void* vertices;    /* device; MeshVertex {vec3 position, vec3 normal, vec4 tangent, vec2 texCoords} = 48 bytes
                              (MeshVertex.hpp:14-19) */
  const uint16_t* indices; /* device */
  uint32_t first_index, index_count; /* drawIndexed(indexCount, 1, firstIndex, vertexOffset, 0), PbrRenderSystem.cpp:458-459 */
  int32_t vertex_offset;
  uint32_t vertex_count;  /* PrimitiveSpan.vertexCount (Mesh.hpp): vertices this draw may address, counted from
                             vertex_offset -- every index of the draw must be < vertex_count.  The vertex stage then runs
                             once per vertex (post-transform cache in the context's scratch).  0 = unknown: the vertex
                             stage is re-run per triangle and per pixel instead (same results, slower) */
  float model[16];        /* ModelPushConstant.model, column-major (ModelPushConstant.hpp:13-17) */
  float normal_model[12]; /* ModelPushConstant.normalModel: three vec4 columns */
  pbrcu_material material;
} pbrcu_draw;
/* view / proj: CameraData.view / .proj, column-major (CameraData.hpp:14-18).  draws: HOST array (read before the call
 * returns).  prim_id_dev (optional): width*height uint32, the winning triangle's sequence number over all draws, or
 * 0xFFFFFFFF.  Scratch (visibility buffer, piece set-ups) lives in the context and grows on demand. */
int pbrcu_geometry_pass(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, const float view[16], const float proj[16],
                        const pbrcu_draw* draws, uint32_t n_draws, uint32_t* prim_id_dev, pbrcu_stream s);

/* The geometry pass of n_views cameras over one set of draws (BASELINE configs[4]: many views of one scene): view v -- camera
 * views16[16 v ..], projs16[16 v ..], column-major like pbrcu_geometry_pass -- fills the G-buffer view_stride8 / view_stride4
 * bytes after view v-1's (all planes g0 has, depth included).  The draw tables are built and uploaded once, the vertex stage's
 * camera-independent outputs (world position, N / T / B, texture coordinates) are computed by view 0 and shared by the others, and
 * consecutive views run on internal streams (4; environment PBRCU_RASTER_SLOTS = 1..8) with a scratch set each, so their small
 * kernels overlap on the device.  Bit-identical per view to pbrcu_geometry_pass.  Replaces n_views iterations of
 * PbrRenderSystem::recordGeometryPass (PbrRenderSystem.cpp:367-465). */
int pbrcu_geometry_pass_batched(pbrcu_ctx* ctx, const pbrcu_gbuffer* g0, size_t view_stride8, size_t view_stride4,
                                uint32_t n_views, const float* views16, const float* projs16, const pbrcu_draw* draws,
                                uint32_t n_draws, pbrcu_stream s);

/* ---- multi-GPU host: ONE process drives the N GPUs of a box (csrc/mg.cu) ---------------------------------------------
 * No counterpart in the reference, which renders on one device from one host thread (core/GpuHandle.cpp:77-81,
 * gltf_viewer/App.cpp:444-483).  A group holds one pbrcu_ctx and one stream per member device with peer access between
 * every pair (cudaDeviceEnablePeerAccess: memory from pbrcu_malloc on one member is addressable from kernels on another).
 * The calls ENQUEUE on the members' streams and end with an event barrier: every member's stream waits for all the
 * others', so work enqueued afterwards on any member's stream sees the whole result; pbrcu_mg_synchronize waits on the host.
 * Allocate a member's buffers with its device current (pbrcu_mg_set_device) or through pbrcu_mg_ctx(mg, i).
 * The same device may be listed more than once (members sharing a GPU): used by the tests to run N > 1 on one device. */
typedef struct pbrcu_mg pbrcu_mg;
int pbrcu_mg_create(const int* devices, int n, pbrcu_mg** out); /* 1 <= n <= 16 */
void pbrcu_mg_destroy(pbrcu_mg* mg);
int pbrcu_mg_size(const pbrcu_mg* mg);
int pbrcu_mg_device(const pbrcu_mg* mg, int member);
pbrcu_ctx* pbrcu_mg_ctx(pbrcu_mg* mg, int member);
pbrcu_stream pbrcu_mg_stream(pbrcu_mg* mg, int member);
int pbrcu_mg_set_device(pbrcu_mg* mg, int member); /* cudaSetDevice(member's device) */
int pbrcu_mg_barrier(pbrcu_mg* mg);                /* device-side: every stream waits for all the others (events) */
int pbrcu_mg_synchronize(pbrcu_mg* mg);            /* host-side: all streams drained */
/* Row bands [y0, y1) of a frame for n members, cut on PBRCU_TILE boundaries (per-band light clustering == whole-frame
 * clustering).  tile_row_costs: one estimate per tile row -- e.g. from the previous frame's lists, sum over the row's tiles
 * of (9 + lights in the tile) -- cuts bands of equal COST; NULL cuts equal numbers of tile rows.  bands_y0y1: 2 n values. */
int pbrcu_mg_row_bands(uint32_t height, int n, const float* tile_row_costs, uint32_t* bands_y0y1);
#define PBRCU_MG_GATHER_NONE 0 /* member i's frame holds its own band only */
#define PBRCU_MG_GATHER_P2P 1  /* fused: K3 stores each finished texel into every member's frame (NVLink P2P stores) */
#define PBRCU_MG_GATHER_NCCL 2 /* after the shading, one grouped ncclBroadcast per band.  libnccl is resolved with dlopen at first
                                * use: the file PBRCU_NCCL_LIB names, else the copy already loaded in the process (e.g. PyTorch's),
                                * else the system's libnccl.so.2 -- a host that loads its own NCCL LATER must set PBRCU_NCCL_LIB */
/* One frame: member i clusters (tiles != NULL) and shades rows bands[2i] .. bands[2i+1] of ITS G-buffer g[i] with ITS IBL set
 * ibl[i] into ITS frame frames[i] (+ the gather).  g, tiles, ibl, frames: arrays of n entries; lights_host: HOST array. */
int pbrcu_mg_shade_frame(pbrcu_mg* mg, const pbrcu_gbuffer* g, const float cam_pos[3], const pbrcu_light* lights_host,
                         uint32_t n_lights, const pbrcu_light_tiles* tiles, const pbrcu_ibl* ibl, const uint32_t* bands_y0y1,
                         void* const* frames, size_t pitch, int gather);
/* pbrcu_mg_shade_frame with RGBA8 frames: the tonemap fused into K3's stores (pbrcu_shade_ggx_ibl_ldr); pitch in bytes of 4-byte texels */
int pbrcu_mg_shade_frame_ldr(pbrcu_mg* mg, const pbrcu_gbuffer* g, const float cam_pos[3], const pbrcu_light* lights_host,
                             uint32_t n_lights, const pbrcu_light_tiles* tiles, const pbrcu_ibl* ibl, const uint32_t* bands_y0y1,
                             void* const* frames, size_t pitch, int gather);
/* Batched views: member i shades its block of n_views[i] views (pbrcu_shade_ggx_ibl_batched per member); no exchange. */
int pbrcu_mg_shade_views(pbrcu_mg* mg, const pbrcu_gbuffer* g0, size_t view_stride8, size_t view_stride4,
                         const uint32_t* n_views, const float* const* cam_pos_xyz, const pbrcu_light* lights_host,
                         uint32_t n_lights, const pbrcu_ibl* ibl, void* const* outs, size_t out_pitch, size_t out_view_stride);
/* The geometry pass of view blocks: member i rasterises n_views[i] cameras (views16[i] / projs16[i], host, 16 floats per view) over its own copy
 * of the scene (draws[i]: buffers in member i's memory) into the G-buffers starting at g0[i] (pbrcu_geometry_pass_batched per member). */
int pbrcu_mg_geometry_views(pbrcu_mg* mg, const pbrcu_gbuffer* g0, size_t view_stride8, size_t view_stride4, const uint32_t* n_views,
                            const float* const* views16, const float* const* projs16, const pbrcu_draw* const* draws,
                            const uint32_t* n_draws);
/* GGX prefilter: member i convolves row block i of n of levels 1.. from its prepared source; gather != 0 stores every
 * finished texel into every member's chain, so all n chains are complete (and bit-identical to the unsharded one). */
int pbrcu_mg_prefilter(pbrcu_mg* mg, const pbrcu_ibl_source* const* sources, uint32_t nmips, uint32_t nsamples,
                       void* const* dst_chains, int gather);

/* ---- synthetic inputs for benchmarks (generated on the device; never timed) ------------------ */
/* kind 0 = gbuf_noise (every texel a hit, no coherence), 1 = gbuf_spheres (ray-cast spheres + plane,
 * ~25 % cleared background).  Fills every plane the G-buffer has. */
int pbrcu_synth_gbuffer(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, int kind, uint64_t seed,
                        const float cam_pos[3], const float cam_target[3], pbrcu_stream s);
int pbrcu_synth_environment(pbrcu_ctx* ctx, float* equirect_rgba32f_dev, uint32_t sw, uint32_t sh,
                            uint64_t seed, pbrcu_stream s);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* PBRCU_H */
