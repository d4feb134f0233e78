// common.cuh -- shared device helpers for the sm_100a kernels (vector ld/st, number formats, status).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pbrcu.h"

namespace pbrcu {
constexpr int kRasterSlots = 8;  // most concurrent views of pbrcu_geometry_pass_batched (scratch sets are created on first use)
// a cube-map mip chain as a hardware texture with contrast flags in alpha (cube_tex.cuh)
struct CubeTex {
  cudaMipmappedArray_t array;
  cudaTextureObject_t tex;
  uint32_t res, mips;
};
}  // namespace pbrcu

struct pbrcu_ctx {
  int device;
  int sm_count;
  int max_smem_optin;
  // staging for pbrcu_frame_host (grown on demand)
  void* stage[8];
  size_t stage_bytes[8];
  // two internal streams (+ events) that pipeline row bands of a host frame: H2D | shade | D2H
  cudaStream_t pipe[2];
  cudaEvent_t pipe_done[2];
  cudaEvent_t pipe_start;
  cudaEvent_t band_ready[64];  // band b's planes have arrived (recorded on the copy-in stream)
  pbrcu_light* lights_dev;
  uint32_t lights_cap;
  // per-view camera positions for the batched multi-view launch (grown on demand)
  float* cams_dev;
  uint32_t cams_cap;
  // cube-map texture (mipmapped CUDA array + texture object) the IBL convolutions sample; kept between
  // calls and re-created only when the chain shape changes (array allocation costs milliseconds)
  pbrcu::CubeTex cube;
  void *cube_stage, *cube_mm;  // staging copy with contrast flags in alpha; per-texel 3x3 min/max
  size_t cube_stage_bytes, cube_mm_bytes;
  bool exact_ibl_filter;  // K6/K7: software fp32-weight filtering instead of the texture unit
  void* raster_scratch;  // K10: visibility buffer, piece set-ups, chunk queue, draw table (grown on demand)
  size_t raster_bytes;
  size_t raster_clean_px;  // leading entries of the scratch's visibility buffer that the last pass left empty (its resolve kernel resets what it reads)
  // pbrcu_geometry_pass_batched: consecutive views run on these internal streams, one scratch set each
  void* raster_slot_scratch[pbrcu::kRasterSlots];
  size_t raster_slot_bytes[pbrcu::kRasterSlots];
  size_t raster_slot_clean_px[pbrcu::kRasterSlots];
  cudaStream_t raster_slot_stream[pbrcu::kRasterSlots];
  cudaEvent_t raster_slot_done[pbrcu::kRasterSlots];
  cudaEvent_t raster_start;
  bool shade_tex;         // K3: texture-unit fetches of the prefiltered / irradiance cubes (pbrcu_shade_set_texture_unit)
};

namespace pbrcu {

void set_error(const char* fmt, ...);
// true for pageable or pinned HOST memory (entry points that take "device or host" arrays of small records)
bool is_host_pointer(const void* p);
int cuda_fail(cudaError_t e, const char* what);

#define PBRCU_CUDA(call)                                         \
  do {                                                           \
    cudaError_t _e = (call);                                     \
    if (_e != cudaSuccess) return ::pbrcu::cuda_fail(_e, #call); \
  } while (0)

#define PBRCU_REQUIRE(cond, msg)                        \
  do {                                                  \
    if (!(cond)) {                                      \
      ::pbrcu::set_error("invalid argument: %s", msg);  \
      return PBRCU_ERR_INVALID;                         \
    }                                                   \
  } while (0)

inline int launch_check(const char* name) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, name);
  return PBRCU_OK;
}

// ---- streaming 128-bit global access: the G-buffer is read once and the output written once, so
// keep them out of L1 and mark them evict-first in L2 (the cube maps / LUT are what should stay).
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ uint2 ldg_stream8(const void* p) {
  uint2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];"
               : "=r"(v.x), "=r"(v.y)
               : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t ldg_stream4(const void* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void stg_stream(void* p, uint4 v) {
  asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void stg_stream8(void* p, uint2 v) {
  asm volatile("st.global.cs.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void stg_stream4(void* p, uint32_t v) {
  asm volatile("st.global.cs.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---- number formats --------------------------------------------------------------------------
__device__ __forceinline__ float2 h2_to_f2(uint32_t packed) {
  __half2 h = *reinterpret_cast<__half2*>(&packed);
  return __half22float2(h);
}
__device__ __forceinline__ uint32_t f2_to_h2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);  // one RNE rounding per component
  return *reinterpret_cast<uint32_t*>(&h);
}
// R16G16B16A16_UNORM decode: c / 65535
__device__ __forceinline__ float unorm16_lo(uint32_t packed) {
  return __uint2float_rn(packed & 0xFFFFu) * (1.0f / 65535.0f);
}
__device__ __forceinline__ float unorm16_hi(uint32_t packed) {
  return __uint2float_rn(packed >> 16) * (1.0f / 65535.0f);
}
__device__ __forceinline__ uint32_t to_unorm8(float v) {
  // clamp with maxNum/minNum semantics (NaN -> 0), then round to nearest even
  float c = fminf(fmaxf(v, 0.0f), 1.0f);
  return __float2uint_rn(c * 255.0f);
}

struct Px8 {  // one 8-byte texel as two packed words
  uint32_t lo, hi;
};

}  // namespace pbrcu
