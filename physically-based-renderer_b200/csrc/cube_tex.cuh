// cube_tex.cuh -- a cube-map mip chain as a hardware texture (mipmapped cube-map CUDA array + texture
// object: per-face clamp-to-edge bilinear, linear between levels) whose ALPHA carries a contrast flag.
// The texture unit filters with 8-bit fixed-point weights, so a trilinear sample is off by up to 3/512 of
// the value range of its 8-texel footprint; a texel is flagged (alpha = 1) when that range can exceed
// `rel_range` x the footprint's minimum, and a sample whose filtered alpha is > 0 must be re-done with
// fp32 weights (cube.cuh / ibl_quads.cuh).  Used by K6/K7 on the source environment (rel_range 0.25,
// error averaged over the sample set) and by K3 on the prefiltered / irradiance maps (rel_range 0.12).
#pragma once
#include "cube.cuh"

namespace pbrcu {

// (Re)creates t for this chain shape if needed and copies the chain into it with the flags in alpha
// (device-to-device on `st`; staging buffers are the context's and are grown on demand).
// A sample between levels l and l+1 reads the flags of both; the flag of a level-l texel always covers its
// own 3x3 neighbourhood and its parent's (which hold the sample's footprints on l and l+1), so the finer
// level's flags alone decide correctly.  guard_coarser additionally flags a texel for the contrast of its
// children (its use as the coarser level), which callers need only if they cannot rule out a level
// fraction that the unit rounds to exactly 1.
int cube_tex_upload(pbrcu_ctx* ctx, CubeTex* t, const void* chain, uint32_t base_res, uint32_t nmips, float rel_range,
                    bool guard_coarser, cudaStream_t st);
void cube_tex_destroy(CubeTex* t);
inline size_t cube_tex_bytes(const CubeTex& t) { return t.array ? cube_mip_offset(t.res, t.mips) * 8 : 0; }

}  // namespace pbrcu
