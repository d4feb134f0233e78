/* TEST INFRASTRUCTURE ONLY -- the CPU oracle.  Nothing under oracle/ is linked into, imported by or
 * executed from the product library (libpbrcu.so / physically-based-renderer_b200); only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it, as the checker.
 *
 * Tier A (pbr_oracle.c): a plain-C restatement of the reference's GLSL for the deferred lighting pass
 *   and the tonemapper.  PARITY PINNED: every function is checked bit-for-bit / to 1 ulp against the
 *   reference's own shader sources compiled for the CPU (oracle/_ref/libpbr_ref_shaders.so, built by
 *   oracle/Makefile from /root/reference/assets/shaders where they lie) and against the golden
 *   vectors those produced (tests/golden/).
 * Tier B (pbr_oracle_ibl.c): GGX/Smith/Schlick shading + the IBL precompute.  The reference has NO
 *   counterpart for these -> "PARITY UNPINNED vs. the reference"; the functions
 *   restate the published formulas (glTF 2.0 Appendix B; Karis 2013; Hammersley via bit reversal).
 */
#ifndef PBR_ORACLE_H
#define PBR_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- number formats (RNE both ways; bit-identical to __float2half_rn / __half2float) ---- */
float pbro_half_to_float(uint16_t h);
uint16_t pbro_float_to_half(float f);
void pbro_half_to_float_n(const uint16_t* h, float* f, size_t n);
void pbro_float_to_half_n(const float* f, uint16_t* h, size_t n);

/* ---- Tier A: deferred Blinn-Phong lighting (lighting.glsl:19-29, BlinnPhong.lib.glsl:4-10) ---- */
void pbro_blinn_phong(const float frag[4], const float N[3], const float V[3], const float L[3],
                      float out4[4]);
void pbro_lighting_pixel(const uint16_t pos_h[4], const uint16_t nrm_h[4], const uint16_t alb_u[4],
                         const float cam_pos[3], float out4[4]);
/* planes: positions/normals RGBA16F, albedo RGBA16_UNORM, pitch8 in bytes; shades the rectangle
 * [x0,x0+w) x [y0,y0+h); out_f32 (optional) is tightly packed w*h*4; out_f16 (optional) is the
 * RGBA16F render target addressed with the FULL-image coordinates and out_pitch bytes per row. */
void pbro_lighting_image(const void* positions, const void* normals, const void* albedo,
                         size_t pitch8, int x0, int y0, int w, int h, const float cam_pos[3],
                         float* out_f32, void* out_f16, size_t out_pitch, int nthreads);

/* ---- Tier A: ACES fit + gamma tonemapper (aces+gamma.glsl:10-40, Gamma.lib.glsl:4-6) ---- */
void pbro_aces_tonemap(const float rgb[3], float out3[3]);             /* pre-gamma */
void pbro_gamma_correct(const float rgb[3], float gamma, float out3[3]); /* pinned: pow(max(c,0),1/g) */
void pbro_tonemap_pixel(const float hdr[4], uint8_t out4[4]);
void pbro_tonemap_image(const void* hdr_rgba16f, size_t hdr_pitch, int x0, int y0, int w, int h,
                        void* out_rgba8, size_t out_pitch, int nthreads);

/* ---- Tier B: deterministic math + sample sequences (bit-exact host vs device) ---- */
void pbro_sincos2pi(float t, float* s, float* c);
uint32_t pbro_radical_inverse_bits(uint32_t i);
void pbro_hammersley(uint32_t i, uint32_t n, float out2[2]);
void pbro_hammersley_table(uint32_t n, float* out_xy /* n*2 */);

/* ---- Tier B: split-sum BRDF LUT (Karis 2013 IntegrateBRDF) -> RG32F res*res, bit-exact ---- */
void pbro_brdf_lut(int res, int nsamples, float* out_rg, int nthreads);
void pbro_brdf_lut_texel(int res, int nsamples, int i, int j, float out2[2]);

/* ---- Tier B: cube maps.  Chain layout: for mip in [0,nmips): for face in +X,-X,+Y,-Y,+Z,-Z:
 *      row-major (res>>mip)^2 RGBA16F texels, all levels contiguous. ---- */
size_t pbro_cube_chain_texels(int base_res, int nmips);
size_t pbro_cube_mip_offset(int base_res, int mip); /* in texels */
void pbro_cube_texel_dir(int face, int x, int y, int res, float dir[3]); /* unnormalised centre dir */
void pbro_cube_dir_to_face_uv(const float dir[3], int* face, float* u, float* v);
void pbro_cube_sample(const uint16_t* chain, int base_res, int nmips, const float dir[3], float lod,
                      float out4[4]);
void pbro_equirect_to_cube(const float* src_rgba32f, int sw, int sh, int res, uint16_t* out_base,
                           int nthreads);
void pbro_cube_mipgen(uint16_t* chain, int base_res, int nmips, int nthreads);

/* GGX importance-sampled specular prefilter (Karis 2013 PrefilterEnvMap, N=V=R, filtered IS).
 * table: per kept sample (Lx,Ly,Lz,lod) in the local frame; returns the kept-sample count. */
int pbro_prefilter_table(float roughness, int nsamples, int base_res, int nmips, float* table4);
void pbro_prefilter_texel(const uint16_t* src_chain, int base_res, int nmips, int mip, int nsamples,
                          int face, int x, int y, float out4[4]);
void pbro_prefilter_ggx(const uint16_t* src_chain, int base_res, int nmips, int nsamples,
                        uint16_t* dst_chain, int first_mip, int last_mip, int nthreads);
/* cosine-weighted irradiance (stores E/pi), filtered importance sampling, res*res*6 RGBA16F */
void pbro_irradiance_texel(const uint16_t* src_chain, int base_res, int nmips, int res, int nsamples,
                           int face, int x, int y, float out4[4]);
void pbro_irradiance(const uint16_t* src_chain, int base_res, int nmips, int res, int nsamples,
                     uint16_t* out, int nthreads);

/* ---- Tier B: GGX + IBL metallic-roughness shading over the extended G-buffer ---- */
typedef struct {
  float px, py, pz, range; /* range <= 0: unlimited */
  float r, g, b, intensity;
} pbro_light;
typedef struct {
  const uint16_t* irradiance; int irr_res;                  /* 6*irr_res^2 RGBA16F */
  const uint16_t* prefiltered; int pref_res; int pref_mips; /* chain */
  const float* lut; int lut_res;                            /* RG32F */
} pbro_ibl;
void pbro_shade_pixel(const uint16_t pos_h[4], const uint16_t nrm_h[4], const uint16_t alb_u[4],
                      const uint8_t orm[4], const uint16_t emi_h[4], const float cam_pos[3],
                      const pbro_light* lights, int nlights, const pbro_ibl* ibl, float out4[4]);
void pbro_shade_image(const void* positions, const void* normals, const void* albedo,
                      const void* orm, const void* emissive, size_t pitch8, size_t pitch4, int x0,
                      int y0, int w, int h, const float cam_pos[3], const pbro_light* lights,
                      int nlights, const pbro_ibl* ibl, float* out_f32, void* out_f16,
                      size_t out_pitch, int nthreads);

/* ---- geometry pass (pbr_oracle_raster.c): vertex.glsl / fragment.glsl + the fixed-function rasteriser ---- */
typedef struct { const void* texels; int width, height; size_t pitch; } pbro_texture; /* R8G8B8A8_UNORM */
typedef struct {
  const void* vertices;         /* MeshVertex: pos3 nrm3 tan4 uv2, 48 bytes (MeshVertex.hpp:14-19) */
  const uint16_t* indices;
  uint32_t first_index, index_count;
  int32_t vertex_offset;
  float model[16];              /* ModelPushConstant.model, column-major */
  float normal_model[12];       /* ModelPushConstant.normalModel: three vec4 columns */
  float color[4];               /* MaterialData.color */
  pbro_texture color_tex, normal_tex;
  float occlusion, roughness, metallic, emissive[3]; /* Tier B planes (no reference counterpart) */
} pbro_draw;
void pbro_geom_vertex(const float model[16], const float normal_model[12], const float view[16], const float proj[16],
                      const float vertex[12], float clip[4], float attrs[14]);
void pbro_geom_fragment(const float attrs[14], const float color[4], const pbro_texture* color_tex, const pbro_texture* normal_tex,
                        float out_pos[4], float out_nrm[4], float out_alb[4]);
void pbro_geometry_pass(const pbro_draw* draws, int n_draws, const float view[16], const float proj[16], int W, int H,
                        uint16_t* positions, uint16_t* normals, uint16_t* albedo, float* depth, uint8_t* orm, uint16_t* emissive,
                        uint32_t* prim_id);

int pbro_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
