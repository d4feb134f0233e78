// synth.cu -- seeded synthetic inputs generated on the device for benchmarks:
// gbuf_noise, gbuf_spheres and a procedural HDR environment.  Generation is never timed; parity
// checks always run the oracle on the bits these kernels actually wrote (downloaded), so nothing
// here needs to match a host generator.
#include "cube.cuh"

namespace pbrcu {

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
struct Rng {
  uint64_t s;
  __device__ explicit Rng(uint64_t seed) : s(seed) {}
  __device__ uint32_t u32() { s = splitmix64(s); return uint32_t(s >> 32); }
  __device__ float uni() { return float(u32() >> 8) * (1.0f / 16777216.0f); }  // [0,1), 24 bits
  __device__ float range(float a, float b) { return a + (b - a) * uni(); }
};

struct SynthParams {
  uint8_t *pos, *nrm, *alb, *orm, *emi;
  float* depth;
  uint32_t w, h;
  size_t pitch8, pitch4;
  uint64_t seed;
  int kind;
  float cam[3], fwd[3], right[3], up[3];
  float tan_half_fov, aspect;
};

__device__ __forceinline__ uint2 pack_h4(float a, float b, float c, float d) {
  return make_uint2(f2_to_h2(a, b), f2_to_h2(c, d));
}

__device__ __forceinline__ void write_material(const SynthParams& p, size_t o8, size_t o4, Rng& r, float3 tint) {
  const uint32_t ar = uint32_t(fminf(fmaxf(tint.x * r.range(0.6f, 1.0f), 0.f), 1.f) * 65535.0f + 0.5f);
  const uint32_t ag = uint32_t(fminf(fmaxf(tint.y * r.range(0.6f, 1.0f), 0.f), 1.f) * 65535.0f + 0.5f);
  const uint32_t ab = uint32_t(fminf(fmaxf(tint.z * r.range(0.6f, 1.0f), 0.f), 1.f) * 65535.0f + 0.5f);
  const uint32_t aa = r.u32() & 0xFFFFu;
  *reinterpret_cast<uint2*>(p.alb + o8) = make_uint2(ar | (ag << 16), ab | (aa << 16));
  if (p.orm) {
    const uint32_t occ = uint32_t(r.range(0.5f, 1.0f) * 255.0f + 0.5f);
    const uint32_t rough = uint32_t(r.range(0.045f, 1.0f) * 255.0f + 0.5f);
    const float m = r.uni();
    const uint32_t metal = m < 0.45f ? 0u : (m < 0.9f ? 255u : uint32_t(r.uni() * 255.0f + 0.5f));
    *reinterpret_cast<uint32_t*>(p.orm + o4) = occ | (rough << 8) | (metal << 16) | (255u << 24);
  }
  if (p.emi) {
    uint2 e = make_uint2(0u, 0u);
    if (r.uni() < 0.05f) e = pack_h4(r.range(0.f, 4.f), r.range(0.f, 4.f), r.range(0.f, 4.f), 0.f);
    *reinterpret_cast<uint2*>(p.emi + o8) = e;
  }
}

constexpr int kSpheres = 64;

__global__ void __launch_bounds__(256) synth_gbuffer_kernel(const __grid_constant__ SynthParams p) {
  __shared__ float4 s_sphere[kSpheres];   // centre, radius
  __shared__ float4 s_tint[kSpheres];
  if (threadIdx.x < kSpheres) {
    Rng r(p.seed * 0x100000001B3ull + 7919ull * threadIdx.x);
    s_sphere[threadIdx.x] = make_float4(r.range(-8.f, 8.f), r.range(-8.f, 1.f), r.range(-8.f, 8.f), r.range(0.5f, 1.5f));
    s_tint[threadIdx.x] = make_float4(r.range(0.1f, 1.f), r.range(0.1f, 1.f), r.range(0.1f, 1.f), 0.f);
  }
  __syncthreads();
  const size_t total = size_t(p.w) * p.h;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const uint32_t x = uint32_t(i % p.w), y = uint32_t(i / p.w);
    const size_t o8 = size_t(y) * p.pitch8 + size_t(x) * 8, o4 = size_t(y) * p.pitch4 + size_t(x) * 4;
    Rng r(p.seed ^ (uint64_t(i) * 0x9E3779B97F4A7C15ull));
    if (p.kind == 0) {  // gbuf_noise: every texel a hit, no spatial coherence
      *reinterpret_cast<uint2*>(p.pos + o8) = pack_h4(r.range(-10.f, 10.f), r.range(-10.f, 10.f), r.range(-10.f, 10.f), 0.f);
      const float z = r.range(-1.f, 1.f), phi = r.range(0.f, 6.2831853f), s = sqrtf(fmaxf(1.f - z * z, 0.f));
      *reinterpret_cast<uint2*>(p.nrm + o8) = pack_h4(s * cosf(phi), s * sinf(phi), z, 0.f);
      write_material(p, o8, o4, r, make_float3(r.uni(), r.uni(), r.uni()));
      if (p.depth) p.depth[o4 / 4] = r.uni();
      continue;
    }
    // gbuf_spheres: closed-form ray cast of 64 spheres + the plane y = 2 (world up is -Y like the
    // reference's camera, CameraData.hpp:28)
    const float nx = (2.0f * (float(x) + 0.5f) / float(p.w) - 1.0f) * p.tan_half_fov * p.aspect;
    const float ny = (2.0f * (float(y) + 0.5f) / float(p.h) - 1.0f) * p.tan_half_fov;
    const float3 d = normalize3(make_float3(p.fwd[0] + nx * p.right[0] - ny * p.up[0], p.fwd[1] + nx * p.right[1] - ny * p.up[1],
                                            p.fwd[2] + nx * p.right[2] - ny * p.up[2]));
    const float3 o = make_float3(p.cam[0], p.cam[1], p.cam[2]);
    float tbest = 1.0e30f;
    int hit = -1;
    for (int k = 0; k < kSpheres; ++k) {
      const float4 sp = s_sphere[k];
      const float ox = o.x - sp.x, oy = o.y - sp.y, oz = o.z - sp.z;
      const float b = ox * d.x + oy * d.y + oz * d.z, c = ox * ox + oy * oy + oz * oz - sp.w * sp.w;
      const float disc = b * b - c;
      if (disc > 0.f) {
        const float t = -b - sqrtf(disc);
        if (t > 0.01f && t < tbest) { tbest = t; hit = k; }
      }
    }
    if (d.y > 1e-4f) {
      const float t = (2.0f - o.y) / d.y;
      if (t > 0.01f && t < tbest && t < 60.f) { tbest = t; hit = kSpheres; }
    }
    if (hit < 0) {  // miss: the reference's clear values (PbrRenderSystem.cpp:370-415)
      *reinterpret_cast<uint2*>(p.pos + o8) = make_uint2(0u, 0u);
      *reinterpret_cast<uint2*>(p.nrm + o8) = make_uint2(0u, 0u);
      *reinterpret_cast<uint2*>(p.alb + o8) = make_uint2(0u, 0u);
      if (p.orm) *reinterpret_cast<uint32_t*>(p.orm + o4) = 0u;
      if (p.emi) *reinterpret_cast<uint2*>(p.emi + o8) = make_uint2(0u, 0u);
      if (p.depth) p.depth[o4 / 4] = 1.0f;
      continue;
    }
    const float3 P = make_float3(o.x + tbest * d.x, o.y + tbest * d.y, o.z + tbest * d.z);
    float3 N, tint;
    if (hit == kSpheres) { N = make_float3(0.f, -1.f, 0.f); const float ch = ((int(floorf(P.x)) + int(floorf(P.z))) & 1) ? 0.8f : 0.3f; tint = make_float3(ch, ch, ch); }
    else { const float4 sp = s_sphere[hit]; N = normalize3(make_float3(P.x - sp.x, P.y - sp.y, P.z - sp.z)); const float4 tt = s_tint[hit]; tint = make_float3(tt.x, tt.y, tt.z); }
    N = normalize3(make_float3(N.x + r.range(-0.2f, 0.2f), N.y + r.range(-0.2f, 0.2f), N.z + r.range(-0.2f, 0.2f)));  // normal-map noise
    *reinterpret_cast<uint2*>(p.pos + o8) = pack_h4(P.x, P.y, P.z, 0.f);
    *reinterpret_cast<uint2*>(p.nrm + o8) = pack_h4(N.x, N.y, N.z, 0.f);
    write_material(p, o8, o4, r, tint);
    if (p.depth) p.depth[o4 / 4] = fminf(tbest / 1024.0f, 1.0f);
  }
}

// procedural sky: vertical gradient + sun disc (radiance 5e4, 0.5 deg radius) + low-amplitude noise
__global__ void __launch_bounds__(256) synth_environment_kernel(float4* out, uint32_t sw, uint32_t sh, uint64_t seed) {
  const size_t total = size_t(sw) * sh;
  const float3 sun = normalize3(make_float3(0.35f, 0.8f, 0.45f));
  const float cos_sun = cosf(0.5f * 3.14159265f / 180.0f);
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const uint32_t x = uint32_t(i % sw), y = uint32_t(i / sw);
    const float u = (float(x) + 0.5f) / float(sw), v = (float(y) + 0.5f) / float(sh);
    const float theta = v * 3.14159265f, phi = (u - 0.5f) * 6.2831853f;
    const float3 d = make_float3(sinf(theta) * cosf(phi), cosf(theta), sinf(theta) * sinf(phi));
    Rng r(seed ^ (uint64_t(i) * 0x9E3779B97F4A7C15ull));
    const float t = 0.5f * (d.y + 1.0f);
    float cr = 0.15f + 0.5f * t * 0.6f, cg = 0.2f + 0.7f * t * 0.7f, cb = 0.25f + 1.4f * t;
    const float n = 1.0f + 0.1f * (r.uni() - 0.5f);
    cr *= n; cg *= n; cb *= n;
    if (d.x * sun.x + d.y * sun.y + d.z * sun.z > cos_sun) { cr = 5.0e4f; cg = 4.5e4f; cb = 4.0e4f; }
    out[i] = make_float4(cr, cg, cb, 1.0f);
  }
}

}  // namespace pbrcu

using namespace pbrcu;

extern "C" int pbrcu_synth_gbuffer(pbrcu_ctx* ctx, const pbrcu_gbuffer* g, int kind, uint64_t seed, const float cam_pos[3],
                                   const float cam_target[3], pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && g && g->positions && g->normals && g->albedo, "NULL argument");
  PBRCU_REQUIRE(kind == 0 || kind == 1, "kind must be 0 (noise) or 1 (spheres)");
  PBRCU_REQUIRE((g->orm == nullptr) == (g->emissive == nullptr), "orm and emissive must both be present or both absent");
  SynthParams p{};
  p.pos = static_cast<uint8_t*>(g->positions); p.nrm = static_cast<uint8_t*>(g->normals); p.alb = static_cast<uint8_t*>(g->albedo);
  p.orm = static_cast<uint8_t*>(g->orm); p.emi = static_cast<uint8_t*>(g->emissive); p.depth = g->depth;
  p.w = g->width; p.h = g->height; p.pitch8 = g->pitch8; p.pitch4 = g->pitch4; p.seed = seed; p.kind = kind;
  float cam[3] = {0.f, -1.f, -12.f}, tgt[3] = {0.f, 0.f, 0.f};
  if (cam_pos) for (int i = 0; i < 3; ++i) cam[i] = cam_pos[i];
  if (cam_target) for (int i = 0; i < 3; ++i) tgt[i] = cam_target[i];
  // the reference's camera basis: lookAtRH(position, target, up = (0,-1,0)) (CameraData.hpp:28), fov pi/2
  float f[3] = {tgt[0] - cam[0], tgt[1] - cam[1], tgt[2] - cam[2]};
  float fl = sqrtf(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
  PBRCU_REQUIRE(fl > 0.f, "camera position equals target");
  for (float& c : f) c /= fl;
  const float upw[3] = {0.f, -1.f, 0.f};
  float r[3] = {f[1] * upw[2] - f[2] * upw[1], f[2] * upw[0] - f[0] * upw[2], f[0] * upw[1] - f[1] * upw[0]};
  float rl = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  PBRCU_REQUIRE(rl > 1e-6f, "camera looks along the up axis");
  for (float& c : r) c /= rl;
  const float u[3] = {r[1] * f[2] - r[2] * f[1], r[2] * f[0] - r[0] * f[2], r[0] * f[1] - r[1] * f[0]};
  for (int i = 0; i < 3; ++i) { p.cam[i] = cam[i]; p.fwd[i] = f[i]; p.right[i] = r[i]; p.up[i] = u[i]; }
  p.tan_half_fov = 1.0f;
  p.aspect = float(g->width) / float(g->height);
  synth_gbuffer_kernel<<<uint32_t(ctx->sm_count) * 8, 256, 0, static_cast<cudaStream_t>(s)>>>(p);
  return launch_check("synth_gbuffer_kernel");
}

extern "C" int pbrcu_synth_environment(pbrcu_ctx* ctx, float* equirect_rgba32f_dev, uint32_t sw, uint32_t sh, uint64_t seed,
                                       pbrcu_stream s) {
  PBRCU_REQUIRE(ctx && equirect_rgba32f_dev && sw > 0 && sh > 0, "NULL argument or empty extent");
  synth_environment_kernel<<<uint32_t(ctx->sm_count) * 8, 256, 0, static_cast<cudaStream_t>(s)>>>(
      reinterpret_cast<float4*>(equirect_rgba32f_dev), sw, sh, seed);
  return launch_check("synth_environment_kernel");
}
