/* TEST INFRASTRUCTURE ONLY -- CPU oracle of the geometry pass: the pass that
 * FILLS the G-buffer the hot path shades.  Never linked into libpbrcu.so.
 *
 * What it restates, and how each part is pinned:
 *   vertex stage   = assets/shaders/geometry_pass/vertex.glsl:25-37          PINNED: checked bit-for-bit against
 *   fragment stage = assets/shaders/geometry_pass/fragment.glsl:19-31        the reference's own GLSL compiled for
 *                                                                            the CPU (oracle/_ref, tests/test_oracle_raster.py)
 *   draw loop / state = src/pbr_engine/engine/pbr/PbrRenderSystem.cpp:367-465 (clear 0/0/0/depth 1, viewport =
 *                    G-buffer extent, one drawIndexed per primitive, uint16 indices + vertexOffset), pipeline state
 *                    :176-187 + core/PipelineBuilder.hpp:92-107 (back-face cull, front = CLOCKWISE, depth test LESS,
 *                    depth write), material sampler = all-zero VkSamplerCreateInfo (gltf/Asset.cpp:116-117: nearest,
 *                    repeat), textures R8G8B8A8_UNORM (image/LoadImage.cpp:27)
 *   fixed-function rasterisation (clipping, viewport transform, coverage, interpolation) lives in the Vulkan
 *                    driver / hardware, not in /root/reference: PARITY UNPINNED -- restated from the Vulkan 1.3
 *                    specification (ch. "Fixed-Function Vertex Post-Processing", "Rasterization") with the
 *                    implementation-defined parts pinned as follows:
 *     - clip volume 0 <= z <= w: the near plane clips geometrically (Sutherland-Hodgman, new vertex =
 *       inside + t (outside - inside), t = d_in / (d_in - d_out), every attribute); the far plane discards fragments
 *       with z_fb >= 1 (equal coverage, and what LESS against the clear value 1.0 does anyway);
 *     - x/y are clipped against a guard band |x|,|y| <= 1024 w instead of the view volume (no visible effect;
 *       bounds the fixed-point coordinates);
 *     - framebuffer coordinates are snapped to 1/256 pixel (RNE); coverage = exact 64-bit edge functions at
 *       pixel centres with the top-left rule;
 *     - perspective divide by one reciprocal per vertex: iw = 1 / w, ndc = clip * iw;
 *     - barycentrics l_i = float(E_i) / float(2A); z_fb = (l0 z0 + l1 z1) + l2 z2; perspective-correct attributes
 *       a = ((p0 a0 + p1 a1) + p2 a2) * (1 / ((p0 + p1) + p2)), p_i = l_i * iw_i;
 *     - every fp32 step is ONE correctly rounded operation in the order written here (compile with
 *       -ffp-contract=off); the device kernels use __fmul_rn/__fadd_rn/__fdiv_rn/__fsqrt_rn in the same order, so
 *       coverage, depth and all three colour planes are BIT-EXACT between this file and csrc/raster.cu.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "pbr_oracle.h"

#define RAST_SUB 256         /* subpixel steps per pixel */
#define RAST_GUARD 1024.0f   /* guard band, in units of w */
#define RAST_MAX_POLY 8      /* a triangle clipped by 5 planes has at most 8 vertices */

typedef struct {
  float clip[4];                /* gl_Position */
  float a[14];                  /* outPosition(3) outNormal(3) outTangent(3) outBitangent(3) outTexCoords(2) */
} rvert;

static void mat4_mul_vec4(const float* m, const float v[4], float r[4]) { /* column-major; shim order */
  for (int i = 0; i < 4; ++i) r[i] = ((m[0 + i] * v[0] + m[4 + i] * v[1]) + m[8 + i] * v[2]) + m[12 + i] * v[3];
}
static void mat3_mul_vec3(const float* m /* 3 columns, stride 4 */, const float v[3], float r[3]) {
  for (int i = 0; i < 3; ++i) r[i] = (m[0 + i] * v[0] + m[4 + i] * v[1]) + m[8 + i] * v[2];
}
static void normalize3(float v[3]) {
  const float d = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
  const float inv = 1.0f / sqrtf(d);
  v[0] = v[0] * inv; v[1] = v[1] * inv; v[2] = v[2] * inv;
}

/* vertex.glsl:25-37 */
void pbro_geom_vertex(const float model[16], const float normal_model[12], const float view[16], const float proj[16],
                      const float vertex[12] /* MeshVertex: pos3 nrm3 tan4 uv2 */, float clip[4], float attrs[14]) {
  const float p[4] = {vertex[0], vertex[1], vertex[2], 1.0f};
  float world[4];
  mat4_mul_vec4(model, p, world);                       /* :26 */
  float pv[16];
  for (int j = 0; j < 4; ++j) mat4_mul_vec4(proj, view + 4 * j, pv + 4 * j);  /* cam.proj * cam.view */
  mat4_mul_vec4(pv, world, clip);                       /* :28 */
  attrs[0] = world[0]; attrs[1] = world[1]; attrs[2] = world[2];  /* :30 */
  const float* n = vertex + 3;
  const float* t = vertex + 6;
  float N[3], T[3], B[3];
  mat3_mul_vec3(normal_model, n, N); normalize3(N);     /* :32 */
  mat3_mul_vec3(normal_model, t, T); normalize3(T);     /* :33 */
  float c[3] = {n[1] * t[2] - t[1] * n[2], n[2] * t[0] - t[2] * n[0], n[0] * t[1] - t[0] * n[1]};
  c[0] = c[0] * t[3]; c[1] = c[1] * t[3]; c[2] = c[2] * t[3];
  mat3_mul_vec3(normal_model, c, B); normalize3(B);     /* :34 */
  memcpy(attrs + 3, N, 12); memcpy(attrs + 6, T, 12); memcpy(attrs + 9, B, 12);
  attrs[12] = vertex[10]; attrs[13] = vertex[11];       /* :36 */
}

/* nearest filter, repeat addressing, R8G8B8A8_UNORM */
static void fetch_tex(const pbro_texture* t, float u, float v, float out[4]) {
  int ix = (int)floorf(u * (float)t->width), iy = (int)floorf(v * (float)t->height);
  ix %= t->width; if (ix < 0) ix += t->width;
  iy %= t->height; if (iy < 0) iy += t->height;
  const uint8_t* px = (const uint8_t*)t->texels + (size_t)iy * t->pitch + 4 * (size_t)ix;
  for (int c = 0; c < 4; ++c) out[c] = (float)px[c] / 255.0f;
}

/* fragment.glsl:19-31; outputs the fp32 shader values */
void pbro_geom_fragment(const float attrs[14], const float color[4], const pbro_texture* color_tex, const pbro_texture* normal_tex,
                        float out_pos[4], float out_nrm[4], float out_alb[4]) {
  out_pos[0] = attrs[0]; out_pos[1] = attrs[1]; out_pos[2] = attrs[2]; out_pos[3] = 0.0f;  /* :21 */
  float tn[4], tc[4];
  fetch_tex(normal_tex, attrs[12], attrs[13], tn);
  const float nx = tn[0] * 2.0f - 1.0f, ny = tn[1] * 2.0f - 1.0f, nz = tn[2] * 2.0f - 1.0f;  /* :25 */
  const float* N = attrs + 3; const float* T = attrs + 6; const float* B = attrs + 9;
  float w[3];
  for (int i = 0; i < 3; ++i) w[i] = (T[i] * nx + B[i] * ny) + N[i] * nz;  /* mat3(T,B,N) * normal, :24,26 */
  normalize3(w);
  out_nrm[0] = w[0]; out_nrm[1] = w[1]; out_nrm[2] = w[2]; out_nrm[3] = 0.0f;  /* :27 */
  fetch_tex(color_tex, attrs[12], attrs[13], tc);
  for (int c = 0; c < 4; ++c) out_alb[c] = color[c] * tc[c];  /* :30 */
}

static uint16_t to_unorm16(float v) {
  const float c = fminf(fmaxf(v, 0.0f), 1.0f);  /* NaN -> 0 */
  return (uint16_t)lrintf(c * 65535.0f);
}
static uint8_t to_unorm8(float v) {
  const float c = fminf(fmaxf(v, 0.0f), 1.0f);
  return (uint8_t)lrintf(c * 255.0f);
}

static void lerp_vert(const rvert* in, const rvert* outv, float t, rvert* r) {
  for (int i = 0; i < 4; ++i) r->clip[i] = in->clip[i] + t * (outv->clip[i] - in->clip[i]);
  for (int i = 0; i < 14; ++i) r->a[i] = in->a[i] + t * (outv->a[i] - in->a[i]);
}
static float plane_dist(const rvert* v, int plane) {
  switch (plane) {
    case 0: return v->clip[2];                                   /* z >= 0 */
    case 1: return v->clip[0] + RAST_GUARD * v->clip[3];         /* x >= -G w */
    case 2: return RAST_GUARD * v->clip[3] - v->clip[0];         /* x <=  G w */
    case 3: return v->clip[1] + RAST_GUARD * v->clip[3];
    default: return RAST_GUARD * v->clip[3] - v->clip[1];
  }
}
/* Sutherland-Hodgman against the five planes; returns the vertex count (0 = fully clipped) */
static int clip_polygon(rvert* poly, int n) {
  rvert tmp[RAST_MAX_POLY + 2];
  for (int plane = 0; plane < 5 && n > 0; ++plane) {
    int m = 0, all_in = 1;
    float d[RAST_MAX_POLY + 2];
    for (int i = 0; i < n; ++i) { d[i] = plane_dist(&poly[i], plane); if (!(d[i] >= 0.0f)) all_in = 0; }
    if (all_in) continue;
    for (int i = 0; i < n; ++i) {
      const int j = (i + 1) % n;
      const int in_i = d[i] >= 0.0f, in_j = d[j] >= 0.0f;
      if (in_i) tmp[m++] = poly[i];
      if (in_i != in_j) {  /* always from the inside vertex towards the outside one: shared edges stay watertight */
        if (in_i) lerp_vert(&poly[i], &poly[j], d[i] / (d[i] - d[j]), &tmp[m++]);
        else lerp_vert(&poly[j], &poly[i], d[j] / (d[j] - d[i]), &tmp[m++]);
      }
    }
    n = m;
    memcpy(poly, tmp, sizeof(rvert) * (size_t)n);
  }
  return n;
}

typedef struct { int64_t X[3], Y[3]; float z[3], iw[3]; const rvert* v[3]; int64_t S; int x0, y0, x1, y1; } rtri;

static int setup_tri(const rvert* a, const rvert* b, const rvert* c, int W, int H, rtri* t) {
  const rvert* v[3] = {a, b, c};
  for (int i = 0; i < 3; ++i)
    if (!(v[i]->clip[3] > 0.0f)) return 0;  /* w <= 0 survives the clip planes only at the degenerate point x = y = z = w = 0 */
  for (int i = 0; i < 3; ++i) {
    const float iw = 1.0f / v[i]->clip[3];   /* one reciprocal per vertex; the perspective divide multiplies by it */
    const float xn = v[i]->clip[0] * iw, yn = v[i]->clip[1] * iw;
    const float xf = ((float)W * 0.5f) * xn + (float)W * 0.5f, yf = ((float)H * 0.5f) * yn + (float)H * 0.5f;
    t->X[i] = (int64_t)lrintf(xf * (float)RAST_SUB); t->Y[i] = (int64_t)lrintf(yf * (float)RAST_SUB);
    t->z[i] = v[i]->clip[2] * iw;
    t->iw[i] = iw;
    t->v[i] = v[i];
  }
  t->S = (t->X[1] - t->X[0]) * (t->Y[2] - t->Y[0]) - (t->X[2] - t->X[0]) * (t->Y[1] - t->Y[0]);
  if (t->S <= 0) return 0;  /* back-facing (front = clockwise = positive S with y down) or degenerate */
  int64_t minx = t->X[0], maxx = t->X[0], miny = t->Y[0], maxy = t->Y[0];
  for (int i = 1; i < 3; ++i) {
    if (t->X[i] < minx) minx = t->X[i];
    if (t->X[i] > maxx) maxx = t->X[i];
    if (t->Y[i] < miny) miny = t->Y[i];
    if (t->Y[i] > maxy) maxy = t->Y[i];
  }
  /* pixel (x,y) has its centre at (x*256+128): the covered candidates are ceil((min-128)/256) .. floor((max-128)/256) */
  int64_t x0 = (minx - 128 + 255) >> 8, x1 = (maxx - 128) >> 8, y0 = (miny - 128 + 255) >> 8, y1 = (maxy - 128) >> 8;
  if (x0 < 0) x0 = 0;
  if (y0 < 0) y0 = 0;
  if (x1 > W - 1) x1 = W - 1;
  if (y1 > H - 1) y1 = H - 1;
  if (x0 > x1 || y0 > y1) return 0;
  t->x0 = (int)x0; t->y0 = (int)y0; t->x1 = (int)x1; t->y1 = (int)y1;
  return 1;
}

static int edge_inside(int64_t E, int64_t dx, int64_t dy) {  /* top-left rule, clockwise triangle, y down */
  if (E != 0) return E > 0;
  return dy < 0 || (dy == 0 && dx > 0);
}

void pbro_geometry_pass(const pbro_draw* draws, int n_draws, const float view[16], const float proj[16], int W, int H,
                        uint16_t* positions, uint16_t* normals, uint16_t* albedo, float* depth, uint8_t* orm, uint16_t* emissive,
                        uint32_t* prim_id /* optional: winning triangle sequence number per pixel, 0xFFFFFFFF = none */) {
  const size_t npx = (size_t)W * (size_t)H;
  memset(positions, 0, npx * 8); memset(normals, 0, npx * 8); memset(albedo, 0, npx * 8);
  if (orm) memset(orm, 0, npx * 4);
  if (emissive) memset(emissive, 0, npx * 8);
  for (size_t i = 0; i < npx; ++i) depth[i] = 1.0f;
  if (prim_id) memset(prim_id, 0xFF, npx * 4);
  uint32_t seq = 0;
  for (int di = 0; di < n_draws; ++di) {
    const pbro_draw* d = &draws[di];
    const float* verts = (const float*)d->vertices;
    for (uint32_t k = 0; k + 3 <= d->index_count; k += 3, ++seq) {
      rvert poly[RAST_MAX_POLY + 2];
      for (int c = 0; c < 3; ++c) {
        const int64_t vi = (int64_t)d->indices[d->first_index + k + (uint32_t)c] + d->vertex_offset;
        pbro_geom_vertex(d->model, d->normal_model, view, proj, verts + 12 * vi, poly[c].clip, poly[c].a);
      }
      const int n = clip_polygon(poly, 3);
      for (int piece = 0; piece + 2 < n; ++piece) {
        rtri t;
        if (!setup_tri(&poly[0], &poly[piece + 1], &poly[piece + 2], W, H, &t)) continue;
        const float fS = (float)t.S;
        for (int y = t.y0; y <= t.y1; ++y)
          for (int x = t.x0; x <= t.x1; ++x) {
            const int64_t px = (int64_t)x * RAST_SUB + 128, py = (int64_t)y * RAST_SUB + 128;
            int64_t E[3];
            int inside = 1;
            for (int i = 0; i < 3 && inside; ++i) {
              const int a = (i + 1) % 3, b = (i + 2) % 3;
              const int64_t dx = t.X[b] - t.X[a], dy = t.Y[b] - t.Y[a];
              E[i] = dx * (py - t.Y[a]) - dy * (px - t.X[a]);
              inside = edge_inside(E[i], dx, dy);
            }
            if (!inside) continue;
            const float l0 = (float)E[0] / fS, l1 = (float)E[1] / fS, l2 = (float)E[2] / fS;
            const float z = (l0 * t.z[0] + l1 * t.z[1]) + l2 * t.z[2];
            const size_t o = (size_t)y * (size_t)W + (size_t)x;
            if (!(z >= 0.0f && z < 1.0f)) continue;      /* depth clip (far) */
            if (!(z < depth[o])) continue;               /* LESS; ties keep the earlier triangle */
            depth[o] = z;
            if (prim_id) prim_id[o] = seq;
            const float p0 = l0 * t.iw[0], p1 = l1 * t.iw[1], p2 = l2 * t.iw[2];
            const float inv = 1.0f / ((p0 + p1) + p2);
            float at[14];
            for (int i = 0; i < 14; ++i) at[i] = ((p0 * t.v[0]->a[i] + p1 * t.v[1]->a[i]) + p2 * t.v[2]->a[i]) * inv;
            float op[4], on[4], oa[4];
            pbro_geom_fragment(at, d->color, &d->color_tex, &d->normal_tex, op, on, oa);
            for (int c = 0; c < 4; ++c) {
              positions[4 * o + c] = pbro_float_to_half(op[c]);
              normals[4 * o + c] = pbro_float_to_half(on[c]);
              albedo[4 * o + c] = to_unorm16(oa[c]);
            }
            if (orm) { orm[4 * o] = to_unorm8(d->occlusion); orm[4 * o + 1] = to_unorm8(d->roughness); orm[4 * o + 2] = to_unorm8(d->metallic); orm[4 * o + 3] = 0; }
            if (emissive) { for (int c = 0; c < 3; ++c) emissive[4 * o + c] = pbro_float_to_half(d->emissive[c]); emissive[4 * o + 3] = 0; }
          }
      }
    }
  }
}
