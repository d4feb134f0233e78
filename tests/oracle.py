"""ctypes bindings of the CPU oracle (oracle/libpbr_oracle.so) and of the reference's own shaders
compiled for the CPU (oracle/_ref/libpbr_ref_shaders.so).  TEST INFRASTRUCTURE: imported only by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np
import sys


ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))
ORACLE_DIR = ROOT / "oracle"
ORACLE_SO = ORACLE_DIR / "libpbr_oracle.so"
REF_SO = ORACLE_DIR / "_ref" / "libpbr_ref_shaders.so"
REFERENCE = Path("/root/reference")

vp, sz, ci, cf = C.c_void_p, C.c_size_t, C.c_int, C.c_float


class Light(C.Structure):
    _fields_ = [("px", cf), ("py", cf), ("pz", cf), ("range", cf), ("r", cf), ("g", cf), ("b", cf), ("intensity", cf)]


class Ibl(C.Structure):
    _fields_ = [("irradiance", vp), ("irr_res", ci), ("prefiltered", vp), ("pref_res", ci), ("pref_mips", ci),
                ("lut", vp), ("lut_res", ci)]


class Texture(C.Structure):
    _fields_ = [("texels", vp), ("width", ci), ("height", ci), ("pitch", sz)]


class Draw(C.Structure):
    """pbro_draw: one drawIndexed of the geometry pass (PbrRenderSystem.cpp:441-461)."""
    _fields_ = [("vertices", vp), ("indices", vp), ("first_index", C.c_uint32), ("index_count", C.c_uint32),
                ("vertex_offset", C.c_int32), ("model", cf * 16), ("normal_model", cf * 12), ("color", cf * 4),
                ("color_tex", Texture), ("normal_tex", Texture), ("occlusion", cf), ("roughness", cf), ("metallic", cf),
                ("emissive", cf * 3)]


def build_oracle(force: bool = False) -> Path:
    if force or not ORACLE_SO.exists() or any(
            p.stat().st_mtime > ORACLE_SO.stat().st_mtime
            for p in [ORACLE_DIR / n for n in ("pbr_oracle.c", "pbr_oracle_ibl.c", "pbr_oracle_raster.c", "pbr_oracle.h", "pbr_detmath.h")]):
        subprocess.run(["make", "-C", str(ORACLE_DIR), "oracle"], check=True, capture_output=True)
    return ORACLE_SO


def build_ref() -> Path | None:
    """Compiles the reference's GLSL for the host when /root/reference is mounted (the CPU container);
    on the GPU box the prebuilt oracle/_ref/*.so that travelled with the snapshot is used as is."""
    if (REFERENCE / "assets" / "shaders").is_dir():
        subprocess.run(["make", "-C", str(ORACLE_DIR), "ref"], check=True, capture_output=True)
    return REF_SO if REF_SO.exists() else None


def _p(a: np.ndarray):
    return a.ctypes.data_as(vp)


class Oracle:
    def __init__(self):
        self.lib = C.CDLL(str(build_oracle()))
        L = self.lib
        L.pbro_half_to_float.restype = cf
        L.pbro_half_to_float.argtypes = [C.c_uint16]
        L.pbro_float_to_half.restype = C.c_uint16
        L.pbro_float_to_half.argtypes = [cf]
        L.pbro_cube_chain_texels.restype = sz
        L.pbro_cube_chain_texels.argtypes = [ci, ci]
        L.pbro_cube_mip_offset.restype = sz
        L.pbro_cube_mip_offset.argtypes = [ci, ci]
        L.pbro_lighting_image.argtypes = [vp, vp, vp, sz, ci, ci, ci, ci, vp, vp, vp, sz, ci]
        L.pbro_tonemap_image.argtypes = [vp, sz, ci, ci, ci, ci, vp, sz, ci]
        L.pbro_brdf_lut.argtypes = [ci, ci, vp, ci]
        L.pbro_equirect_to_cube.argtypes = [vp, ci, ci, ci, vp, ci]
        L.pbro_cube_mipgen.argtypes = [vp, ci, ci, ci]
        L.pbro_cube_sample.argtypes = [vp, ci, ci, vp, cf, vp]
        L.pbro_prefilter_table.argtypes = [cf, ci, ci, ci, vp]
        L.pbro_prefilter_table.restype = ci
        L.pbro_prefilter_texel.argtypes = [vp, ci, ci, ci, ci, ci, ci, ci, vp]
        L.pbro_prefilter_ggx.argtypes = [vp, ci, ci, ci, vp, ci, ci, ci]
        L.pbro_irradiance_texel.argtypes = [vp, ci, ci, ci, ci, ci, ci, ci, vp]
        L.pbro_irradiance.argtypes = [vp, ci, ci, ci, ci, vp, ci]
        L.pbro_shade_image.argtypes = [vp, vp, vp, vp, vp, sz, sz, ci, ci, ci, ci, vp, vp, ci, vp, vp, vp, sz, ci]
        L.pbro_hammersley_table.argtypes = [C.c_uint32, vp]
        L.pbro_sincos2pi.argtypes = [cf, C.POINTER(cf), C.POINTER(cf)]
        L.pbro_blinn_phong.argtypes = [vp, vp, vp, vp, vp]
        L.pbro_aces_tonemap.argtypes = [vp, vp]
        L.pbro_gamma_correct.argtypes = [vp, cf, vp]
        L.pbro_max_threads.restype = ci
        L.pbro_cube_texel_dir.argtypes = [ci, ci, ci, ci, vp]
        L.pbro_cube_dir_to_face_uv.argtypes = [vp, vp, vp, vp]
        L.pbro_brdf_lut_texel.argtypes = [ci, ci, ci, ci, vp]
        L.pbro_tonemap_pixel.argtypes = [vp, vp]
        L.pbro_geom_vertex.argtypes = [vp] * 7
        L.pbro_geom_fragment.argtypes = [vp, vp, C.POINTER(Texture), C.POINTER(Texture), vp, vp, vp]
        L.pbro_geometry_pass.argtypes = [C.POINTER(Draw), ci, vp, vp, ci, ci, vp, vp, vp, vp, vp, vp, vp]

    def max_threads(self) -> int:
        return self.lib.pbro_max_threads()

    # ---- geometry pass ----
    def geom_vertex(self, model, normal_model, view, proj, vertex):
        clip, attrs = np.zeros(4, np.float32), np.zeros(14, np.float32)
        a = [np.ascontiguousarray(v, np.float32) for v in (model, normal_model, view, proj, vertex)]
        self.lib.pbro_geom_vertex(_p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(a[4]), _p(clip), _p(attrs))
        return clip, attrs

    def geom_fragment(self, attrs, color, color_tex, normal_tex):
        o = [np.zeros(4, np.float32) for _ in range(3)]
        ct, nt = _texture(color_tex), _texture(normal_tex)
        self.lib.pbro_geom_fragment(_p(np.ascontiguousarray(attrs, np.float32)), _p(np.ascontiguousarray(color, np.float32)),
                                    C.byref(ct), C.byref(nt), _p(o[0]), _p(o[1]), _p(o[2]))
        return o

    def geometry_pass(self, draws, view, proj, W, H, extended=False):
        """draws: list of dicts (see make_draw).  Returns the planes the reference's geometry pass writes."""
        arr = (Draw * max(1, len(draws)))()
        for i, d in enumerate(draws):
            fill_draw(arr[i], d, lambda a: a.ctypes.data)
        out = dict(pos=np.zeros((H, W, 4), np.uint16), nrm=np.zeros((H, W, 4), np.uint16), alb=np.zeros((H, W, 4), np.uint16),
                   depth=np.zeros((H, W), np.float32), prim=np.zeros((H, W), np.uint32))
        if extended:
            out.update(orm=np.zeros((H, W, 4), np.uint8), emi=np.zeros((H, W, 4), np.uint16))
        self.lib.pbro_geometry_pass(arr, len(draws), _p(np.ascontiguousarray(view, np.float32)), _p(np.ascontiguousarray(proj, np.float32)),
                                    W, H, _p(out["pos"]), _p(out["nrm"]), _p(out["alb"]), _p(out["depth"]),
                                    _p(out["orm"]) if extended else None, _p(out["emi"]) if extended else None, _p(out["prim"]))
        return out

    # ---- Tier A ----
    def lighting(self, pos, nrm, alb, cam, rect=None, nthreads=0):
        """pos/nrm: (H,W,4) float16, alb: (H,W,4) uint16.  Returns (f32 (h,w,4), f16-bits full image)."""
        H, W = pos.shape[:2]
        x0, y0, w, h = rect or (0, 0, W, H)
        cam = np.asarray(cam, np.float32)
        o32 = np.zeros((h, w, 4), np.float32)
        o16 = np.zeros((H, W, 4), np.uint16)
        self.lib.pbro_lighting_image(_p(pos), _p(nrm), _p(alb), W * 8, x0, y0, w, h, _p(cam), _p(o32), _p(o16), W * 8, nthreads)
        return o32, o16

    def tonemap(self, hdr16, rect=None, nthreads=0):
        """hdr16: (H,W,4) uint16 (fp16 bits).  Returns (H,W,4) uint8."""
        H, W = hdr16.shape[:2]
        x0, y0, w, h = rect or (0, 0, W, H)
        out = np.zeros((H, W, 4), np.uint8)
        self.lib.pbro_tonemap_image(_p(hdr16), W * 8, x0, y0, w, h, _p(out), W * 4, nthreads)
        return out

    # ---- Tier B ----
    def hammersley(self, n):
        out = np.zeros((n, 2), np.float32)
        self.lib.pbro_hammersley_table(n, _p(out))
        return out

    def brdf_lut(self, res, nsamples, nthreads=0):
        out = np.zeros((res, res, 2), np.float32)
        self.lib.pbro_brdf_lut(res, nsamples, _p(out), nthreads)
        return out

    def chain_texels(self, base_res, nmips):
        return self.lib.pbro_cube_chain_texels(base_res, nmips)

    def mip_offset(self, base_res, mip):
        return self.lib.pbro_cube_mip_offset(base_res, mip)

    def equirect_to_cube(self, env, res, nthreads=0):
        sh, sw = env.shape[:2]
        out = np.zeros((6, res, res, 4), np.uint16)
        self.lib.pbro_equirect_to_cube(_p(np.ascontiguousarray(env, np.float32)), sw, sh, res, _p(out), nthreads)
        return out

    def make_chain(self, base, nmips, nthreads=0):
        res = base.shape[1]
        chain = np.zeros((self.chain_texels(res, nmips), 4), np.uint16)
        chain[: 6 * res * res] = base.reshape(-1, 4)
        self.lib.pbro_cube_mipgen(_p(chain), res, nmips, nthreads)
        return chain

    def cube_sample(self, chain, base_res, nmips, d, lod):
        out = np.zeros(4, np.float32)
        d = np.asarray(d, np.float32)
        self.lib.pbro_cube_sample(_p(chain), base_res, nmips, _p(d), lod, _p(out))
        return out

    def prefilter(self, chain, base_res, nmips, nsamples, first=0, last=None, nthreads=0):
        dst = np.zeros_like(chain)
        self.lib.pbro_prefilter_ggx(_p(chain), base_res, nmips, nsamples, _p(dst), first, nmips - 1 if last is None else last, nthreads)
        return dst

    def prefilter_texel(self, chain, base_res, nmips, mip, nsamples, face, x, y):
        out = np.zeros(4, np.float32)
        self.lib.pbro_prefilter_texel(_p(chain), base_res, nmips, mip, nsamples, face, x, y, _p(out))
        return out

    def irradiance(self, chain, base_res, nmips, res, nsamples, nthreads=0):
        out = np.zeros((6, res, res, 4), np.uint16)
        self.lib.pbro_irradiance(_p(chain), base_res, nmips, res, nsamples, _p(out), nthreads)
        return out

    def shade(self, pos, nrm, alb, orm, emi, cam, lights, ibl=None, rect=None, nthreads=0):
        """lights: (n,8) float32 rows (px,py,pz,range,r,g,b,intensity); ibl: dict(irr, pref, pref_mips, lut)."""
        H, W = pos.shape[:2]
        x0, y0, w, h = rect or (0, 0, W, H)
        cam = np.asarray(cam, np.float32)
        lights = np.ascontiguousarray(lights, np.float32).reshape(-1, 8)
        o32 = np.zeros((h, w, 4), np.float32)
        o16 = np.zeros((H, W, 4), np.uint16)
        iblp = None
        if ibl is not None:
            s = Ibl()
            s.irradiance, s.irr_res = ibl["irr"].ctypes.data, ibl["irr"].shape[1]
            s.prefiltered, s.pref_res, s.pref_mips = ibl["pref"].ctypes.data, ibl["pref_res"], ibl["pref_mips"]
            s.lut, s.lut_res = ibl["lut"].ctypes.data, ibl["lut"].shape[0]
            iblp = C.byref(s)
        self.lib.pbro_shade_image(_p(pos), _p(nrm), _p(alb), _p(orm), _p(emi), W * 8, W * 4, x0, y0, w, h, _p(cam),
                                  _p(lights), len(lights), iblp, _p(o32), _p(o16), W * 8, nthreads)
        return o32, o16


class RefShaders:
    """The reference's assets/shaders compiled for the CPU (oracle/_ref)."""

    def __init__(self):
        path = build_ref()
        if path is None:
            raise FileNotFoundError("oracle/_ref/libpbr_ref_shaders.so is missing and /root/reference is not mounted")
        self.lib = C.CDLL(str(path))
        self.lib.pbr_ref_lighting_image.argtypes = [vp, vp, vp, vp, ci, ci, sz, sz, vp, vp, vp, sz, ci]
        self.lib.pbr_ref_tonemap_image.argtypes = [vp, sz, ci, ci, vp, sz, ci]
        self.lib.pbr_ref_blinn_phong.argtypes = [vp, vp, vp, vp, vp]
        self.lib.pbr_ref_gamma_correct.argtypes = [vp, cf, vp]
        self.lib.pbr_ref_aces_tonemap.argtypes = [vp, vp]
        self.lib.pbr_ref_fullscreen_vertex.argtypes = [ci, vp]
        self.lib.pbr_ref_geom_vertex.argtypes = [vp] * 7
        self.lib.pbr_ref_geom_fragment.argtypes = [vp, vp, vp, ci, ci, sz, vp, ci, ci, sz, vp, vp, vp]

    def geom_vertex(self, model, normal_model, view, proj, vertex):
        """assets/shaders/geometry_pass/vertex.glsl::main, compiled for the CPU."""
        clip, attrs = np.zeros(4, np.float32), np.zeros(14, np.float32)
        a = [np.ascontiguousarray(v, np.float32) for v in (model, normal_model, view, proj, vertex)]
        self.lib.pbr_ref_geom_vertex(_p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(a[4]), _p(clip), _p(attrs))
        return clip, attrs

    def geom_fragment(self, attrs, color, color_tex, normal_tex):
        """assets/shaders/geometry_pass/fragment.glsl::main, compiled for the CPU."""
        o = [np.zeros(4, np.float32) for _ in range(3)]
        self.lib.pbr_ref_geom_fragment(_p(np.ascontiguousarray(attrs, np.float32)), _p(np.ascontiguousarray(color, np.float32)),
                                       _p(color_tex), color_tex.shape[1], color_tex.shape[0], color_tex.strides[0],
                                       _p(normal_tex), normal_tex.shape[1], normal_tex.shape[0], normal_tex.strides[0],
                                       _p(o[0]), _p(o[1]), _p(o[2]))
        return o

    def lighting(self, pos, nrm, alb, cam, nthreads=0):
        H, W = pos.shape[:2]
        cam = np.asarray(cam, np.float32)
        depth = np.ones((H, W), np.float32)
        o32 = np.zeros((H, W, 4), np.float32)
        o16 = np.zeros((H, W, 4), np.uint16)
        self.lib.pbr_ref_lighting_image(_p(pos), _p(nrm), _p(alb), _p(depth), W, H, W * 8, W * 4, _p(cam), _p(o32), _p(o16), W * 8, nthreads)
        return o32, o16

    def tonemap(self, hdr16, nthreads=0):
        H, W = hdr16.shape[:2]
        out = np.zeros((H, W, 4), np.uint8)
        self.lib.pbr_ref_tonemap_image(_p(hdr16), W * 8, W, H, _p(out), W * 4, nthreads)
        return out

    def blinn_phong(self, frag, N, V, L):
        out = np.zeros(4, np.float32)
        a = [np.asarray(v, np.float32) for v in (frag, N, V, L)]
        self.lib.pbr_ref_blinn_phong(_p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(out))
        return out

    def gamma_correct(self, rgb, gamma):
        out = np.zeros(3, np.float32)
        self.lib.pbr_ref_gamma_correct(_p(np.asarray(rgb, np.float32)), gamma, _p(out))
        return out

    def aces(self, rgb):
        out = np.zeros(3, np.float32)
        self.lib.pbr_ref_aces_tonemap(_p(np.asarray(rgb, np.float32)), _p(out))
        return out

    def fullscreen_vertex(self, i):
        out = np.zeros(6, np.float32)
        self.lib.pbr_ref_fullscreen_vertex(i, _p(out))
        return out


def _texture(img: np.ndarray) -> Texture:
    assert img.dtype == np.uint8 and img.ndim == 3 and img.shape[2] == 4 and img.flags.c_contiguous
    return Texture(img.ctypes.data, img.shape[1], img.shape[0], img.strides[0])


def fill_draw(dst, d: dict, addr):
    """Fills a pbro_draw / pbrcu_draw-shaped ctypes struct from a draw dict; `addr(x)` maps the dict's arrays
    (numpy on the host, device buffers on the GPU) to addresses."""
    dst.vertices, dst.indices = addr(d["vertices"]), addr(d["indices"])
    dst.first_index, dst.index_count, dst.vertex_offset = d["first_index"], d["index_count"], d["vertex_offset"]
    dst.model[:] = np.asarray(d["model"], np.float32).reshape(-1).tolist()
    dst.normal_model[:] = np.asarray(d["normal_model"], np.float32).reshape(-1).tolist()
    dst.color[:] = [float(c) for c in d["color"]]
    for name in ("color_tex", "normal_tex"):
        t, img = getattr(dst, name), d[name]
        t.texels, t.width, t.height, t.pitch = addr(img), d[name + "_shape"][1], d[name + "_shape"][0], d[name + "_shape"][1] * 4
    dst.occlusion, dst.roughness, dst.metallic = d.get("occlusion", 1.0), d.get("roughness", 1.0), d.get("metallic", 1.0)
    dst.emissive[:] = [float(c) for c in d.get("emissive", (0.0, 0.0, 0.0))]


# ---- column-major 4x4 maths the way glm builds them (CameraData.hpp:23-32, ModelPushConstant.hpp:33-46); float64
# here, rounded to fp32 once: these are INPUTS of the geometry pass, identical for the oracle and the device ----
from pbr_b200.synth import (checker_texture, col_major, look_at_rh, mesh_vertex, model_push_constant, perspective_rh_no,  # noqa: E402,F401
                            synth_environment, synth_gbuffer, synth_lights, uv_sphere)


def make_draw(vertices, indices, model=None, color=(1, 1, 1, 1), color_tex=None, normal_tex=None, first_index=0, index_count=None,
              vertex_offset=0, **pbr):
    if model is None:
        model = model_push_constant()
    if color_tex is None:
        color_tex = np.full((1, 1, 4), 255, np.uint8)
    if normal_tex is None:
        normal_tex = np.array([[[128, 128, 255, 255]]], np.uint8)
    vertices = np.ascontiguousarray(vertices, np.float32).reshape(-1, 12)
    indices = np.ascontiguousarray(indices, np.uint16).reshape(-1)
    d = dict(vertices=vertices, indices=indices, first_index=first_index, index_count=len(indices) - first_index if index_count is None else index_count,
             vertex_offset=vertex_offset, model=model[0], normal_model=model[1], color=color,
             color_tex=np.ascontiguousarray(color_tex), normal_tex=np.ascontiguousarray(normal_tex))
    d["color_tex_shape"], d["normal_tex_shape"] = d["color_tex"].shape, d["normal_tex"].shape
    d.update(pbr)
    return d


def raster_scene(seed=1, n_spheres=6):
    """A small seeded scene for the geometry pass: textured spheres over a ground quad."""
    rng = np.random.default_rng(seed)
    sv, si = uv_sphere(1.0, 12, 24)
    draws = []
    for k in range(n_spheres):
        m = model_push_constant(translation=rng.uniform(-4, 4, 3), scale=rng.uniform(0.5, 1.5, 3), rot_axis=rng.standard_normal(3),
                                rot_angle=rng.uniform(0, 6.28))
        draws.append(make_draw(sv, si, m, color=rng.uniform(0.3, 1.0, 4), color_tex=checker_texture(16, 8, seed + k),
                               normal_tex=checker_texture(8, 8, seed + 100 + k, normal_map=True),
                               occlusion=float(rng.uniform(0.5, 1)), roughness=float(rng.uniform(0.05, 1)), metallic=float(rng.integers(0, 2)),
                               emissive=tuple(rng.uniform(0, 0.5, 3))))
    g = 30.0
    quad = np.array([mesh_vertex((-g, 2.0, -g), (0, -1, 0), (1, 0, 0, 1), (0, 0)), mesh_vertex((g, 2.0, -g), (0, -1, 0), (1, 0, 0, 1), (8, 0)),
                     mesh_vertex((g, 2.0, g), (0, -1, 0), (1, 0, 0, 1), (8, 8)), mesh_vertex((-g, 2.0, g), (0, -1, 0), (1, 0, 0, 1), (0, 8))], np.float32)
    draws.append(make_draw(quad, [0, 1, 2, 0, 2, 3], color=(0.8, 0.8, 0.8, 1.0), color_tex=checker_texture(32, 32, seed + 50)))
    draws.append(make_draw(quad, [0, 2, 1, 0, 3, 2], color=(0.8, 0.8, 0.8, 1.0), color_tex=checker_texture(32, 32, seed + 50)))  # both windings: one is culled
    return draws


# ---- seeded synthetic inputs shared by tests, smoke and bench's cpu legs --------------------------
