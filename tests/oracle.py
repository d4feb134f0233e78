"""ctypes bindings of the CPU oracle (oracle/libpbr_oracle.so) and of the reference's own shaders
compiled for the CPU (oracle/_ref/libpbr_ref_shaders.so).  TEST INFRASTRUCTURE: imported only by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
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


def build_oracle(force: bool = False) -> Path:
    if force or not ORACLE_SO.exists() or any(
            p.stat().st_mtime > ORACLE_SO.stat().st_mtime
            for p in [ORACLE_DIR / n for n in ("pbr_oracle.c", "pbr_oracle_ibl.c", "pbr_oracle.h", "pbr_detmath.h")]):
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

    def max_threads(self) -> int:
        return self.lib.pbro_max_threads()

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


# ---- seeded synthetic inputs shared by tests, smoke and bench's cpu legs --------------------------
def synth_gbuffer(w, h, seed, background=0.2, tier_b=False, cam_at_origin_bg=True):
    """A seeded G-buffer in the reference's formats: unit normals, positions in [-10,10]^3, random
    UNORM16 albedo, `background` fraction of cleared (all-zero) texels."""
    rng = np.random.default_rng(seed)
    pos = np.zeros((h, w, 4), np.float16)
    pos[..., :3] = rng.uniform(-10, 10, (h, w, 3))
    n = rng.standard_normal((h, w, 3))
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    nrm = np.zeros((h, w, 4), np.float16)
    nrm[..., :3] = n
    alb = rng.integers(0, 65536, (h, w, 4), dtype=np.uint16)
    bg = rng.random((h, w)) < background
    pos[bg] = 0
    nrm[bg] = 0
    alb[bg] = 0
    out = dict(pos=pos, nrm=nrm, alb=alb, bg=bg)
    if tier_b:
        orm = np.zeros((h, w, 4), np.uint8)
        orm[..., 0] = rng.integers(128, 256, (h, w))
        orm[..., 1] = rng.integers(0, 256, (h, w))
        m = rng.random((h, w))
        orm[..., 2] = np.where(m < 0.45, 0, np.where(m < 0.9, 255, rng.integers(0, 256, (h, w))))
        orm[..., 3] = 255
        emi = np.zeros((h, w, 4), np.float16)
        e = rng.random((h, w)) < 0.05
        emi[e, :3] = rng.uniform(0, 4, (int(e.sum()), 3))
        orm[bg] = 0
        emi[bg] = 0
        out.update(orm=orm, emi=emi)
    return out


def synth_environment(sw, sh, seed, sun=50.0):
    """Small procedural lat-long HDR sky: gradient + bright sun disc + noise (RGBA32F)."""
    rng = np.random.default_rng(seed)
    v = (np.arange(sh) + 0.5) / sh
    u = (np.arange(sw) + 0.5) / sw
    theta = v[:, None] * np.pi
    phi = (u[None, :] - 0.5) * 2 * np.pi
    d = np.stack([np.sin(theta) * np.cos(phi), np.cos(theta) * np.ones_like(phi), np.sin(theta) * np.sin(phi)], -1)
    t = 0.5 * (d[..., 1] + 1)
    env = np.ones((sh, sw, 4), np.float32)
    env[..., 0] = 0.15 + 0.3 * t
    env[..., 1] = 0.2 + 0.5 * t
    env[..., 2] = 0.25 + 1.4 * t
    env[..., :3] *= (1 + 0.1 * (rng.random((sh, sw, 1)) - 0.5)).astype(np.float32)
    s = np.array([0.35, 0.8, 0.45])
    s /= np.linalg.norm(s)
    env[(d @ s) > np.cos(np.radians(4.0)), :3] = sun
    return env


def synth_lights(n, seed, ranged=False):
    rng = np.random.default_rng(seed)
    L = np.zeros((n, 8), np.float32)
    L[:, :3] = rng.uniform(-8, 8, (n, 3))
    L[:, 3] = rng.uniform(2, 6, n) if ranged else 0.0
    L[:, 4:7] = rng.uniform(0.2, 1.0, (n, 3))
    L[:, 7] = rng.uniform(50, 200, n)
    return L
