"""physically-based-renderer_b200 -- B200-native (sm_100a) shading hot path behind a C ABI.

This Python layer is only a ctypes binding of ``lib/libpbrcu.so`` (include/pbrcu.h) for tests and
benchmarks; the product is the CUDA library and the C++ facade in ``include/pbr``.  There is no CPU
fallback: importing works without a GPU (so symbols can be checked), but every compute call needs a
CUDA device, and a missing library raises immediately.

The directory name has a hyphen, so import it through the root-level ``pbr_b200`` shim::

    import pbr_b200 as pbr
    lib = pbr.load()
"""
from __future__ import annotations

import ctypes as C
import re
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent
LIB_PATH = HERE / "lib" / "libpbrcu.so"
HEADER = ROOT / "include" / "pbrcu.h"

PBRCU_TILE = 16
IBL_EXACT, IBL_TEXTURE_UNIT, IBL_GATHER = 1, 2, 4  # pbrcu_ibl.flags


class PbrcuError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"pbrcu status {code}: {message}")
        self.code = code


class GBuffer(C.Structure):
    _fields_ = [
        ("positions", C.c_void_p), ("normals", C.c_void_p), ("albedo", C.c_void_p),
        ("depth", C.c_void_p), ("orm", C.c_void_p), ("emissive", C.c_void_p),
        ("width", C.c_uint32), ("height", C.c_uint32),
        ("pitch8", C.c_size_t), ("pitch4", C.c_size_t),
    ]


class Rect(C.Structure):
    _fields_ = [("x0", C.c_uint32), ("y0", C.c_uint32), ("w", C.c_uint32), ("h", C.c_uint32)]


class Light(C.Structure):
    _fields_ = [("px", C.c_float), ("py", C.c_float), ("pz", C.c_float), ("range", C.c_float),
                ("r", C.c_float), ("g", C.c_float), ("b", C.c_float), ("intensity", C.c_float)]


class Ibl(C.Structure):
    _fields_ = [("irradiance", C.c_void_p), ("irr_res", C.c_uint32),
                ("prefiltered", C.c_void_p), ("pref_res", C.c_uint32), ("pref_mips", C.c_uint32),
                ("lut", C.c_void_p), ("lut_res", C.c_uint32), ("accel", C.c_void_p), ("flags", C.c_uint32)]


class Texture(C.Structure):
    _fields_ = [("texels", C.c_void_p), ("width", C.c_uint32), ("height", C.c_uint32), ("pitch", C.c_size_t)]


class Material(C.Structure):
    _fields_ = [("color", C.c_float * 4), ("color_tex", Texture), ("normal_tex", Texture), ("occlusion", C.c_float),
                ("roughness", C.c_float), ("metallic", C.c_float), ("emissive", C.c_float * 3)]


class Draw(C.Structure):
    """pbrcu_draw: one drawIndexed of the geometry pass."""
    _fields_ = [("vertices", C.c_void_p), ("indices", C.c_void_p), ("first_index", C.c_uint32), ("index_count", C.c_uint32),
                ("vertex_offset", C.c_int32), ("vertex_count", C.c_uint32), ("model", C.c_float * 16), ("normal_model", C.c_float * 12), ("material", Material)]


class LightTiles(C.Structure):
    _fields_ = [("counts", C.c_void_p), ("indices", C.c_void_p),
                ("tiles_x", C.c_uint32), ("tiles_y", C.c_uint32), ("max_per_tile", C.c_uint32)]


def declared_symbols(header: Path = HEADER) -> list[str]:
    """Every function the C-ABI header declares (used by the 'library exports all symbols' test)."""
    text = re.sub(r"/\*.*?\*/", "", header.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(pbrcu_[a-z0-9_]+)\s*\(", text)))


_vp, _sz, _u32, _u64, _i, _f3 = C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint64, C.c_int, C.POINTER(C.c_float)
_PG, _PI, _PT = C.POINTER(GBuffer), C.POINTER(Ibl), C.POINTER(LightTiles)

_SIGNATURES = {
    "pbrcu_version": (C.c_int, []),
    "pbrcu_last_error": (C.c_char_p, []),
    "pbrcu_device_count": (_i, [C.POINTER(C.c_int)]),
    "pbrcu_ctx_create": (_i, [_i, C.POINTER(_vp)]),
    "pbrcu_ctx_destroy": (None, [_vp]),
    "pbrcu_ctx_sm_count": (_i, [_vp]),
    "pbrcu_malloc": (_i, [_vp, _sz, C.POINTER(_vp)]),
    "pbrcu_free": (_i, [_vp, _vp]),
    "pbrcu_host_alloc": (_i, [_sz, C.POINTER(_vp)]),
    "pbrcu_host_free": (_i, [_vp]),
    "pbrcu_memcpy_h2d": (_i, [_vp, _vp, _sz, _vp]),
    "pbrcu_memcpy_d2h": (_i, [_vp, _vp, _sz, _vp]),
    "pbrcu_memset": (_i, [_vp, _i, _sz, _vp]),
    "pbrcu_memset_2d": (_i, [_vp, _sz, _i, _sz, _sz, _vp]),
    "pbrcu_memcpy_h2d_2d": (_i, [_vp, _sz, _vp, _sz, _sz, _sz, _vp]),
    "pbrcu_stream_create": (_i, [C.POINTER(_vp)]),
    "pbrcu_stream_destroy": (_i, [_vp]),
    "pbrcu_stream_synchronize": (_i, [_vp]),
    "pbrcu_timer_create": (_i, [C.POINTER(_vp)]),
    "pbrcu_timer_destroy": (_i, [_vp]),
    "pbrcu_timer_start": (_i, [_vp, _vp]),
    "pbrcu_timer_stop": (_i, [_vp, _vp]),
    "pbrcu_timer_elapsed_ms": (_i, [_vp, C.POINTER(C.c_float)]),
    "pbrcu_gbuffer_alloc": (_i, [_vp, _u32, _u32, _i, _PG]),
    "pbrcu_gbuffer_free": (_i, [_vp, _PG]),
    "pbrcu_gbuffer_clear": (_i, [_PG, _vp]),
    "pbrcu_image_alloc": (_i, [_vp, _u32, _u32, _u32, C.POINTER(_vp), C.POINTER(_sz)]),
    "pbrcu_lighting_blinnphong": (_i, [_vp, _PG, _f3, Rect, _vp, _sz, _vp]),
    "pbrcu_tonemap_aces_gamma": (_i, [_vp, _vp, _sz, _u32, _u32, Rect, _vp, _sz, _vp]),
    "pbrcu_lighting_tonemap_fused": (_i, [_vp, _PG, _f3, Rect, _vp, _sz, _vp, _sz, _vp]),
    "pbrcu_lighting_blinnphong_batched": (_i, [_vp, _PG, _sz, _sz, _u32, _f3, _vp, _sz, _sz, _vp]),
    "pbrcu_lighting_tonemap_fused_batched": (_i, [_vp, _PG, _sz, _u32, _f3, _vp, _sz, _sz, _vp, _sz, _sz, _vp]),
    "pbrcu_frame_host": (_i, [_vp, _vp, _vp, _vp, _u32, _u32, _f3, _vp, _vp, _vp]),
    "pbrcu_cube_chain_texels": (_sz, [_u32, _u32]),
    "pbrcu_cube_mip_offset": (_sz, [_u32, _u32]),
    "pbrcu_shade_ggx_ibl": (_i, [_vp, _PG, _f3, _vp, _u32, _PT, _PI, Rect, _vp, _sz, _vp]),
    "pbrcu_shade_ggx_ibl_multicast": (_i, [_vp, _PG, _f3, _vp, _u32, _PT, _PI, Rect, _vp, _sz, _vp]),
    "pbrcu_shade_ggx_ibl_replicated": (_i, [_vp, _PG, _f3, _vp, _u32, _PT, _PI, Rect, _vp, _sz, C.POINTER(_vp), _u32, _vp]),
    "pbrcu_shade_ggx_ibl_ldr": (_i, [_vp, _PG, _f3, _vp, _u32, _PT, _PI, Rect, _vp, _sz, C.POINTER(_vp), _u32, _vp]),
    "pbrcu_shade_ggx_ibl_batched": (_i, [_vp, _PG, _sz, _sz, _u32, _f3, _vp, _u32, _PI, _vp, _sz, _sz, _vp]),
    "pbrcu_shade_ggx_ibl_batched_static": (_i, [_vp, _PG, _sz, _sz, _u32, _f3, _vp, _u32, _PI, _vp, _sz, _sz, _vp]),
    "pbrcu_shade_frame_host": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _u32, _u32, _f3, _vp, _u32, _PI, _vp, _u32, _vp]),
    "pbrcu_ibl_accel_build": (_i, [_vp, _PI, C.POINTER(_vp), _vp]),
    "pbrcu_ibl_accel_free": (_i, [_vp, _vp]),
    "pbrcu_ibl_accel_bytes": (_sz, [_vp]),
    "pbrcu_light_tiles_alloc": (_i, [_vp, _u32, _u32, _u32, _PT]),
    "pbrcu_light_tiles_free": (_i, [_vp, _PT]),
    "pbrcu_cluster_lights": (_i, [_vp, _PG, _vp, _u32, _PT, _vp]),
    "pbrcu_ibl_hammersley": (_i, [_vp, _u32, _vp, _vp]),
    "pbrcu_ibl_brdf_lut": (_i, [_vp, _u32, _u32, _vp, _vp]),
    "pbrcu_ibl_equirect_to_cube": (_i, [_vp, _vp, _u32, _u32, _u32, _vp, _vp]),
    "pbrcu_ibl_cube_mipgen": (_i, [_vp, _vp, _u32, _u32, _vp]),
    "pbrcu_ibl_set_exact_filtering": (_i, [_vp, _i]),
    "pbrcu_shade_set_texture_unit": (_i, [_vp, _i]),
    "pbrcu_geometry_pass": (_i, [_vp, C.POINTER(GBuffer), _vp, _vp, C.POINTER(Draw), C.c_uint32, _vp, _vp]),
    "pbrcu_geometry_pass_batched": (_i, [_vp, C.POINTER(GBuffer), _sz, _sz, _u32, _vp, _vp, C.POINTER(Draw), _u32, _vp]),
    "pbrcu_ibl_prefilter_ggx": (_i, [_vp, _vp, _u32, _u32, _u32, _vp, _u32, _u32, _u32, _u32, _vp]),
    "pbrcu_ibl_prefilter_ggx_replicated": (_i, [_vp, _vp, _u32, _u32, _u32, _vp, _u32, _u32, _u32, _u32, C.POINTER(_vp), _u32, _vp]),
    "pbrcu_ibl_source_create": (_i, [_vp, _vp, _u32, _u32, _vp, C.POINTER(_vp)]),
    "pbrcu_ibl_source_free": (_i, [_vp, _vp]),
    "pbrcu_ibl_prefilter_ggx_prepared": (_i, [_vp, _vp, _u32, _vp, _u32, _u32, _u32, _u32, C.POINTER(_vp), _u32, _vp]),
    "pbrcu_ibl_irradiance_prepared": (_i, [_vp, _vp, _u32, _u32, _vp, _vp]),
    "pbrcu_ibl_irradiance": (_i, [_vp, _vp, _u32, _u32, _u32, _u32, _vp, _vp]),
    "pbrcu_mg_create": (_i, [C.POINTER(C.c_int), _i, C.POINTER(_vp)]),
    "pbrcu_mg_destroy": (None, [_vp]),
    "pbrcu_mg_size": (_i, [_vp]),
    "pbrcu_mg_device": (_i, [_vp, _i]),
    "pbrcu_mg_ctx": (_vp, [_vp, _i]),
    "pbrcu_mg_stream": (_vp, [_vp, _i]),
    "pbrcu_mg_set_device": (_i, [_vp, _i]),
    "pbrcu_mg_barrier": (_i, [_vp]),
    "pbrcu_mg_synchronize": (_i, [_vp]),
    "pbrcu_mg_row_bands": (_i, [_u32, _i, _f3, C.POINTER(_u32)]),
    "pbrcu_mg_shade_frame": (_i, [_vp, _PG, _f3, _vp, _u32, _PT, _PI, C.POINTER(_u32), C.POINTER(_vp), _sz, _i]),
    "pbrcu_mg_shade_frame_ldr": (_i, [_vp, _PG, _f3, _vp, _u32, _PT, _PI, C.POINTER(_u32), C.POINTER(_vp), _sz, _i]),
    "pbrcu_mg_shade_views": (_i, [_vp, _PG, _sz, _sz, C.POINTER(_u32), C.POINTER(_f3), _vp, _u32, _PI, C.POINTER(_vp), _sz, _sz]),
    "pbrcu_mg_geometry_views": (_i, [_vp, _PG, _sz, _sz, C.POINTER(_u32), C.POINTER(_f3), C.POINTER(_f3), C.POINTER(C.POINTER(Draw)), C.POINTER(_u32)]),
    "pbrcu_mg_prefilter": (_i, [_vp, C.POINTER(_vp), _u32, _u32, C.POINTER(_vp), _i]),
    "pbrcu_synth_gbuffer": (_i, [_vp, _PG, _i, _u64, _f3, _f3, _vp]),
    "pbrcu_synth_environment": (_i, [_vp, _vp, _u32, _u32, _u64, _vp]),
}

_NO_STATUS = {"pbrcu_mg_destroy", "pbrcu_mg_size", "pbrcu_mg_device", "pbrcu_mg_ctx", "pbrcu_mg_stream", "pbrcu_version", "pbrcu_last_error", "pbrcu_ctx_destroy", "pbrcu_ctx_sm_count",
              "pbrcu_cube_chain_texels", "pbrcu_cube_mip_offset", "pbrcu_ibl_accel_bytes"}


class Lib:
    """ctypes view of libpbrcu.so.  ``lib.<name>(...)`` calls ``pbrcu_<name>`` and raises PbrcuError
    on a non-zero status."""

    def __init__(self, path: Path = LIB_PATH):
        if not Path(path).exists():
            raise FileNotFoundError(
                f"{path} is missing: build it with `python physically-based-renderer_b200/build.py` "
                "(there is no CPU fallback for the shading path)")
        self.path = Path(path)
        self.cdll = C.CDLL(str(path))
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(self.cdll, name)  # AttributeError if the library lacks a declared symbol
            fn.restype = res
            fn.argtypes = args

    def last_error(self) -> str:
        return self.cdll.pbrcu_last_error().decode()

    def __getattr__(self, short: str):
        name = "pbrcu_" + short
        if name not in _SIGNATURES:
            raise AttributeError(short)
        fn = getattr(self.cdll, name)
        if name in _NO_STATUS:
            return fn

        def call(*a):
            rc = fn(*a)
            if rc != 0:
                raise PbrcuError(rc, self.last_error())
            return rc

        call.__name__ = name
        return call


_LIB: Lib | None = None


def load() -> Lib:
    global _LIB
    if _LIB is None:
        _LIB = Lib()
    return _LIB


def f3(x, y, z):
    return (C.c_float * 3)(float(x), float(y), float(z))


WHOLE = Rect(0, 0, 0, 0)

from .device import Device, DeviceBuffer  # noqa: E402  (numpy-level conveniences for tests / bench)
from . import partition  # noqa: E402
from . import synth  # noqa: E402  (seeded synthetic inputs for tests / bench)
