"""CPU tests of the Tier-B oracle (published formulas; NO reference counterpart -> parity unpinned vs. the
reference).  These pin the oracle to closed-form / textbook properties so that "GPU == oracle" means
something: Hammersley and the deterministic sincos, cube-map addressing round trips, white-furnace
behaviour of the prefilter / irradiance, analytic limits of the BRDF LUT, and the shading model."""
import ctypes as C

import numpy as np
import pytest
from oracle import Light, synth_environment


def test_hammersley_and_radical_inverse(oracle):
    h = oracle.hammersley(1024)
    assert np.array_equal(h[:, 0], np.arange(1024, dtype=np.float32) / 1024)
    assert list(h[:5, 1]) == [0.0, 0.5, 0.25, 0.75, 0.125]
    assert len(set(h[:, 1].tolist())) == 1024 and h[:, 1].max() < 1
    assert oracle.lib.pbro_radical_inverse_bits(1) & 0xFFFFFFFF == 0x80000000


def test_sincos2pi_accuracy(oracle):
    s, c = C.c_float(), C.c_float()
    worst = 0.0
    for t in np.linspace(0, 1, 4097, endpoint=False, dtype=np.float32):
        oracle.lib.pbro_sincos2pi(float(t), C.byref(s), C.byref(c))
        worst = max(worst, abs(s.value - np.sin(2 * np.pi * float(t))), abs(c.value - np.cos(2 * np.pi * float(t))))
    assert worst < 3e-7


def test_cube_addressing_round_trip(oracle):
    res = 8
    d = (C.c_float * 3)()
    face, u, v = C.c_int(), C.c_float(), C.c_float()
    for f in range(6):
        for y in range(res):
            for x in range(res):
                oracle.lib.pbro_cube_texel_dir(f, x, y, res, d)
                oracle.lib.pbro_cube_dir_to_face_uv(d, C.byref(face), C.byref(u), C.byref(v))
                assert face.value == f
                assert abs(u.value * res - 0.5 - x) < 1e-5 and abs(v.value * res - 0.5 - y) < 1e-5
    # axes hit face centres in the +X,-X,+Y,-Y,+Z,-Z order
    for f, axis in enumerate([(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]):
        oracle.lib.pbro_cube_dir_to_face_uv((C.c_float * 3)(*axis), C.byref(face), C.byref(u), C.byref(v))
        assert (face.value, u.value, v.value) == (f, 0.5, 0.5)


def test_cube_sample_is_exact_at_texel_centres_and_linear_in_lod(oracle):
    res, nmips = 16, 5
    rng = np.random.default_rng(3)
    base = rng.random((6, res, res, 4)).astype(np.float16).view(np.uint16)
    chain = oracle.make_chain(base, nmips)
    d = (C.c_float * 3)()
    for f, x, y in [(0, 0, 0), (3, 7, 9), (5, 15, 15)]:
        oracle.lib.pbro_cube_texel_dir(f, x, y, res, d)
        got = oracle.cube_sample(chain, res, nmips, list(d), 0.0)
        assert np.array_equal(got, base[f, y, x].view(np.float16).astype(np.float32))
    a, b = oracle.cube_sample(chain, res, nmips, [0.3, -0.5, 0.8], 1.0), oracle.cube_sample(chain, res, nmips, [0.3, -0.5, 0.8], 2.0)
    mid = oracle.cube_sample(chain, res, nmips, [0.3, -0.5, 0.8], 1.25)
    assert np.allclose(mid, a + 0.25 * (b - a), atol=1e-6)
    # box mips preserve the mean exactly up to fp16 rounding
    top = chain[oracle.mip_offset(res, nmips - 1):].view(np.float16).astype(np.float32).reshape(6, 4)
    assert np.allclose(top, base.view(np.float16).astype(np.float32).mean(axis=(1, 2)), atol=2e-3)


def test_white_furnace(oracle):
    """A constant environment must prefilter / irradiance-convolve to the same constant at every level."""
    res, nmips = 16, 5
    base = np.full((6, res, res, 4), 0.75, np.float16).view(np.uint16)
    chain = oracle.make_chain(base, nmips)
    pref = oracle.prefilter(chain, res, nmips, 64).view(np.float16).astype(np.float32)
    assert np.allclose(pref[:, :3], 0.75, atol=1e-3)
    irr = oracle.irradiance(chain, res, nmips, 4, 64).view(np.float16).astype(np.float32)
    assert np.allclose(irr[..., :3], 0.75, atol=1e-3)


def test_equirect_to_cube_directions(oracle):
    env = np.zeros((64, 128, 4), np.float32)
    env[:32, :, 0] = 1.0   # upper hemisphere (+Y) red
    env[32:, :, 2] = 1.0   # lower hemisphere blue
    cube = oracle.equirect_to_cube(env, 8).view(np.float16).astype(np.float32)
    assert np.allclose(cube[2, ..., 0], 1, atol=1e-3) and np.allclose(cube[2, ..., 2], 0, atol=1e-3)  # +Y face
    assert np.allclose(cube[3, ..., 2], 1, atol=1e-3) and np.allclose(cube[3, ..., 0], 0, atol=1e-3)  # -Y face


def test_brdf_lut_limits(oracle):
    lut = oracle.brdf_lut(32, 256)
    assert lut.min() >= 0 and (lut.sum(-1) <= 1.0001).all()
    # smooth + facing: scale -> 1, bias -> 0 (Karis 2013 fig. 6); rough + grazing loses energy
    assert lut[0, -1, 0] > 0.95 and lut[0, -1, 1] < 0.02
    assert lut[-1, 0].sum() < lut[0, -1].sum()
    # deterministic: the same texel recomputed alone is bit-identical
    one = np.zeros(2, np.float32)
    oracle.lib.pbro_brdf_lut_texel(32, 256, 5, 9, one.ctypes.data_as(C.c_void_p))
    assert np.array_equal(one.view(np.uint32), lut[9, 5].view(np.uint32))


def shade_one(oracle, base, orm, n=(0, 0, 1), p=(0, 0, 0), cam=(0, 0, 5), lights=(), emi=(0, 0, 0), ibl=None):
    pos = np.zeros((1, 1, 4), np.float16); pos[0, 0, :3] = p
    nrm = np.zeros((1, 1, 4), np.float16); nrm[0, 0, :3] = n
    alb = np.zeros((1, 1, 4), np.uint16); alb[0, 0] = [int(round(c * 65535)) for c in base]
    o = np.zeros((1, 1, 4), np.uint8); o[0, 0] = orm
    e = np.zeros((1, 1, 4), np.float16); e[0, 0, :3] = emi
    f32, _ = oracle.shade(pos, nrm, alb, o, e, cam, np.array(lights, np.float32).reshape(-1, 8), ibl)
    return f32[0, 0]


def test_shading_model(oracle):
    white, rough_diel = (1, 1, 1, 0.5), (255, 255, 0, 255)
    light = [0, 0, 2, 0, 1, 1, 1, 10]  # straight above the surface point, d = 2
    # no light, no IBL: emissive only, alpha = base alpha; cleared texel -> 0
    assert np.allclose(shade_one(oracle, white, rough_diel, emi=(0.5, 0.25, 0.125)), [0.5, 0.25, 0.125, 0.5], atol=1e-3)
    assert (shade_one(oracle, white, rough_diel, n=(0, 0, 0)) == 0).all()
    # inverse-square law: 2x the distance -> 1/4 the radiance (N.L = N.V = 1 in both)
    a = shade_one(oracle, white, rough_diel, lights=[light])
    far = [0, 0, 4, 0, 1, 1, 1, 10]
    b = shade_one(oracle, white, rough_diel, lights=[far], cam=(0, 0, 5))
    assert np.allclose(a[:3] / b[:3], 4.0, rtol=1e-4)
    # rough dielectric lit head-on: Lambert term c_diff/pi * (1-F0) * E  +  specular D*Vis*F0 * E, E = I/d^2
    E, a2 = 10 / 4, 1.0
    expect = (1 - 0.04) / np.pi * E + (a2 / (np.pi)) * (0.5 / 2.0) * 0.04 * E
    assert np.allclose(a[:3], expect, rtol=1e-4)
    # range window: a light whose range ends before the surface contributes nothing; lights add linearly
    assert (shade_one(oracle, white, rough_diel, lights=[[0, 0, 2, 1.5, 1, 1, 1, 10]])[:3] == 0).all()
    two = shade_one(oracle, white, rough_diel, lights=[light, far])
    assert np.allclose(two[:3], a[:3] + b[:3], rtol=1e-5)
    # back-facing light contributes nothing; metals have no diffuse term
    assert (shade_one(oracle, white, rough_diel, lights=[[0, 0, -2, 0, 1, 1, 1, 10]])[:3] == 0).all()
    metal = shade_one(oracle, (1, 0.5, 0.25, 1), (255, 255, 255, 255), lights=[light])
    assert np.allclose(metal[:3], (a2 / np.pi) * 0.25 * np.array([1, 0.5, 0.25]) * E, rtol=2e-3)


def test_shading_ibl_white_furnace(oracle):
    """Constant environment L, metallic 0, base white: ambient = L*(c_diff + F0*A + B) * occlusion."""
    res, nmips = 8, 4
    chain = oracle.make_chain(np.full((6, res, res, 4), 0.5, np.float16).view(np.uint16), nmips)
    ibl = dict(irr=oracle.irradiance(chain, res, nmips, 4, 32), pref=oracle.prefilter(chain, res, nmips, 32), pref_res=res,
               pref_mips=nmips, lut=oracle.brdf_lut(16, 128))
    out = shade_one(oracle, (1, 1, 1, 1), (255, 128, 0, 255), ibl=ibl, cam=(0, 3, 4))
    nov, rough = 0.8, 128 / 255
    lut = ibl["lut"]
    u, v = nov * 16 - 0.5, rough * 16 - 0.5
    x0, y0 = int(np.floor(u)), int(np.floor(v)); tx, ty = u - x0, v - y0
    ab = (lut[y0, x0] * (1 - tx) + lut[y0, x0 + 1] * tx) * (1 - ty) + (lut[y0 + 1, x0] * (1 - tx) + lut[y0 + 1, x0 + 1] * tx) * ty
    assert np.allclose(out[:3], 0.5 * (1.0 + 0.04 * ab[0] + ab[1]), rtol=2e-3)
