"""GPU: closed-form cases on the CUDA path itself -- evidence for Tier B that does not pass through the in-repo oracle (the
reference has no GGX / IBL code: parity unpinned vs the reference).  Every case is checked on every K3 kernel a caller can reach
(device light list -> shade_ggx_one_kernel, host light list -> shade_ggx_cl_kernel, tile lists -> shade_ggx_kernel, plain textures
without footprint records) and in both filter modes; the expected values come from the published formulas (glTF 2.0 Appendix B,
KHR_lights_punctual, Karis 2013) evaluated in numpy.
Tolerances: default path 1e-3 + 2^-9 |ref| (irradiance and LUT through the texture unit's 8-bit weights), exact path
1e-3 + 2^-10 |ref| (one fp16 ulp of the output format)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H = 32, 16
ROUGH_BYTES = [255, 200, 128, 90, 64, 32, 16, 12]


def constant_ibl(pbr, dev, level=0.5, res=16, nmips=5, ires=8, lres=32, ns=128, accel=True):
    """An IBL set computed ON THE DEVICE from a constant environment: box mip chain, GGX prefilter, irradiance, LUT."""
    base = np.full((6, res, res, 4), level, np.float16).view(np.uint16)
    n_chain = dev.lib.cube_chain_texels(res, nmips)
    chain = dev.alloc(n_chain * 8)
    dev.lib.memcpy_h2d(chain.ptr, base.ctypes.data, base.nbytes, None)
    dev.sync()
    dev.lib.ibl_cube_mipgen(dev.ctx, chain.ptr, res, nmips, None)
    pref, irr, lut = dev.alloc(n_chain * 8), dev.alloc(6 * ires * ires * 8), dev.alloc(lres * lres * 8)
    dev.lib.ibl_prefilter_ggx(dev.ctx, chain.ptr, res, nmips, ns, pref.ptr, 0, nmips - 1, 0, 1, None)
    dev.lib.ibl_irradiance(dev.ctx, chain.ptr, res, nmips, ires, ns, irr.ptr, None)
    dev.lib.ibl_brdf_lut(dev.ctx, lres, ns, lut.ptr, None)
    dev.sync()
    s = pbr.Ibl()
    s.irradiance, s.irr_res = irr.ptr, ires
    s.prefiltered, s.pref_res, s.pref_mips = pref.ptr, res, nmips
    s.lut, s.lut_res = lut.ptr, lres
    if accel:
        a = C.c_void_p()
        dev.lib.ibl_accel_build(dev.ctx, C.byref(s), C.byref(a), None)
        s.accel = a
    host = dict(chain=chain.download((n_chain, 4), np.uint16), pref=pref.download((n_chain, 4), np.uint16), irr=irr.download((6, ires, ires, 4), np.uint16),
                lut=lut.download((lres, lres, 2), np.float32))
    return s, host


def f16(a):
    return a.view(np.float16).astype(np.float32)


@pytest.mark.parametrize("exact", [0, 1])
def test_constant_environment_stays_constant(pbr, dev, exact):
    """Box mips, GGX prefilter (every level) and cosine irradiance of a constant environment are that constant."""
    dev.lib.ibl_set_exact_filtering(dev.ctx, exact)
    try:
        _, host = constant_ibl(pbr, dev, level=0.75, accel=False)
    finally:
        dev.lib.ibl_set_exact_filtering(dev.ctx, 0)
    assert (f16(host["chain"])[:, :3] == 0.75).all(), "box mips of a constant are exact"
    assert np.abs(f16(host["pref"])[:, :3] - 0.75).max() <= 1e-3
    assert np.abs(f16(host["irr"])[..., :3] - 0.75).max() <= 1e-3


def gbuffer_plane(dev, pbr, rough_bytes, metal=0, base=(1.0, 1.0, 1.0, 0.5), occ=255, emi=(0.0, 0.0, 0.0), spread=3.0):
    """N = +Z everywhere, P = (x, 0, 0) with x across the row (so N.V varies under a camera on the Z axis), one roughness per row."""
    pos = np.zeros((H, W, 4), np.float16)
    pos[..., 0] = np.linspace(-spread, spread, W, dtype=np.float32)[None, :]
    nrm = np.zeros((H, W, 4), np.float16)
    nrm[..., 2] = 1.0
    alb = np.empty((H, W, 4), np.uint16)
    alb[...] = np.round(np.array(base) * 65535).astype(np.uint16)
    orm = np.zeros((H, W, 4), np.uint8)
    orm[..., 0], orm[..., 2] = occ, metal
    orm[..., 1] = np.array([rough_bytes[y % len(rough_bytes)] for y in range(H)], np.uint8)[:, None]
    emi_p = np.zeros((H, W, 4), np.float16)
    emi_p[..., :3] = emi
    g = dict(pos=pos, nrm=nrm, alb=alb, orm=orm, emi=emi_p)
    return g, dev.gbuffer(pos, nrm, alb, orm=orm, emissive=emi_p)


def shade_paths(pbr, dev, gb, cam, lights, ibl_accel, ibl_plain):
    """The same frame through every kernel path: yields (name, exact, HDR frame as float32)."""
    nl = len(lights)
    out = dev.alloc(W * H * 8)
    ld = dev.upload(lights) if nl else None
    tiles = pbr.LightTiles()
    dev.lib.light_tiles_alloc(dev.ctx, W, H, max(nl, 1), C.byref(tiles))

    def run(ptr, ibl, use_tiles=False):
        dev.lib.memset(out.ptr, 0x5A, W * H * 8, None)
        if use_tiles:
            dev.lib.cluster_lights(dev.ctx, C.byref(gb), ptr, nl, C.byref(tiles), None)
        dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ptr, nl, C.byref(tiles) if use_tiles else None, C.byref(ibl) if ibl else None, pbr.WHOLE,
                              out.ptr, W * 8, None)
        dev.sync()
        return f16(out.download((H, W, 4), np.uint16))

    for exact in (0, 1):
        dev.lib.ibl_set_exact_filtering(dev.ctx, exact)
        try:
            yield "device list (shade_ggx_one_kernel)", exact, run(ld.ptr if ld else None, ibl_accel)
            yield "host list (shade_ggx_cl_kernel)", exact, run(lights.ctypes.data if nl else None, ibl_accel)
            if nl:
                yield "tile lists (shade_ggx_kernel)", exact, run(ld.ptr, ibl_accel, use_tiles=True)
        finally:
            dev.lib.ibl_set_exact_filtering(dev.ctx, 0)
    yield "plain textures (shade_ggx_kernel, four-load filter)", 1, run(ld.ptr if ld else None, ibl_plain)
    dev.lib.light_tiles_free(dev.ctx, C.byref(tiles))


def within(got, want, exact):
    bound = 1e-3 + 2.0 ** (-10 if exact else -9) * np.abs(want)
    return int((np.abs(got - want) > bound).sum()), float((np.abs(got - want) / bound).max())


def lut_lookup(lut, u, v):
    """The LUT's clamp-to-edge bilinear in numpy fp32 (u = N.V, v = roughness)."""
    res = lut.shape[0]
    fx, fy = np.float32(u) * res - np.float32(0.5), np.float32(v) * res - np.float32(0.5)
    x0, y0 = np.floor(fx), np.floor(fy)
    tx, ty = (fx - x0)[..., None], (fy - y0)[..., None]
    c = lambda a: np.clip(a.astype(int), 0, res - 1)
    t00, t10, t01, t11 = lut[c(y0), c(x0)], lut[c(y0), c(x0 + 1)], lut[c(y0 + 1), c(x0)], lut[c(y0 + 1), c(x0 + 1)]
    top, bot = t00 + tx * (t10 - t00), t01 + tx * (t11 - t01)
    return top + ty * (bot - top)


def test_white_furnace_on_the_gpu(pbr, dev):
    """Constant environment L, no lights, dielectric white base: out = L (c_diff + F0 A + B) occlusion + emissive, (A, B) from the
    (bit-exact) split-sum LUT at (N.V, roughness) -- over a sweep of N.V and roughness, on every kernel path."""
    L, occ_byte, emi = 0.5, 204, (0.25, 0.125, 0.0625)
    ibl_a, host = constant_ibl(pbr, dev, level=L)
    ibl_p, _ = constant_ibl(pbr, dev, level=L, accel=False)
    g, gb = gbuffer_plane(dev, pbr, ROUGH_BYTES, occ=occ_byte, emi=emi)
    cam = (0.0, 0.0, 2.0)
    x = g["pos"][..., 0].astype(np.float32)
    nov = np.clip(cam[2] / np.sqrt(x * x + cam[2] ** 2), 1e-4, 1.0).astype(np.float32)
    rough = np.clip(g["orm"][..., 1].astype(np.float32) / 255.0, 0.045, 1.0)
    ab = lut_lookup(host["lut"], nov, rough)
    env = np.float32(f16(host["pref"])[0, 0])  # the device's own (constant) prefiltered value: L up to its fp16 rounding
    irr = np.float32(f16(host["irr"])[0, 0, 0, 0])
    occ = occ_byte / 255.0
    want = np.empty((H, W, 4), np.float32)
    for c in range(3):
        want[..., c] = (irr * 1.0 + env * (0.04 * ab[..., 0] + ab[..., 1])) * occ + emi[c]
    want[..., 3] = 0.5
    assert abs(env - L) <= 1e-3 and abs(irr - L) <= 1e-3
    for name, exact, got in shade_paths(pbr, dev, gb, cam, np.zeros((0, 8), np.float32), ibl_a, ibl_p):
        bad, worst = within(got, want, exact)
        assert bad == 0, f"white furnace, {name}, exact={exact}: {bad} values outside the tolerance (worst {worst:.2f}x)"
    dev.lib.ibl_accel_free(dev.ctx, ibl_a.accel)


def test_normal_incidence_light_on_the_gpu(pbr, dev):
    """One unlimited-range point light straight above a surface seen from straight above (N = V = L = H): D = 1/(pi a^2), Vis = 1/4,
    F = F0, so out = ((1 - F0) c_diff / pi + F0 / (4 pi a^2)) I / d^2 -- for dielectrics and metals over a roughness sweep, incl. the
    smoothest surface the material model allows (roughness 0.045: a^2 = 4.1e-6, the case the |N x H|^2 form of D exists for)."""
    for metal_byte, base in ((0, (1.0, 1.0, 1.0, 1.0)), (255, (1.0, 0.5, 0.25, 1.0))):
        g, gb = gbuffer_plane(dev, pbr, ROUGH_BYTES, metal=metal_byte, base=base, spread=0.0)  # every pixel the same point
        cam, d, intensity = (0.0, 0.0, 5.0), 2.0, 10.0
        lights = np.array([[0, 0, d, 0, 1.0, 0.5, 0.25, intensity]], np.float32)
        rough = np.clip(g["orm"][..., 1].astype(np.float64) / 255.0, 0.045, 1.0)
        a2 = rough ** 4
        E = intensity / d ** 2
        m = metal_byte / 255.0
        base_q = np.round(np.array(base) * 65535) / 65535
        want = np.empty((H, W, 4), np.float64)
        for c, lc in enumerate((1.0, 0.5, 0.25)):
            f0 = 0.04 * (1 - m) + base_q[c] * m
            cdiff = base_q[c] * (1 - m)
            want[..., c] = np.minimum(((1 - f0) * cdiff / np.pi + f0 / (4 * np.pi * a2)) * E * lc, 65504.0)
        want[..., 3] = base_q[3]
        for name, exact, got in shade_paths(pbr, dev, gb, cam, lights, None, None):
            # the kernels use single MUFU approximations (rsqrt, rcp: 1-2 ulp fp32 each) on a chain ending in 1 / a^4-sized terms:
            # allow 2^-9 relative in both modes here (the oracle-parity tests hold the tighter bound on ordinary inputs)
            bad, worst = within(got, want.astype(np.float32), 0)
            assert bad == 0, f"normal incidence, metal={metal_byte}, {name}: {bad} values outside the tolerance (worst {worst:.2f}x)"


def test_cleared_frame_is_exactly_zero(pbr, dev):
    """A G-buffer with the reference's clear values (all planes 0) shades to exactly 0 in all four channels, whatever the lights/IBL."""
    ibl_a, _ = constant_ibl(pbr, dev, level=2.0)
    ibl_p, _ = constant_ibl(pbr, dev, level=2.0, accel=False)
    z16, z8 = np.zeros((H, W, 4), np.uint16), np.zeros((H, W, 4), np.uint8)
    gb = dev.gbuffer(z16.view(np.float16), z16.view(np.float16), z16, orm=z8, emissive=z16.view(np.float16))
    lights = np.array([[0, 0, 2, 0, 1, 1, 1, 100], [1, 1, 1, 3, 1, 0, 0, 50]], np.float32)
    for name, exact, got in shade_paths(pbr, dev, gb, (0.0, 0.0, 0.0), lights, ibl_a, ibl_p):  # camera AT the cleared positions
        assert (got.view(np.uint32) == 0).all(), f"cleared frame, {name}, exact={exact}: not exactly zero"
    dev.lib.ibl_accel_free(dev.ctx, ibl_a.accel)
