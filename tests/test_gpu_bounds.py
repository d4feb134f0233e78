"""GPU bounds checks of our own (compute-sanitizer is closed on this pool): every output of the hot path is
placed between two guard bands filled with a byte pattern, at extents that are NOT multiples of the CTA /
tile / vector sizes, and the bands must come back untouched.  Values are covered by the parity tests; this
file only asks "did a kernel write outside the extent its C-ABI call was given"."""
import ctypes as C

import numpy as np
import pytest
from oracle import synth_environment, synth_gbuffer, synth_lights

pytestmark = pytest.mark.gpu
GUARD = 8192
PATTERN = 0xA5


class Guarded:
    """nbytes of device memory with GUARD pattern bytes on either side; `ptr` is the start of the payload."""

    def __init__(self, dev, nbytes, fill=None):
        self.nbytes = nbytes
        host = np.full(nbytes + 2 * GUARD, PATTERN, np.uint8)
        if fill is not None:
            host[GUARD:GUARD + nbytes] = np.ascontiguousarray(fill).view(np.uint8).reshape(-1)
        self.buf = dev.upload(host)
        self.ptr = self.buf.ptr + GUARD

    def check(self, what):
        host = self.buf.download((self.nbytes + 2 * GUARD,), np.uint8)
        lo, hi = host[:GUARD], host[GUARD + self.nbytes:]
        assert (lo == PATTERN).all(), f"{what}: {np.count_nonzero(lo != PATTERN)} bytes written BEFORE the buffer"
        assert (hi == PATTERN).all(), f"{what}: {np.count_nonzero(hi != PATTERN)} bytes written AFTER the buffer (first at +{int(np.argmax(hi != PATTERN))})"
        return host[GUARD:GUARD + self.nbytes]

    def free(self):
        self.buf.free()


def test_ibl_precompute_outputs_stay_inside(pbr, dev, oracle):
    """K4-K9 at extents that leave partial CTAs in every grid dimension."""
    lib, ctx = dev.lib, dev.ctx
    ham = Guarded(dev, 33 * 8)
    lib.ibl_hammersley(ctx, 33, ham.ptr, None)
    lut = Guarded(dev, 24 * 24 * 8)
    lib.ibl_brdf_lut(ctx, 24, 64, lut.ptr, None)
    dev.sync()
    ham.check("hammersley n=33"); lut.check("brdf_lut 24^2")

    env = synth_environment(100, 50, 2)  # odd-sized lat-long source
    src = dev.upload(env)
    base = Guarded(dev, 6 * 24 * 24 * 8)  # K4 alone takes any resolution
    lib.ibl_equirect_to_cube(ctx, src.ptr, 100, 50, 24, base.ptr, None)
    dev.sync()
    texels = base.check("equirect_to_cube res=24").view(np.uint16)
    assert texels.any() and np.isfinite(texels.view(np.float16).astype(np.float32)).all()

    for res, nmips in ((8, 4), (16, 5), (64, 7)):
        n = oracle.chain_texels(res, nmips)
        chain = Guarded(dev, n * 8)
        lib.ibl_equirect_to_cube(ctx, src.ptr, 100, 50, res, chain.ptr, None)
        lib.ibl_cube_mipgen(ctx, chain.ptr, res, nmips, None)
        dev.sync()
        host_chain = chain.check(f"mipgen {res}/{nmips}").view(np.uint16).reshape(n, 4)
        assert np.array_equal(host_chain, oracle.make_chain(host_chain[: 6 * res * res].reshape(6, res, res, 4), nmips))
        for exact in (0, 1):
            lib.ibl_set_exact_filtering(ctx, exact)
            pref = Guarded(dev, n * 8)
            lib.ibl_prefilter_ggx(ctx, chain.ptr, res, nmips, 64, pref.ptr, 0, nmips - 1, 0, 1, None)
            irr = Guarded(dev, 6 * 12 * 12 * 8)
            lib.ibl_irradiance(ctx, chain.ptr, res, nmips, 12, 64, irr.ptr, None)
            dev.sync()
            pref.check(f"prefilter {res}/{nmips} exact={exact}"); irr.check(f"irradiance 12^2 from {res} exact={exact}")
            # a middle shard of three writes only its own row blocks: everything else keeps the pattern
            part = Guarded(dev, n * 8)
            lib.ibl_prefilter_ggx(ctx, chain.ptr, res, nmips, 64, part.ptr, 1, nmips - 1, 1, 3, None)
            dev.sync()
            got, whole = part.check(f"prefilter shard {res}/{nmips}").view(np.uint16).reshape(n, 4), pref.check("").view(np.uint16).reshape(n, 4)
            written = (got != (PATTERN | PATTERN << 8)).any(axis=1)
            assert np.array_equal(got[written], whole[written]) and not written[: 6 * res * res].any()
            pref.free(); irr.free(); part.free()
        lib.ibl_set_exact_filtering(ctx, 0)
        chain.free()
    for b in (ham, lut, base, src):
        b.free()


@pytest.mark.parametrize("w,h", [(1, 1), (37, 19), (130, 17), (255, 33)])
def test_frame_outputs_stay_inside(pbr, dev, oracle, w, h):
    """K1, K2, fused K1+K2, K3c and K3 (one list / tile lists / no IBL) at ragged extents with tight pitches."""
    lib, ctx = dev.lib, dev.ctx
    g = synth_gbuffer(w, h, seed=3 * w + h, tier_b=True)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    cam = pbr.f3(0.5, -1.0, 12.0)
    hdr, ldr, hdr2, ldr2 = Guarded(dev, w * h * 8), Guarded(dev, w * h * 4), Guarded(dev, w * h * 8), Guarded(dev, w * h * 4)
    lib.lighting_blinnphong(ctx, C.byref(gb), cam, pbr.WHOLE, hdr.ptr, w * 8, None)
    lib.tonemap_aces_gamma(ctx, hdr.ptr, w * 8, w, h, pbr.WHOLE, ldr.ptr, w * 4, None)
    lib.lighting_tonemap_fused(ctx, C.byref(gb), cam, pbr.WHOLE, hdr2.ptr, w * 8, ldr2.ptr, w * 4, None)
    dev.sync()
    a, b = hdr.check("K1"), ldr.check("K2")
    assert np.array_equal(a, hdr2.check("fused HDR")) and np.array_equal(b, ldr2.check("fused LDR"))

    nl = 9
    lights = synth_lights(nl, seed=4)
    ld = dev.upload(lights)
    tiles = pbr.LightTiles()
    tiles.tiles_x, tiles.tiles_y, tiles.max_per_tile = (w + 15) // 16, (h + 15) // 16, 4  # fewer slots than lights: truncation
    nt = tiles.tiles_x * tiles.tiles_y
    counts, indices = Guarded(dev, nt * 4), Guarded(dev, nt * 4 * 2)
    tiles.counts, tiles.indices = counts.ptr, indices.ptr
    lib.cluster_lights(ctx, C.byref(gb), ld.ptr, nl, C.byref(tiles), None)
    dev.sync()
    assert (counts.check("K3c counts").view(np.uint32) <= 4).all()
    indices.check("K3c indices")

    from test_gpu_tier_b import build_ibl, make_ibl_struct
    host, bufs = build_ibl(dev, oracle, ns=64)
    ibl = make_ibl_struct(pbr, host, bufs, dev, accel=True)
    outs = []
    for what, ibl_arg, tile_arg in (("K3 one list", C.byref(ibl), None), ("K3 tile lists", C.byref(ibl), C.byref(tiles)), ("K3 no IBL", None, None)):
        out = Guarded(dev, w * h * 8)
        lib.shade_ggx_ibl(ctx, C.byref(gb), cam, ld.ptr, nl, tile_arg, ibl_arg, pbr.WHOLE, out.ptr, w * 8, None)
        dev.sync()
        outs.append(out.check(f"{what} {w}x{h}"))
        out.free()
    assert outs[0].any() or g["bg"].all()
    lib.ibl_accel_free(ctx, ibl.accel)
    for x in (hdr, ldr, hdr2, ldr2, counts, indices, ld, *bufs):
        x.free()
