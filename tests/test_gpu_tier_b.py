"""GPU parity, Tier B: GGX+IBL shading and the IBL precompute vs the in-repo CPU oracle of the published
formulas.  NO REFERENCE COUNTERPART (the reference shades with Blinn-Phong and has no IBL) -- parity is
"unpinned vs. the reference, pinned vs. oracle/pbr_oracle_ibl.c".  Bit-exact: Hammersley, BRDF LUT,
cube mip chain.  Tolerance elsewhere: |err| <= 1e-3 + 2^-9 |ref| on the fp16 outputs (fp16 resolves
2^-10 relative; HDR environments exceed the range where 1e-3 absolute is representable), PSNR >= 60 dB."""
import ctypes as C

import numpy as np
import pytest
from conftest import assert_hdr_parity, f16
from oracle import synth_environment, synth_gbuffer, synth_lights

pytestmark = pytest.mark.gpu
REL = 2.0 ** -9


def test_hammersley_bit_exact(dev, oracle):
    for n in (1, 16, 1024, 4096):
        out = dev.alloc(n * 8)
        dev.lib.ibl_hammersley(dev.ctx, n, out.ptr, None)
        got = out.download((n, 2), np.float32)
        assert np.array_equal(got.view(np.uint32), oracle.hammersley(n).view(np.uint32))
        out.free()


@pytest.mark.parametrize("res,ns", [(32, 64), (64, 1024), (512, 1024)])
def test_brdf_lut_bit_exact(dev, oracle, res, ns):
    out = dev.alloc(res * res * 8)
    dev.lib.ibl_brdf_lut(dev.ctx, res, ns, out.ptr, None)
    got = out.download((res, res, 2), np.float32)
    want = oracle.brdf_lut(res, ns)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), f"max diff {np.abs(got - want).max()}"
    assert got.min() >= 0 and got.max() <= 1.0001
    out.free()


def make_env_chain(dev, oracle, res, nmips, seed=1):
    env = synth_environment(256, 128, seed)
    src = dev.upload(env)
    chain_d = dev.alloc(oracle.chain_texels(res, nmips) * 8)
    dev.lib.ibl_equirect_to_cube(dev.ctx, src.ptr, 256, 128, res, chain_d.ptr, None)
    dev.lib.ibl_cube_mipgen(dev.ctx, chain_d.ptr, res, nmips, None)
    dev.sync()
    chain = chain_d.download((oracle.chain_texels(res, nmips), 4), np.uint16)
    return env, chain_d, chain


def test_equirect_to_cube_and_mipgen(dev, oracle):
    res, nmips = 64, 7
    env, chain_d, chain = make_env_chain(dev, oracle, res, nmips)
    base_want = oracle.equirect_to_cube(env, res)
    assert_hdr_parity(chain[: 6 * res * res].reshape(6, res, res, 4), base_want, peak=50.0, rel_tol=REL, what="equirect->cube")
    # mip chain: bit-exact when the oracle starts from the device's own base level
    want_chain = oracle.make_chain(chain[: 6 * res * res].reshape(6, res, res, 4), nmips)
    assert np.array_equal(chain, want_chain), "cube mip chain is not bit-exact"


@pytest.mark.parametrize("exact", [0, 1], ids=["texture-unit", "fp32-weights"])
def test_prefilter_and_irradiance(dev, oracle, exact):
    """Both K6/K7 filter paths (hardware trilinear cube fetches; software fp32 weights) against the oracle."""
    res, nmips, ns = 32, 6, 1024
    env, chain_d, chain = make_env_chain(dev, oracle, res, nmips)
    dev.lib.ibl_set_exact_filtering(dev.ctx, exact)
    dst = dev.alloc(chain.nbytes)
    dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, dst.ptr, 0, nmips - 1, 0, 1, None)
    dev.sync()
    got = dst.download(chain.shape, np.uint16)
    want = oracle.prefilter(chain, res, nmips, ns)
    n0 = 6 * res * res
    assert np.array_equal(got[:n0], chain[:n0]), "level 0 must be an exact copy"
    for m in range(1, nmips):
        a, b = oracle.mip_offset(res, m), oracle.mip_offset(res, m + 1)
        assert_hdr_parity(got[a:b], want[a:b], peak=50.0, rel_tol=REL, what=f"prefilter mip {m}")
    # sharded == unsharded (the multi-GPU partition of every level into row blocks)
    dst2 = dev.alloc(chain.nbytes); dst2.zero()
    for s in range(3):
        dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, dst2.ptr, 0, nmips - 1, s, 3, None)
    dev.sync()
    assert np.array_equal(dst2.download(chain.shape, np.uint16), got)
    # irradiance
    ires = 16
    irr = dev.alloc(6 * ires * ires * 8)
    dev.lib.ibl_irradiance(dev.ctx, chain_d.ptr, res, nmips, ires, ns, irr.ptr, None)
    dev.sync()
    assert_hdr_parity(irr.download((6, ires, ires, 4), np.uint16), oracle.irradiance(chain, res, nmips, ires, ns), peak=50.0, rel_tol=REL, what="irradiance")
    dev.lib.ibl_set_exact_filtering(dev.ctx, 0)


def test_prefilter_texture_unit_on_sharp_environment(dev, oracle):
    """The hardware-filtered prefilter on a 256^2 chain whose base level holds a hard-edged sun (the worst case
    for the unit's 8-bit weights): every level within tolerance of the oracle at randomly chosen texels, and
    within tolerance of the fp32-weight path everywhere."""
    res, nmips, ns = 256, 9, 1024
    env = synth_environment(1024, 512, 3, sun=5.0e3)
    src = dev.upload(env)
    n = oracle.chain_texels(res, nmips)
    chain_d = dev.alloc(n * 8)
    dev.lib.ibl_equirect_to_cube(dev.ctx, src.ptr, 1024, 512, res, chain_d.ptr, None)
    dev.lib.ibl_cube_mipgen(dev.ctx, chain_d.ptr, res, nmips, None)
    dev.sync()
    chain = chain_d.download((n, 4), np.uint16)
    outs = []
    for exact in (0, 1):
        dev.lib.ibl_set_exact_filtering(dev.ctx, exact)
        dst = dev.alloc(n * 8)
        dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, dst.ptr, 0, nmips - 1, 0, 1, None)
        dev.sync()
        outs.append(dst.download((n, 4), np.uint16))
        dst.free()
    dev.lib.ibl_set_exact_filtering(dev.ctx, 0)
    fast, exact_out = outs
    a = oracle.mip_offset(res, 1)
    assert_hdr_parity(fast[a:], exact_out[a:], peak=float(f16(exact_out).max()), rel_tol=REL, what="texture-unit vs fp32-weight prefilter")
    rng = np.random.default_rng(5)
    for m in range(1, nmips):
        r = res >> m
        off = oracle.mip_offset(res, m)
        for _ in range(24):
            f, x, y = int(rng.integers(6)), int(rng.integers(r)), int(rng.integers(r))
            want = oracle.prefilter_texel(chain, res, nmips, m, ns, f, x, y)[:3]
            got = f16(fast[off + (f * r + y) * r + x])[:3]
            assert (np.abs(got - want) <= 1e-3 + REL * np.abs(want)).all(), (m, f, x, y, got, want)


def build_ibl(dev, oracle, res=32, nmips=6, ns=256, ires=16, lres=32):
    env, chain_d, chain = make_env_chain(dev, oracle, res, nmips)
    pref_d = dev.alloc(chain.nbytes)
    dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, pref_d.ptr, 0, nmips - 1, 0, 1, None)
    irr_d = dev.alloc(6 * ires * ires * 8)
    dev.lib.ibl_irradiance(dev.ctx, chain_d.ptr, res, nmips, ires, ns, irr_d.ptr, None)
    lut_d = dev.alloc(lres * lres * 8)
    dev.lib.ibl_brdf_lut(dev.ctx, lres, ns, lut_d.ptr, None)
    dev.sync()
    host = dict(irr=irr_d.download((6, ires, ires, 4), np.uint16), pref=pref_d.download(chain.shape, np.uint16),
                pref_res=res, pref_mips=nmips, lut=lut_d.download((lres, lres, 2), np.float32))
    return host, (irr_d, pref_d, lut_d)


def make_ibl_struct(pbr, host, bufs, dev=None, accel=False, gather=False):
    """accel=True also builds the footprint-record copies (pbrcu_ibl_accel_build) the fast K3 path samples; gather=True the
    texture-gather atlases as well (PBRCU_IBL_GATHER; the flag is cleared again: a test sets it per call)."""
    s = pbr.Ibl()
    s.irradiance, s.irr_res = bufs[0].ptr, host["irr"].shape[1]
    s.prefiltered, s.pref_res, s.pref_mips = bufs[1].ptr, host["pref_res"], host["pref_mips"]
    s.lut, s.lut_res = bufs[2].ptr, host["lut"].shape[0]
    if accel:
        h = C.c_void_p()
        s.flags = pbr.IBL_GATHER if gather else 0
        dev.lib.ibl_accel_build(dev.ctx, C.byref(s), C.byref(h), None)
        s.flags = 0
        s.accel = h
    return s


@pytest.mark.parametrize("w,h,nl,use_ibl", [(1, 1, 0, True), (37, 19, 4, True), (160, 90, 4, False), (256, 144, 4, True), (64, 64, 32, True)])
def test_shade_matches_oracle(pbr, dev, oracle, w, h, nl, use_ibl):
    host, bufs = build_ibl(dev, oracle)
    g = synth_gbuffer(w, h, seed=w + h, tier_b=True)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    lights = synth_lights(nl, seed=9)
    ld = dev.upload(lights) if nl else None
    ibl = make_ibl_struct(pbr, host, bufs) if use_ibl else None
    cam = (0.5, -1.0, 12.0)
    out = dev.alloc(w * h * 8)
    dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ld.ptr if ld else None, nl, None, C.byref(ibl) if ibl else None, pbr.WHOLE, out.ptr, w * 8, None)
    dev.sync()
    got = out.download((h, w, 4), np.uint16)
    _, want = oracle.shade(g["pos"], g["nrm"], g["alb"], g["orm"], g["emi"], cam, lights, host if use_ibl else None)
    peak = float(max(f16(want).max(), 1.0))
    assert_hdr_parity(got, want, peak=peak, rel_tol=REL, what=f"shade {w}x{h} nl={nl}")
    assert (got[g["bg"]] == 0).all()
    if use_ibl:
        ibl2 = make_ibl_struct(pbr, host, bufs, dev, accel=True)
        n_cube = oracle.chain_texels(host["pref_res"], host["pref_mips"]) + host["irr"].size // 4
        # 32-byte footprint records of the three maps + the flagged RGBA16F texture copies of the two cubes + the LUT as a float2
        # texture (no gather atlases: they are built on request only, PBRCU_IBL_GATHER)
        assert dev.lib.ibl_accel_bytes(ibl2.accel) == 32 * (n_cube + host["lut"].size // 2) + 8 * n_cube + 8 * (host["lut"].size // 2)
        out2 = dev.alloc(w * h * 8)
        # opt-in: texture-unit fetches (8-bit weights), fp32-weight records where the contrast flag is set
        dev.lib.shade_set_texture_unit(dev.ctx, 1)
        dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ld.ptr if ld else None, nl, None, C.byref(ibl2), pbr.WHOLE, out2.ptr, w * 8, None)
        dev.sync()
        dev.lib.shade_set_texture_unit(dev.ctx, 0)
        assert_hdr_parity(out2.download((h, w, 4), np.uint16), want, peak=peak, rel_tol=REL, what=f"shade (texture unit) {w}x{h} nl={nl}")
        # default: prefiltered taps from footprint records (fp32 weights), irradiance and LUT through the texture unit
        dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ld.ptr if ld else None, nl, None, C.byref(ibl2), pbr.WHOLE, out2.ptr, w * 8, None)
        dev.sync()
        assert_hdr_parity(out2.download((h, w, 4), np.uint16), want, peak=peak, rel_tol=REL, what=f"shade (records + texture unit) {w}x{h} nl={nl}")
        # fp32 weights everywhere: one 256-bit footprint-record load per bilinear tap, bit-identical to the plain path
        dev.lib.ibl_set_exact_filtering(dev.ctx, 1)
        dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ld.ptr if ld else None, nl, None, C.byref(ibl2), pbr.WHOLE, out2.ptr, w * 8, None)
        dev.sync()
        dev.lib.ibl_set_exact_filtering(dev.ctx, 0)
        got2 = out2.download((h, w, 4), np.uint16)
        assert np.array_equal(got2, got), f"record path differs from the plain path in {np.count_nonzero(got2 != got)} values"
        dev.lib.ibl_accel_free(dev.ctx, ibl2.accel)


@pytest.mark.parametrize("nl", [1, 2, 3, 5, 6, 7, 9])
def test_shade_single_list_kernel_light_counts_and_ragged_rect(pbr, dev, oracle, nl):
    """shade_ggx_one_kernel (one light list + footprint records) takes FOUR lights per iteration and pads the list with null
    lights: every count around the padding, on a rect that starts and ends inside tiles, with a pitched output whose
    texels outside the rect must stay untouched."""
    host, bufs = build_ibl(dev, oracle)
    w, h = 75, 53
    g = synth_gbuffer(w, h, seed=100 + nl, tier_b=True)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    lights = synth_lights(nl, seed=20 + nl)
    ld = dev.upload(lights)
    ibl = make_ibl_struct(pbr, host, bufs, dev, accel=True)
    cam = (0.5, -1.0, 12.0)
    pitch = (w + 5) * 8
    sentinel = np.full((h, w + 5, 4), 0x7B7B, np.uint16)
    out = dev.upload(sentinel)
    x0, y0, rw, rh = 5, 3, 61, 42
    dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ld.ptr, nl, None, C.byref(ibl), pbr.Rect(x0, y0, rw, rh), out.ptr, pitch, None)
    dev.sync()
    got = out.download((h, w + 5, 4), np.uint16)
    _, want = oracle.shade(g["pos"], g["nrm"], g["alb"], g["orm"], g["emi"], cam, lights, host)
    peak = float(max(f16(want).max(), 1.0))
    assert_hdr_parity(got[y0:y0 + rh, x0:x0 + rw], want[y0:y0 + rh, x0:x0 + rw], peak=peak, rel_tol=REL, what=f"single-list kernel nl={nl}")
    outside = np.ones((h, w + 5), bool)
    outside[y0:y0 + rh, x0:x0 + rw] = False
    assert (got[outside] == 0x7B7B).all(), "texels outside the rect were written"
    assert (got[y0:y0 + rh, x0:x0 + rw][g["bg"][y0:y0 + rh, x0:x0 + rw]] == 0).all()
    dev.lib.ibl_accel_free(dev.ctx, ibl.accel)


def test_shade_default_path_on_cube_face_ties(pbr, dev, oracle):
    """fp16 normals do hit exact ties of the major axis (|x| == |y| ...).  The texture unit breaks them its own way and, without
    seamless filtering, the two faces return different edge texels; the default path (irradiance through the unit) must still
    pick the face the oracle picks (X, then Y, then Z)."""
    host, bufs = build_ibl(dev, oracle)
    w, h = 64, 48
    g = synth_gbuffer(w, h, seed=77, tier_b=True)
    ties = np.array([[1, 1, 0.3], [1, -1, 0.2], [-1, 1, -0.4], [0.5, 0.2, 0.5], [-0.5, 0.1, 0.5], [0.3, 0.7, 0.7], [0.2, -0.7, 0.7],
                     [1, 1, 1], [-1, -1, -1], [1, 0, 1], [0, 1, 1], [1, 1, 0]], np.float32)
    nrm = np.asarray(g["nrm"], np.float32).copy()
    lit = ~g["bg"]
    ys, xs = np.nonzero(lit)
    for k, (y, x) in enumerate(zip(ys, xs)):
        nrm[y, x, :3] = ties[k % len(ties)] * (0.25 + 0.125 * (k % 5))  # un-normalised on purpose: the kernels normalise
    g["nrm"] = nrm.astype(np.float16)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    lights = synth_lights(4, seed=3)
    ld = dev.upload(lights)
    ibl = make_ibl_struct(pbr, host, bufs, dev, accel=True)
    cam = (0.5, -1.0, 12.0)
    out = dev.alloc(w * h * 8)
    dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ld.ptr, 4, None, C.byref(ibl), pbr.WHOLE, out.ptr, w * 8, None)
    dev.sync()
    got = out.download((h, w, 4), np.uint16)
    _, want = oracle.shade(g["pos"], g["nrm"], g["alb"], g["orm"], g["emi"], cam, lights, host)
    assert_hdr_parity(got, want, peak=float(max(f16(want).max(), 1.0)), rel_tol=REL, what="default path on face ties")
    dev.lib.ibl_accel_free(dev.ctx, ibl.accel)


def test_shade_texture_unit_on_sharp_environment(pbr, dev, oracle):
    """K3's texture-unit path with a hard-edged 5e3 sun in the environment (the worst case for the unit's 8-bit
    weights): the contrast flags must route the affected pixels to the fp32-weight records."""
    res, nmips, ns, ires, lres = 128, 8, 256, 32, 128
    env = synth_environment(512, 256, 3, sun=5.0e3)
    src = dev.upload(env)
    n = oracle.chain_texels(res, nmips)
    chain_d = dev.alloc(n * 8)
    dev.lib.ibl_equirect_to_cube(dev.ctx, src.ptr, 512, 256, res, chain_d.ptr, None)
    dev.lib.ibl_cube_mipgen(dev.ctx, chain_d.ptr, res, nmips, None)
    pref_d = dev.alloc(n * 8)
    dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, pref_d.ptr, 0, nmips - 1, 0, 1, None)
    irr_d = dev.alloc(6 * ires * ires * 8)
    dev.lib.ibl_irradiance(dev.ctx, chain_d.ptr, res, nmips, ires, ns, irr_d.ptr, None)
    lut_d = dev.alloc(lres * lres * 8)
    dev.lib.ibl_brdf_lut(dev.ctx, lres, ns, lut_d.ptr, None)
    dev.sync()
    host = dict(irr=irr_d.download((6, ires, ires, 4), np.uint16), pref=pref_d.download((n, 4), np.uint16),
                pref_res=res, pref_mips=nmips, lut=lut_d.download((lres, lres, 2), np.float32))
    w, h = 320, 180
    g = synth_gbuffer(w, h, seed=11, tier_b=True)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    ibl = make_ibl_struct(pbr, host, (irr_d, pref_d, lut_d), dev, accel=True)
    cam = (0.5, -1.0, 12.0)
    out = dev.alloc(w * h * 8)
    dev.lib.shade_set_texture_unit(dev.ctx, 1)
    dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), None, 0, None, C.byref(ibl), pbr.WHOLE, out.ptr, w * 8, None)
    dev.sync()
    dev.lib.shade_set_texture_unit(dev.ctx, 0)
    got = out.download((h, w, 4), np.uint16)
    _, want = oracle.shade(g["pos"], g["nrm"], g["alb"], g["orm"], g["emi"], cam, synth_lights(0, seed=1), host)
    assert f16(want).max() > 50.0, "the sun should be visible in some reflections"
    assert_hdr_parity(got, want, peak=float(f16(want).max()), rel_tol=REL, what="shade (texture unit), sharp environment")
    dev.lib.ibl_accel_free(dev.ctx, ibl.accel)


def test_shade_tiled_lights_equal_brute_force(pbr, dev, oracle):
    w, h, nl = 200, 120, 256
    host, bufs = build_ibl(dev, oracle)
    g = synth_gbuffer(w, h, seed=77, tier_b=True)
    # spatially coherent positions so tiles have tight bounds
    yy, xx = np.mgrid[0:h, 0:w]
    g["pos"][..., 0] = (xx / w * 16 - 8); g["pos"][..., 1] = (yy / h * 16 - 8); g["pos"][..., 2] = np.sin(xx * 0.1) * 2
    g["pos"][g["bg"]] = 0
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    lights = synth_lights(nl, seed=3, ranged=True)
    ld = dev.upload(lights)
    tiles = pbr.LightTiles()
    dev.lib.light_tiles_alloc(dev.ctx, w, h, 256, C.byref(tiles))
    dev.lib.cluster_lights(dev.ctx, C.byref(gb), ld.ptr, nl, C.byref(tiles), None)
    ibl = make_ibl_struct(pbr, host, bufs)
    cam = (0.0, 0.0, 14.0)
    o1 = dev.alloc(w * h * 8); o2 = dev.alloc(w * h * 8)
    dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ld.ptr, nl, C.byref(tiles), C.byref(ibl), pbr.WHOLE, o1.ptr, w * 8, None)
    dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ld.ptr, nl, None, C.byref(ibl), pbr.WHOLE, o2.ptr, w * 8, None)
    dev.sync()
    tiled, brute = o1.download((h, w, 4), np.uint16), o2.download((h, w, 4), np.uint16)
    # K3 sums the even and the odd entries of a light list in separate lanes, so culling a light changes the
    # association of the fp32 sum (not its terms): equal up to fp32 rounding, i.e. at most 1 fp16 ulp apart
    assert (np.abs(tiled.astype(np.int32) - brute.astype(np.int32)) <= 1).all(), "tiled light lists changed the result"
    assert np.mean(tiled != brute) < 0.02
    cnt = pbr.DeviceBuffer.__new__(pbr.DeviceBuffer); cnt.dev, cnt.nbytes, cnt.ptr = dev, tiles.tiles_x * tiles.tiles_y * 4, tiles.counts
    counts = cnt.download((tiles.tiles_y, tiles.tiles_x), np.uint32)
    assert counts.mean() < 0.6 * nl, "clustering culled nothing"
    _, want = oracle.shade(g["pos"], g["nrm"], g["alb"], g["orm"], g["emi"], cam, lights, host)
    assert_hdr_parity(tiled, want, peak=float(max(f16(want).max(), 1.0)), rel_tol=REL, what="tiled shade")
    dev.lib.light_tiles_free(dev.ctx, C.byref(tiles))


@pytest.mark.parametrize("exact", [0, 1])
def test_prefilter_shards_are_bit_identical_to_the_whole(dev, oracle, exact):
    """Multi-GPU partition of K7 (row blocks of every level): at a size where the levels use DIFFERENT numbers of lanes
    per texel, 2 and 8 shards must reproduce the unsharded chain bit for bit (the lane count is taken from the whole
    level, so a texel's samples are summed in the same order whoever computes it)."""
    res, nmips, ns = 256, 9, 256
    _, chain_d, chain = make_env_chain(dev, oracle, res, nmips, seed=3)
    dev.lib.ibl_set_exact_filtering(dev.ctx, exact)
    whole = dev.alloc(chain.nbytes)
    dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, whole.ptr, 0, nmips - 1, 0, 1, None)
    dev.sync()
    want = whole.download(chain.shape, np.uint16)
    for nshards in (2, 8):
        part = dev.alloc(chain.nbytes); part.zero()
        for s in range(nshards):
            dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, part.ptr, 0, nmips - 1, s, nshards, None)
        dev.sync()
        got = part.download(chain.shape, np.uint16)
        assert np.array_equal(got, want), f"{nshards} shards: {np.count_nonzero(got != want)} values differ"
    dev.lib.ibl_set_exact_filtering(dev.ctx, 0)


def test_shade_batched_views_equal_single_launches(pbr, dev, oracle):
    """pbrcu_shade_ggx_ibl_batched (config 5: many views of one shape in ONE launch) writes exactly the frames the
    per-view calls write -- also when the grid has more CTAs than one view has work items, and for a single view."""
    host, bufs = build_ibl(dev, oracle)
    ibl = make_ibl_struct(pbr, host, bufs, dev, accel=True)
    lights = synth_lights(5, seed=9)  # odd count: the padded null light
    ld = dev.upload(lights)
    for (w, h, nv) in ((64, 32, 7), (333, 61, 3), (48, 16, 1)):
        gs = [synth_gbuffer(w, h, seed=300 + v, tier_b=True) for v in range(nv)]
        stack = lambda k, dt: np.ascontiguousarray(np.stack([g[k] for g in gs])).view(dt)
        planes = dict(positions=dev.upload(stack("pos", np.uint16)), normals=dev.upload(stack("nrm", np.uint16)), albedo=dev.upload(stack("alb", np.uint16)),
                      orm=dev.upload(stack("orm", np.uint8)), emissive=dev.upload(stack("emi", np.uint16)))
        gb = pbr.GBuffer()
        gb.width, gb.height, gb.pitch8, gb.pitch4 = w, h, w * 8, w * 4
        for k, b in planes.items():
            setattr(gb, k, b.ptr)
        cams = np.array([[12 * np.cos(v), -1.0 - v, 12 * np.sin(v)] for v in range(nv)], np.float32)
        out = dev.alloc(nv * w * h * 8); out.zero()
        dev.lib.shade_ggx_ibl_batched(dev.ctx, C.byref(gb), w * h * 8, w * h * 4, nv, cams.ctypes.data_as(C.POINTER(C.c_float)), ld.ptr, 5, C.byref(ibl), out.ptr, w * 8,
                                      w * h * 8, None)
        dev.sync()
        got = out.download((nv, h, w, 4), np.uint16)
        one = dev.alloc(w * h * 8)
        for v in range(nv):
            gv = pbr.GBuffer()
            gv.width, gv.height, gv.pitch8, gv.pitch4 = w, h, w * 8, w * 4
            for k, b in planes.items():
                setattr(gv, k, b.ptr + v * w * h * (4 if k == "orm" else 8))
            dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gv), pbr.f3(*cams[v]), ld.ptr, 5, None, C.byref(ibl), pbr.WHOLE, one.ptr, w * 8, None)
            dev.sync()
            assert np.array_equal(got[v], one.download((h, w, 4), np.uint16)), f"{w}x{h} view {v} of {nv}"
        _, want = oracle.shade(gs[-1]["pos"], gs[-1]["nrm"], gs[-1]["alb"], gs[-1]["orm"], gs[-1]["emi"], cams[-1], lights, host)
        assert_hdr_parity(got[-1], want, peak=float(max(f16(want).max(), 1.0)), rel_tol=REL, what="batched shade, last view")
    with pytest.raises(pbr.PbrcuError):  # a view stride smaller than one view
        dev.lib.shade_ggx_ibl_batched(dev.ctx, C.byref(gb), 8, 4, 2, cams.ctypes.data_as(C.POINTER(C.c_float)), ld.ptr, 5, C.byref(ibl), out.ptr, w * 8, w * h * 8, None)
    dev.lib.ibl_accel_free(dev.ctx, ibl.accel)


def test_cluster_lights_on_a_row_band_equals_the_whole_frame(pbr, dev, oracle):
    """Multi-GPU config 4: a rank clusters ONLY its 16-aligned row band, through sub-views of the G-buffer and of the
    tile lists (no extra entry point) -- same counts and indices as the whole-frame clustering has for those tiles."""
    w, h, nl = 200, 120, 256
    g = synth_gbuffer(w, h, seed=78, tier_b=True)
    yy, xx = np.mgrid[0:h, 0:w]
    g["pos"][..., 0] = (xx / w * 16 - 8); g["pos"][..., 1] = (yy / h * 16 - 8); g["pos"][..., 2] = np.cos(xx * 0.07) * 2
    g["pos"][g["bg"]] = 0
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    ld = dev.upload(synth_lights(nl, seed=5, ranged=True))
    whole, part = pbr.LightTiles(), pbr.LightTiles()
    dev.lib.light_tiles_alloc(dev.ctx, w, h, 64, C.byref(whole))
    dev.lib.light_tiles_alloc(dev.ctx, w, h, 64, C.byref(part))
    dev.lib.cluster_lights(dev.ctx, C.byref(gb), ld.ptr, nl, C.byref(whole), None)

    def fetch(t):
        c = pbr.DeviceBuffer.__new__(pbr.DeviceBuffer); c.dev, c.nbytes, c.ptr = dev, t.tiles_x * t.tiles_y * 4, t.counts
        i = pbr.DeviceBuffer.__new__(pbr.DeviceBuffer); i.dev, i.nbytes, i.ptr = dev, t.tiles_x * t.tiles_y * t.max_per_tile * 2, t.indices
        return c.download((t.tiles_y, t.tiles_x), np.uint32), i.download((t.tiles_y, t.tiles_x, t.max_per_tile), np.uint16)

    for rank in range(3):
        y0, y1 = pbr.partition.row_band(h, 3, rank, align=16)
        band = pbr.partition.band_views(pbr, gb, part, y0, y1)
        dev.lib.cluster_lights(dev.ctx, C.byref(band.gbuffer), ld.ptr, nl, C.byref(band.tiles), None)
    dev.sync()
    (cw, iw), (cp, ip) = fetch(whole), fetch(part)
    assert np.array_equal(cw, cp) and cw.max() > 0
    for ty in range(whole.tiles_y):
        for tx in range(whole.tiles_x):
            assert np.array_equal(iw[ty, tx, :cw[ty, tx]], ip[ty, tx, :cw[ty, tx]])
    dev.lib.light_tiles_free(dev.ctx, C.byref(whole)); dev.lib.light_tiles_free(dev.ctx, C.byref(part))


def _plane(pbr, dev, ptr, nbytes):
    b = pbr.DeviceBuffer.__new__(pbr.DeviceBuffer)
    b.dev, b.nbytes, b.ptr = dev, nbytes, ptr
    return b


def test_shade_full_size_properties(pbr, dev, oracle):
    """BASELINE shapes for K3 through size-independent properties, on device-generated gbuf_spheres:
    3840x2160 with 4 lights + IBL: cleared texels -> exactly 0, finite everywhere, the 8 row bands of the multi-GPU
    partition == the whole frame bit for bit, 4096 random texels vs the oracle run on the very bits the device generated;
    7680x4320 with 256 ranged lights: clustered lists == brute force (<= 1 fp16 ulp), row bands == whole frame bit for bit,
    and the same random-sample check."""
    host, bufs = build_ibl(dev, oracle)
    ibl = make_ibl_struct(pbr, host, bufs, dev, accel=True)
    cam = (0.0, -1.0, -12.0)
    rng = np.random.default_rng(11)
    for (w, h, nl, ranged) in [(3840, 2160, 4, False), (7680, 4320, 256, True)]:
        gb = pbr.GBuffer()
        dev.lib.gbuffer_alloc(dev.ctx, w, h, 1, C.byref(gb))
        dev.lib.synth_gbuffer(dev.ctx, C.byref(gb), 1, 0x5EED0002, pbr.f3(*cam), pbr.f3(0, 0, 0), None)
        lights = synth_lights(nl, seed=2, ranged=ranged)
        ld = dev.upload(lights)
        pitch = gb.pitch8
        out = dev.alloc(pitch * h); out2 = dev.alloc(pitch * h)
        tiles = None
        if ranged:
            tiles = pbr.LightTiles()
            dev.lib.light_tiles_alloc(dev.ctx, w, h, 256, C.byref(tiles))
            dev.lib.cluster_lights(dev.ctx, C.byref(gb), ld.ptr, nl, C.byref(tiles), None)
        tp = C.byref(tiles) if tiles else None
        dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ld.ptr, nl, tp, C.byref(ibl), pbr.WHOLE, out.ptr, pitch, None)
        bands = 8
        for r in range(bands):  # 16-row aligned bands, like partition.row_band
            y0, y1 = (r * h // bands) // 16 * 16, ((r + 1) * h // bands) // 16 * 16 if r + 1 < bands else h
            dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ld.ptr, nl, tp, C.byref(ibl), pbr.Rect(0, y0, w, y1 - y0), out2.ptr, pitch, None)
        dev.sync()
        a = out.download((h, w, 4), np.uint16); b = out2.download((h, w, 4), np.uint16)
        assert np.array_equal(a, b), f"{w}x{h}: row-band result differs from the whole-frame result"
        if ranged:  # brute force over all 256 lights
            dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ld.ptr, nl, None, C.byref(ibl), pbr.WHOLE, out2.ptr, pitch, None)
            dev.sync()
            b = out2.download((h, w, 4), np.uint16)
            assert (np.abs(a.astype(np.int32) - b.astype(np.int32)) <= 1).all(), "clustered light lists changed the result"
        planes = {k: _plane(pbr, dev, getattr(gb, k), (pitch if k != "orm" else gb.pitch4) * h) for k in ("positions", "normals", "albedo", "orm", "emissive")}
        N = planes["normals"].download((h, w, 4), np.uint16)
        bg = (N[..., :3] == 0).all(-1)
        assert 0.02 < bg.mean() < 0.9
        assert (a[bg] == 0).all()
        assert np.isfinite(f16(a)).all()
        ys, xs = rng.integers(0, h, 4096), rng.integers(0, w, 4096)
        P = planes["positions"].download((h, w, 4), np.uint16)[ys, xs][None]
        A = planes["albedo"].download((h, w, 4), np.uint16)[ys, xs][None]
        E = planes["emissive"].download((h, w, 4), np.uint16)[ys, xs][None]
        O = planes["orm"].download((h, w, 4), np.uint8)[ys, xs][None]
        _, want = oracle.shade(P.view(np.float16), N[ys, xs][None].view(np.float16), A, O, E.view(np.float16), cam, lights, host)
        assert_hdr_parity(a[ys, xs][None], want, peak=float(max(f16(want).max(), 1.0)), rel_tol=REL, what=f"{w}x{h} sample")
        if tiles:
            dev.lib.light_tiles_free(dev.ctx, C.byref(tiles))
        out.free(); out2.free(); ld.free()
        dev.lib.gbuffer_free(dev.ctx, C.byref(gb))
    dev.lib.ibl_accel_free(dev.ctx, ibl.accel)


def test_prefilter_full_size_properties(dev, oracle):
    """BASELINE configs[2] (2048^2 cube, 12 mips, 1024 samples per texel) through size-independent properties:
    level 0 is an exact copy; 2 and 8 row-block shards reproduce the unsharded chain bit for bit; the texture-unit result is
    within tolerance of the fp32-weight result on every texel of every level; a few texels per level vs the oracle."""
    res, nmips, ns = 2048, 12, 1024
    n = oracle.chain_texels(res, nmips)
    env = synth_environment(512, 256, 7)
    src = dev.upload(env)
    chain_d = dev.alloc(n * 8)
    dev.lib.ibl_equirect_to_cube(dev.ctx, src.ptr, 512, 256, res, chain_d.ptr, None)
    dev.lib.ibl_cube_mipgen(dev.ctx, chain_d.ptr, res, nmips, None)
    whole = dev.alloc(n * 8); part = dev.alloc(n * 8); exact_d = dev.alloc(n * 8)
    dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, whole.ptr, 0, nmips - 1, 0, 1, None)
    dev.sync()
    chain = chain_d.download((n, 4), np.uint16)
    want = whole.download((n, 4), np.uint16)
    assert np.array_equal(want[: 6 * res * res], chain[: 6 * res * res]), "level 0 (roughness 0) must be a copy"
    for nshards in (2, 8):
        part.zero()
        for s in range(nshards):
            dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, part.ptr, 0, nmips - 1, s, nshards, None)
        dev.sync()
        got = part.download((n, 4), np.uint16)
        assert np.array_equal(got, want), f"{nshards} shards: {np.count_nonzero(got != want)} values differ"
    dev.lib.ibl_set_exact_filtering(dev.ctx, 1)
    dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, exact_d.ptr, 0, nmips - 1, 0, 1, None)
    dev.sync()
    dev.lib.ibl_set_exact_filtering(dev.ctx, 0)
    ex = exact_d.download((n, 4), np.uint16)
    a = oracle.mip_offset(res, 1)
    fe, fw = f16(ex[a:])[:, :3], f16(want[a:])[:, :3]
    assert (np.abs(fw - fe) <= 1e-3 + REL * np.abs(fe)).all(), "texture-unit prefilter outside the tolerance of the fp32-weight one"
    rng = np.random.default_rng(5)
    for m in range(1, nmips):
        r = res >> m
        off = oracle.mip_offset(res, m)
        for _ in range(6):
            f, x, y = int(rng.integers(6)), int(rng.integers(r)), int(rng.integers(r))
            ref = oracle.prefilter_texel(chain, res, nmips, m, ns, f, x, y)[:3]
            got = f16(want[off + (f * r + y) * r + x])[:3]
            assert (np.abs(got - ref) <= 1e-3 + REL * np.abs(ref)).all(), (m, f, x, y, got, ref)
    for b in (src, chain_d, whole, part, exact_d):
        b.free()


@pytest.mark.parametrize("mode", ["one-list", "tile-lists", "no-ibl"])
def test_shade_replicated_stores_equal_the_plain_frame(pbr, dev, oracle, mode):
    """pbrcu_shade_ggx_ibl_replicated (the row-band gather fused into K3's stores): every replica receives exactly the
    texels of the rect, bit-identical to the plain entry point, and nothing outside the rect -- here with the replicas in
    local memory; the N = 2..8 bench runs the same call on peer-mapped frames and checks the gathered frame."""
    w, h, nl = 150, 70, 7
    host, bufs = build_ibl(dev, oracle)
    g = synth_gbuffer(w, h, seed=5, tier_b=True)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    lights = synth_lights(nl, seed=2)
    ld = dev.upload(lights)
    ibl = make_ibl_struct(pbr, host, bufs, dev, accel=True)
    tiles = pbr.LightTiles()
    dev.lib.light_tiles_alloc(dev.ctx, w, h, 16, C.byref(tiles))
    dev.lib.cluster_lights(dev.ctx, C.byref(gb), ld.ptr, nl, C.byref(tiles), None)
    targ = C.byref(tiles) if mode == "tile-lists" else None
    iarg = None if mode == "no-ibl" else C.byref(ibl)
    cam = pbr.f3(0.5, -1.0, 12.0)
    plain = dev.alloc(w * h * 8)
    dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), cam, ld.ptr, nl, targ, iarg, pbr.WHOLE, plain.ptr, w * 8, None)
    dev.sync()
    want = plain.download((h, w, 4), np.uint16)
    sentinel = np.full((h, w, 4), 0x7B7B, np.uint16)
    rect = (3, 17, 140, 37)  # starts and ends inside tiles
    own = dev.upload(sentinel)
    reps = [dev.upload(sentinel) for _ in range(3)]
    ptrs = (C.c_void_p * 3)(*[r.ptr for r in reps])
    dev.lib.shade_ggx_ibl_replicated(dev.ctx, C.byref(gb), cam, ld.ptr, nl, targ, iarg, pbr.Rect(*rect), own.ptr, w * 8, ptrs, 3, None)
    dev.sync()
    expect = sentinel.copy()
    expect[rect[1]:rect[1] + rect[3], rect[0]:rect[0] + rect[2]] = want[rect[1]:rect[1] + rect[3], rect[0]:rect[0] + rect[2]]
    for what, buf in [("own frame", own)] + [(f"replica {k}", r) for k, r in enumerate(reps)]:
        got = buf.download((h, w, 4), np.uint16)
        assert np.array_equal(got, expect), f"{mode}: {what} differs in {np.count_nonzero(got != expect)} values"
    # a replica that is only 8-byte aligned: the kernel must fall back from 16-byte pair stores to 8-byte stores, same bits
    shifted = dev.upload(np.full((h * w + 1, 4), 0x7B7B, np.uint16))
    one = (C.c_void_p * 1)(shifted.ptr + 8)
    dev.lib.shade_ggx_ibl_replicated(dev.ctx, C.byref(gb), cam, ld.ptr, nl, targ, iarg, pbr.Rect(*rect), own.ptr, w * 8, one, 1, None)
    dev.sync()
    got = shifted.download((h * w + 1, 4), np.uint16)
    assert (got[0] == 0x7B7B).all() and np.array_equal(got[1:].reshape(h, w, 4), expect), f"{mode}: 8-byte aligned replica differs"
    # no replicas = the plain call
    dev.lib.shade_ggx_ibl_replicated(dev.ctx, C.byref(gb), cam, ld.ptr, nl, targ, iarg, pbr.WHOLE, own.ptr, w * 8, None, 0, None)
    dev.sync()
    assert np.array_equal(own.download((h, w, 4), np.uint16), want)
    with pytest.raises(pbr.PbrcuError):  # more replicas than PBRCU_MAX_REPLICAS
        many = (C.c_void_p * 16)(*[reps[0].ptr] * 16)
        dev.lib.shade_ggx_ibl_replicated(dev.ctx, C.byref(gb), cam, ld.ptr, nl, targ, iarg, pbr.WHOLE, own.ptr, w * 8, many, 16, None)
    with pytest.raises(pbr.PbrcuError):  # a NULL replica
        bad = (C.c_void_p * 2)(reps[0].ptr, None)
        dev.lib.shade_ggx_ibl_replicated(dev.ctx, C.byref(gb), cam, ld.ptr, nl, targ, iarg, pbr.WHOLE, own.ptr, w * 8, bad, 2, None)
    with pytest.raises(pbr.PbrcuError):  # a misaligned replica
        odd = (C.c_void_p * 1)(reps[0].ptr + 4)
        dev.lib.shade_ggx_ibl_replicated(dev.ctx, C.byref(gb), cam, ld.ptr, nl, targ, iarg, pbr.WHOLE, own.ptr, w * 8, odd, 1, None)
    dev.lib.light_tiles_free(dev.ctx, C.byref(tiles))
    dev.lib.ibl_accel_free(dev.ctx, ibl.accel)
    for b in (plain, own, ld, *reps):
        b.free()


@pytest.mark.parametrize("exact", [0, 1])
def test_prefilter_replicated_stores_build_the_whole_chain(dev, oracle, exact):
    """pbrcu_ibl_prefilter_ggx_replicated (the all-gather of a sharded prefilter fused into K7's stores): three "ranks"
    each run their shard into their own chain with the other two as replicas; afterwards all three chains are complete
    and bit-identical to the unsharded prefilter.  Replicas in local memory here; the N >= 2 bench maps peer memory."""
    res, nmips, ns = 64, 7, 128
    _, chain_d, chain = make_env_chain(dev, oracle, res, nmips, seed=4)
    dev.lib.ibl_set_exact_filtering(dev.ctx, exact)
    whole = dev.alloc(chain.nbytes)
    dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, whole.ptr, 0, nmips - 1, 0, 1, None)
    dev.sync()
    want = whole.download(chain.shape, np.uint16)
    world = 3
    chains = [dev.upload(np.full(chain.shape, 0x7B7B, np.uint16)) for _ in range(world)]
    for r in range(world):
        others = (C.c_void_p * (world - 1))(*[chains[k].ptr for k in range(world) if k != r])
        dev.lib.ibl_prefilter_ggx_replicated(dev.ctx, chain_d.ptr, res, nmips, ns, chains[r].ptr, 0, nmips - 1, r, world, others, world - 1, None)
    dev.sync()
    for r in range(world):
        got = chains[r].download(chain.shape, np.uint16)
        assert np.array_equal(got, want), f"chain of rank {r} differs from the unsharded prefilter in {np.count_nonzero(got != want)} values"
    with pytest.raises(Exception):
        dev.lib.ibl_prefilter_ggx_replicated(dev.ctx, chain_d.ptr, res, nmips, ns, chains[0].ptr, 0, nmips - 1, 0, 1, None, 2, None)
    dev.lib.ibl_set_exact_filtering(dev.ctx, 0)
    for b in (whole, *chains):
        b.free()


@pytest.mark.parametrize("w,h,bands", [(37, 1, 0), (64, 3, 8), (200, 4, 4), (130, 7, 4), (256, 33, 0), (160, 90, 8), (96, 70, 64), (320, 180, 1)])
def test_shade_frame_host_equals_the_device_path(pbr, dev, oracle, w, h, bands):
    """pbrcu_shade_frame_host (the e2e call of the headline metric: host planes in, HDR frame out, tapered row bands pipelined
    over two streams) must return the bits of pbrcu_shade_ggx_ibl on the same inputs -- at heights smaller than the band
    count, bands that come out empty, one band, and the 64-band maximum; and must leave the bytes after the frame alone."""
    host, bufs = build_ibl(dev, oracle)
    g = synth_gbuffer(w, h, seed=11 * w + h, tier_b=True)
    nl = 5
    lights = synth_lights(nl, seed=6)
    ld = dev.upload(lights)
    ibl = make_ibl_struct(pbr, host, bufs, dev, accel=True)
    cam = pbr.f3(0.5, -1.0, 12.0)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    out = dev.alloc(w * h * 8)
    # (the same HOST light list: both calls then take the parameter-block kernel)
    dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), cam, lights.ctypes.data, nl, None, C.byref(ibl), pbr.WHOLE, out.ptr, w * 8, None)
    dev.sync()
    want = out.download((h, w, 4), np.uint16)
    planes = {k: np.ascontiguousarray(g[k]) for k in ("pos", "nrm", "alb", "orm", "emi")}
    got = np.full((h + 2, w, 4), 0x7B7B, np.uint16)  # two guard rows after the frame
    for _ in range(2):  # twice: the staging buffers and events are reused
        dev.lib.shade_frame_host(dev.ctx, planes["pos"].ctypes.data, planes["nrm"].ctypes.data, planes["alb"].ctypes.data, planes["orm"].ctypes.data,
                                 planes["emi"].ctypes.data, w, h, cam, lights.ctypes.data, nl, C.byref(ibl), got.ctypes.data, bands, None)
        assert np.array_equal(got[:h], want), f"{np.count_nonzero(got[:h] != want)} values differ from the device path"
        assert (got[h:] == 0x7B7B).all(), "wrote past the frame"
    dev.lib.ibl_accel_free(dev.ctx, ibl.accel)
    for b in (out, ld, *bufs):
        b.free()


@pytest.mark.parametrize("nl", [0, 1, 3, 4, 7, 16])
def test_shade_host_light_list_kernels(pbr, dev, oracle, nl, monkeypatch):
    """Lights passed as a HOST array (what the facade, the bench and pbrcu_shade_frame_host do) travel in the kernel's parameter
    block (shade_ggx_cl.cu).  Default: IBL taps from the footprint records; pbrcu_ibl.flags & PBRCU_IBL_GATHER: tld4 gathers; PBRCU_K3_CL=0: the list
    is uploaded and the device-list kernel runs, bit for bit.  Every mode against the oracle in both filter modes; the exact
    modes against the device-list kernel within one fp16 ulp (the lights are summed in another contraction order)."""
    host, bufs = build_ibl(dev, oracle)
    w, h = 83, 41
    g = synth_gbuffer(w, h, seed=300 + nl, tier_b=True)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    lights = synth_lights(max(nl, 1), seed=30 + nl)[:nl]
    ld = dev.upload(lights) if nl else None
    ibl = make_ibl_struct(pbr, host, bufs, dev, accel=True, gather=True)
    cam = (0.5, -1.0, 12.0)
    out = dev.alloc(w * h * 8)

    def run(ptr, use_ibl=True, gather=False):
        ibl.flags = pbr.IBL_GATHER if gather else 0
        dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), ptr, nl, None, C.byref(ibl) if use_ibl else None, pbr.WHOLE, out.ptr, w * 8, None)
        dev.sync()
        return out.download((h, w, 4), np.uint16)

    host_ptr = lights.ctypes.data if nl else None
    _, want = oracle.shade(g["pos"], g["nrm"], g["alb"], g["orm"], g["emi"], cam, lights, host)
    peak = float(max(f16(want).max(), 1.0))
    for exact in (0, 1):
        dev.lib.ibl_set_exact_filtering(dev.ctx, exact)
        monkeypatch.delenv("PBRCU_K3_CL", raising=False)
        from_dev = run(ld.ptr if ld else None)
        monkeypatch.setenv("PBRCU_K3_CL", "0")
        assert np.array_equal(run(host_ptr), from_dev), "uploaded host light list != device light list"
        monkeypatch.delenv("PBRCU_K3_CL", raising=False)
        for mode in ("records", "gathers"):
            got = run(host_ptr, gather=mode == "gathers")
            assert_hdr_parity(got, want, peak=peak, rel_tol=REL, what=f"parameter-block kernel nl={nl} exact={exact} mode={mode}")
            assert (got[g["bg"]] == 0).all()
            if exact:
                ulps = np.abs(got.view(np.int16).astype(int) - from_dev.view(np.int16).astype(int))
                assert ulps.max() <= 1, f"mode {mode} (exact) differs from the device-list kernel by {ulps.max()} fp16 ulps"
    monkeypatch.delenv("PBRCU_K3_CL", raising=False)
    dev.lib.ibl_set_exact_filtering(dev.ctx, 0)
    # no IBL at all
    _, want0 = oracle.shade(g["pos"], g["nrm"], g["alb"], g["orm"], g["emi"], cam, lights, None)
    assert_hdr_parity(run(host_ptr, use_ibl=False), want0, peak=float(max(f16(want0).max(), 1.0)), rel_tol=REL, what=f"parameter-block kernel nl={nl}, no IBL")
    dev.lib.ibl_accel_free(dev.ctx, ibl.accel)


def test_prepared_source_equals_the_plain_entry_points(dev, oracle):
    """pbrcu_ibl_source: the flagged texture copy built once; prefilter (whole and as three shards) and irradiance from it must
    return the bits of the entry points that rebuild it per call -- also after the context's own texture was re-used for
    another chain in between (the object owns its copy)."""
    res, nmips, ns, ires = 64, 7, 128, 16
    env, chain_d, chain = make_env_chain(dev, oracle, res, nmips)
    want_pref, want_irr = dev.alloc(chain.nbytes), dev.alloc(6 * ires * ires * 8)
    dev.lib.ibl_prefilter_ggx(dev.ctx, chain_d.ptr, res, nmips, ns, want_pref.ptr, 0, nmips - 1, 0, 1, None)
    dev.lib.ibl_irradiance(dev.ctx, chain_d.ptr, res, nmips, ires, ns, want_irr.ptr, None)
    src = C.c_void_p()
    dev.lib.ibl_source_create(dev.ctx, chain_d.ptr, res, nmips, None, C.byref(src))
    # another chain through the plain entry point: overwrites the context's texture, must not touch the prepared object
    _, other_d, other = make_env_chain(dev, oracle, res, nmips, seed=5)
    scratch = dev.alloc(other.nbytes)
    dev.lib.ibl_prefilter_ggx(dev.ctx, other_d.ptr, res, nmips, ns, scratch.ptr, 0, nmips - 1, 0, 1, None)
    got_pref, got_irr, got_shards = dev.alloc(chain.nbytes), dev.alloc(6 * ires * ires * 8), dev.alloc(chain.nbytes)
    dev.lib.ibl_prefilter_ggx_prepared(dev.ctx, src, ns, got_pref.ptr, 0, nmips - 1, 0, 1, None, 0, None)
    for k in range(3):
        dev.lib.ibl_prefilter_ggx_prepared(dev.ctx, src, ns, got_shards.ptr, 0, nmips - 1, k, 3, None, 0, None)
    dev.lib.ibl_irradiance_prepared(dev.ctx, src, ires, ns, got_irr.ptr, None)
    dev.sync()
    w = want_pref.download(chain.shape, np.uint16)
    assert np.array_equal(got_pref.download(chain.shape, np.uint16), w)
    assert np.array_equal(got_shards.download(chain.shape, np.uint16), w)
    assert np.array_equal(got_irr.download((6, ires, ires, 4), np.uint16), want_irr.download((6, ires, ires, 4), np.uint16))
    dev.lib.ibl_source_free(dev.ctx, src)


def test_per_call_filter_flags_equal_the_context_switches(pbr, dev, oracle):
    """pbrcu_ibl.flags selects the ambient filtering per call: PBRCU_IBL_EXACT == pbrcu_ibl_set_exact_filtering(ctx, 1),
    PBRCU_IBL_TEXTURE_UNIT == pbrcu_shade_set_texture_unit(ctx, 1), bit for bit, on the device-list, host-list and tiled kernels."""
    host, bufs = build_ibl(dev, oracle)
    w, h, nl = 70, 50, 6
    g = synth_gbuffer(w, h, seed=91, tier_b=True)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    lights = synth_lights(nl, seed=3, ranged=True)
    ld = dev.upload(lights)
    ibl = make_ibl_struct(pbr, host, bufs, dev, accel=True)
    tiles = pbr.LightTiles()
    dev.lib.light_tiles_alloc(dev.ctx, w, h, nl, C.byref(tiles))
    dev.lib.cluster_lights(dev.ctx, C.byref(gb), ld.ptr, nl, C.byref(tiles), None)
    out = dev.alloc(w * h * 8)

    def run(ptr, use_tiles, flags):
        ibl.flags = flags
        dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(0.5, -1.0, 12.0), ptr, nl, C.byref(tiles) if use_tiles else None, C.byref(ibl), pbr.WHOLE, out.ptr, w * 8, None)
        dev.sync()
        ibl.flags = 0
        return out.download((h, w, 4), np.uint16)

    for ptr, use_tiles in ((ld.ptr, False), (lights.ctypes.data, False), (ld.ptr, True)):
        default = run(ptr, use_tiles, 0)
        dev.lib.ibl_set_exact_filtering(dev.ctx, 1)
        exact_ctx = run(ptr, use_tiles, 0)
        dev.lib.ibl_set_exact_filtering(dev.ctx, 0)
        assert np.array_equal(run(ptr, use_tiles, pbr.IBL_EXACT), exact_ctx)
        assert not np.array_equal(exact_ctx, default), "the exact mode should differ from the default somewhere on this frame"
        dev.lib.shade_set_texture_unit(dev.ctx, 1)
        tex_ctx = run(ptr, use_tiles, 0)
        dev.lib.shade_set_texture_unit(dev.ctx, 0)
        assert np.array_equal(run(ptr, use_tiles, pbr.IBL_TEXTURE_UNIT), tex_ctx)
        assert np.array_equal(run(ptr, use_tiles, 0), default), "flags must not stick"
    dev.lib.light_tiles_free(dev.ctx, C.byref(tiles))
    dev.lib.ibl_accel_free(dev.ctx, ibl.accel)


@pytest.mark.parametrize("mode", ["device-list", "host-list", "tile-lists", "no-ibl"])
def test_shade_ldr_equals_shade_then_tonemap(pbr, dev, oracle, mode):
    """pbrcu_shade_ggx_ibl_ldr (the reference's tonemap pass, aces+gamma.glsl:30-40, fused into K3's stores) == pbrcu_shade_ggx_ibl followed by
    pbrcu_tonemap_aces_gamma, bit for bit, on every K3 kernel; with two replicas and a rect that starts and ends inside tiles, texels outside
    the rect untouched; and within 1 LSB of the oracle's tonemap of the oracle's frame."""
    w, h, nl = 150, 70, 7
    host, bufs = build_ibl(dev, oracle)
    g = synth_gbuffer(w, h, seed=15, tier_b=True)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    lights = synth_lights(nl, seed=2)
    ld = dev.upload(lights)
    ibl = make_ibl_struct(pbr, host, bufs, dev, accel=True)
    tiles = pbr.LightTiles()
    dev.lib.light_tiles_alloc(dev.ctx, w, h, 16, C.byref(tiles))
    dev.lib.cluster_lights(dev.ctx, C.byref(gb), ld.ptr, nl, C.byref(tiles), None)
    lptr = lights.ctypes.data if mode == "host-list" else ld.ptr
    targ = C.byref(tiles) if mode == "tile-lists" else None
    iarg = None if mode == "no-ibl" else C.byref(ibl)
    cam = (0.5, -1.0, 12.0)
    hdr, ldr2 = dev.alloc(w * h * 8), dev.alloc(w * h * 4)
    dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(*cam), lptr, nl, targ, iarg, pbr.WHOLE, hdr.ptr, w * 8, None)
    dev.lib.tonemap_aces_gamma(dev.ctx, hdr.ptr, w * 8, w, h, pbr.WHOLE, ldr2.ptr, w * 4, None)
    dev.sync()
    want = ldr2.download((h, w, 4), np.uint8)
    got_whole = dev.alloc(w * h * 4)
    dev.lib.shade_ggx_ibl_ldr(dev.ctx, C.byref(gb), pbr.f3(*cam), lptr, nl, targ, iarg, pbr.WHOLE, got_whole.ptr, w * 4, None, 0, None)
    dev.sync()
    got = got_whole.download((h, w, 4), np.uint8)
    assert np.array_equal(got, want), f"{mode}: LDR frame differs from shade + tonemap in {np.count_nonzero(got != want)} values"
    _, o16 = oracle.shade(g["pos"], g["nrm"], g["alb"], g["orm"], g["emi"], cam, lights, None if mode == "no-ibl" else host)
    assert np.abs(got.astype(int) - oracle.tonemap(o16).astype(int)).max() <= 2  # 1 fp16 ulp of HDR + 1 LSB of rounding
    # band with replicas
    rect = (3, 17, 140, 37)
    sentinel = np.full((h, w, 4), 0x7B, np.uint8)
    own, reps = dev.upload(sentinel), [dev.upload(sentinel) for _ in range(2)]
    ptrs = (C.c_void_p * 2)(*[r.ptr for r in reps])
    dev.lib.shade_ggx_ibl_ldr(dev.ctx, C.byref(gb), pbr.f3(*cam), lptr, nl, targ, iarg, pbr.Rect(*rect), own.ptr, w * 4, ptrs, 2, None)
    dev.sync()
    expect = sentinel.copy()
    expect[rect[1]:rect[1] + rect[3], rect[0]:rect[0] + rect[2]] = want[rect[1]:rect[1] + rect[3], rect[0]:rect[0] + rect[2]]
    for what, buf in [("own frame", own)] + [(f"replica {k}", r) for k, r in enumerate(reps)]:
        assert np.array_equal(buf.download((h, w, 4), np.uint8), expect), f"{mode}: {what}"
    with pytest.raises(pbr.PbrcuError):  # a pitch smaller than a row of 4-byte texels
        dev.lib.shade_ggx_ibl_ldr(dev.ctx, C.byref(gb), pbr.f3(*cam), lptr, nl, targ, iarg, pbr.WHOLE, own.ptr, w * 4 - 4, None, 0, None)
    dev.lib.light_tiles_free(dev.ctx, C.byref(tiles))
    dev.lib.ibl_accel_free(dev.ctx, ibl.accel)


def test_shade_ldr_full_hd_equals_two_passes(pbr, dev, oracle):
    """The same identity on a full-HD device-generated frame (2 M texels: rare rounding cases that a small frame does not meet -- the fused
    store tonemaps one texel at a time, pbrcu_tonemap_aces_gamma two at a time in packed arithmetic; both must give the same bits),
    host list and tile lists."""
    w, h, nl = 1920, 1080, 24
    host, bufs = build_ibl(dev, oracle)
    gb = pbr.GBuffer()
    dev.lib.gbuffer_alloc(dev.ctx, w, h, 1, C.byref(gb))
    dev.lib.synth_gbuffer(dev.ctx, C.byref(gb), 1, 0x5EED0099, pbr.f3(0.0, -1.0, -12.0), pbr.f3(0, 0, 0), None)
    ibl = make_ibl_struct(pbr, host, bufs, dev, accel=True)
    lights = synth_lights(nl, seed=4, ranged=True)
    ld = dev.upload(lights)
    tiles = pbr.LightTiles()
    dev.lib.light_tiles_alloc(dev.ctx, w, h, nl, C.byref(tiles))
    dev.lib.cluster_lights(dev.ctx, C.byref(gb), ld.ptr, nl, C.byref(tiles), None)
    hdr, two, one = dev.alloc(gb.pitch8 * h), dev.alloc(w * h * 4), dev.alloc(w * h * 4)
    for lptr, n, targ in ((lights.ctypes.data, 4, None), (ld.ptr, nl, C.byref(tiles))):
        dev.lib.shade_ggx_ibl(dev.ctx, C.byref(gb), pbr.f3(0.0, -1.0, -12.0), lptr, n, targ, C.byref(ibl), pbr.WHOLE, hdr.ptr, gb.pitch8, None)
        dev.lib.tonemap_aces_gamma(dev.ctx, hdr.ptr, gb.pitch8, w, h, pbr.WHOLE, two.ptr, w * 4, None)
        dev.lib.shade_ggx_ibl_ldr(dev.ctx, C.byref(gb), pbr.f3(0.0, -1.0, -12.0), lptr, n, targ, C.byref(ibl), pbr.WHOLE, one.ptr, w * 4, None, 0, None)
        dev.sync()
        a, b = two.download((h, w, 4), np.uint8), one.download((h, w, 4), np.uint8)
        assert a.any() and np.array_equal(a, b), f"{np.count_nonzero(a != b)} bytes differ (max {np.abs(a.astype(int) - b.astype(int)).max()})"
    dev.lib.light_tiles_free(dev.ctx, C.byref(tiles))
    dev.lib.ibl_accel_free(dev.ctx, ibl.accel)
    dev.lib.gbuffer_free(dev.ctx, C.byref(gb))
