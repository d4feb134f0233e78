"""GPU parity, Tier A: the CUDA lighting / tonemap kernels (through the C ABI) vs the CPU oracle that
restates -- and is pinned bit-for-bit to -- the reference's GLSL (lighting.glsl, BlinnPhong.lib.glsl,
aces+gamma.glsl, Gamma.lib.glsl).  Tolerance (north_star): max abs err <= 1e-3 in linear HDR on the
RGBA16F output, PSNR >= 60 dB (peak 2.0 = the shader's range), LDR within +-1 LSB."""
import ctypes as C

import numpy as np
import pytest
from conftest import assert_hdr_parity, f16
from oracle import synth_gbuffer

pytestmark = pytest.mark.gpu

SIZES = [(1, 1), (2, 1), (3, 5), (17, 9), (64, 64), (257, 63), (1024, 1024)]
CAMS = [(0.0, 0.0, 0.0), (1.5, -2.0, 12.0)]


def run_lighting(pbr, dev, g, gb, cam, rect=None, pitch=None):
    w, h = gb.width, gb.height
    pitch = pitch or w * 8
    out = dev.alloc(pitch * h)
    out.zero()
    r = pbr.Rect(*rect) if rect else pbr.WHOLE
    dev.lib.lighting_blinnphong(dev.ctx, C.byref(gb), pbr.f3(*cam), r, out.ptr, pitch, None)
    dev.sync()
    res = out.download((h, pitch // 8, 4), np.uint16)[:, :w]
    out.free()
    return res


@pytest.mark.parametrize("w,h", SIZES)
@pytest.mark.parametrize("cam", CAMS)
def test_lighting_matches_oracle(pbr, dev, oracle, w, h, cam):
    g = synth_gbuffer(w, h, seed=w * 1000 + h)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"])
    got = run_lighting(pbr, dev, g, gb, cam)
    _, want = oracle.lighting(g["pos"], g["nrm"], g["alb"], cam)
    assert_hdr_parity(got, want, peak=2.0, what=f"lighting {w}x{h}")
    # cleared texels shade to exactly (0,0,0,0) -- also for a camera at the origin (normalize(0) -> NaN -> max -> 0)
    assert (got[g["bg"]] == 0).all()


def test_lighting_known_answers(pbr, dev):
    # N.V = 1  ->  out = albedo + 1 in all four channels (SURVEY 8(a) hand-derived vector)
    pos = np.zeros((1, 2, 4), np.float16)
    nrm = np.zeros((1, 2, 4), np.float16)
    nrm[0, 0, :3] = (0, 0, 1)
    alb = np.zeros((1, 2, 4), np.uint16)
    alb[0, 0] = (65535, 32768, 0, 16384)
    gb = dev.gbuffer(pos, nrm, alb)
    got = f16(run_lighting(pbr, dev, None, gb, (0, 0, 5)))
    want = alb[0, 0].astype(np.float32) / 65535.0 + 1.0
    assert np.abs(got[0, 0] - want).max() <= 1e-3
    assert (got[0, 1] == 0).all()  # all-zero texel -> exactly zero


def test_lighting_rect_and_pitch(pbr, dev, oracle):
    w, h = 100, 37
    g = synth_gbuffer(w, h, seed=5)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"])
    cam = (0.3, 0.2, 9.0)
    for rect in [(0, 0, 100, 37), (1, 3, 7, 5), (2, 0, 98, 37), (50, 36, 50, 1), (99, 0, 1, 37)]:
        for pitch in (w * 8, w * 8 + 8, w * 8 + 128):
            got = run_lighting(pbr, dev, g, gb, cam, rect=rect, pitch=pitch)
            _, want = oracle.lighting(g["pos"], g["nrm"], g["alb"], cam, rect=rect)
            x0, y0, rw, rh = rect
            assert_hdr_parity(got[y0:y0 + rh, x0:x0 + rw], want[y0:y0 + rh, x0:x0 + rw], peak=2.0, what=f"rect {rect} pitch {pitch}")
            mask = np.ones((h, w), bool)
            mask[y0:y0 + rh, x0:x0 + rw] = False
            assert (got[mask] == 0).all(), "texels outside the rect were written"


def run_tonemap(pbr, dev, hdr_bits, rect=None):
    h, w = hdr_bits.shape[:2]
    src = dev.upload(hdr_bits)
    out = dev.alloc(w * h * 4)
    out.zero()
    dev.lib.tonemap_aces_gamma(dev.ctx, src.ptr, w * 8, w, h, pbr.Rect(*rect) if rect else pbr.WHOLE, out.ptr, w * 4, None)
    dev.sync()
    res = out.download((h, w, 4), np.uint8)
    src.free(); out.free()
    return res


@pytest.mark.parametrize("w,h", SIZES)
def test_tonemap_matches_oracle(pbr, dev, oracle, w, h):
    rng = np.random.default_rng(w + 7 * h)
    hdr = (rng.random((h, w, 4)) ** 3 * 3.0).astype(np.float16)
    hdr[rng.random((h, w)) < 0.1] = 0  # black pixels: the reference's pow(negative) case, pinned to 0
    got = run_tonemap(pbr, dev, hdr.view(np.uint16))
    want = oracle.tonemap(hdr.view(np.uint16))
    d = np.abs(got.astype(int) - want.astype(int))
    assert d.max() <= 1, f"LDR differs by {d.max()} LSB"
    assert (d > 0).mean() < 0.02, "more than 2 % of LDR channels differ by 1 LSB"
    assert (got[(hdr[..., :3] == 0).all(-1)][:, :3] == 0).all()


def test_tonemap_known_answers(pbr, dev):
    # grey 0.18 -> (0.319164, 0.420219, 0.337710); (2,2,2) -> (1, 0.740818, 0.894574); 0 -> 0
    hdr = np.zeros((1, 4, 4), np.float16)
    hdr[0, 0] = (0.18, 0.18, 0.18, 1.0)
    hdr[0, 1] = (2, 2, 2, 0.5)
    hdr[0, 2] = (1, 1, 1, 2.0)
    got = run_tonemap(pbr, dev, hdr.view(np.uint16))
    # the fp16 input 0.18 is 0.18005; expected values recomputed for it stay within 1 LSB of the table
    exp = np.array([[0.319164, 0.420219, 0.337710, 1.0], [1.0, 0.740818, 0.894574, 0.5],
                    [0.862050, 0.723248, 0.785676, 1.0], [0, 0, 0, 0]])
    assert np.abs(got[0].astype(float) - np.rint(exp * 255)).max() <= 1


@pytest.mark.parametrize("with_hdr", [True, False])
def test_fused_equals_two_passes(pbr, dev, oracle, with_hdr):
    w, h = 333, 77
    g = synth_gbuffer(w, h, seed=11)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"])
    cam = (0.0, 1.0, 11.0)
    hdr = dev.alloc(w * h * 8)
    ldr = dev.alloc(w * h * 4)
    dev.lib.lighting_tonemap_fused(dev.ctx, C.byref(gb), pbr.f3(*cam), pbr.WHOLE, hdr.ptr if with_hdr else None, w * 8, ldr.ptr, w * 4, None)
    dev.sync()
    got_ldr = ldr.download((h, w, 4), np.uint8)
    two_pass_hdr = run_lighting(pbr, dev, g, gb, cam)
    two_pass_ldr = run_tonemap(pbr, dev, two_pass_hdr)
    assert np.array_equal(got_ldr, two_pass_ldr), "fused LDR differs from lighting->tonemap on the GPU"
    if with_hdr:
        assert np.array_equal(hdr.download((h, w, 4), np.uint16), two_pass_hdr)
    _, want_hdr = oracle.lighting(g["pos"], g["nrm"], g["alb"], cam)
    want_ldr = oracle.tonemap(want_hdr)
    assert np.abs(got_ldr.astype(int) - want_ldr.astype(int)).max() <= 2  # 1 fp16 ulp of HDR + 1 LSB rounding


def test_batched_views(pbr, dev, oracle):
    w, h, nv = 64, 32, 5
    gs = [synth_gbuffer(w, h, seed=100 + v) for v in range(nv)]
    pos = np.stack([g["pos"] for g in gs]); nrm = np.stack([g["nrm"] for g in gs]); alb = np.stack([g["alb"] for g in gs])
    gb = pbr.GBuffer()
    gb.width, gb.height, gb.pitch8, gb.pitch4 = w, h, w * 8, w * 4
    gb.positions = dev.upload(pos.view(np.uint16)).ptr
    gb.normals = dev.upload(nrm.view(np.uint16)).ptr
    gb.albedo = dev.upload(alb).ptr
    cams = np.array([[0, 0, 12], [12, 0, 0], [0, 3, -12], [-12, -3, 0], [0, 0, 0]], np.float32)
    out = dev.alloc(nv * w * h * 8)
    dev.lib.lighting_blinnphong_batched(dev.ctx, C.byref(gb), w * h * 8, w * h * 4, nv, cams.ctypes.data_as(C.POINTER(C.c_float)), out.ptr, w * 8, w * h * 8, None)
    dev.sync()
    got = out.download((nv, h, w, 4), np.uint16)
    for v in range(nv):
        _, want = oracle.lighting(gs[v]["pos"], gs[v]["nrm"], gs[v]["alb"], cams[v])
        assert_hdr_parity(got[v], want, peak=2.0, what=f"view {v}")


def test_batched_fused_views_equal_two_passes(pbr, dev, oracle):
    """pbrcu_lighting_tonemap_fused_batched (the reference's per-view lighting -> tonemap, App.cpp:387-392, over a block of views in
    one launch): every view's HDR equals the batched lighting pass bit for bit, its LDR the tonemap pass of that HDR bit for bit
    (so it inherits their pin to the reference's shaders); LDR-only gives the same LDR; odd width exercises the scalar path."""
    for w, h in ((64, 32), (37, 11)):
        nv = 4
        gs = [synth_gbuffer(w, h, seed=200 + v) for v in range(nv)]
        pos = np.stack([g["pos"] for g in gs]); nrm = np.stack([g["nrm"] for g in gs]); alb = np.stack([g["alb"] for g in gs])
        gb = pbr.GBuffer()
        gb.width, gb.height, gb.pitch8, gb.pitch4 = w, h, w * 8, w * 4
        gb.positions, gb.normals, gb.albedo = dev.upload(pos.view(np.uint16)).ptr, dev.upload(nrm.view(np.uint16)).ptr, dev.upload(alb).ptr
        cams = np.array([[0, 0, 12], [12, 0, 0], [0, 3, -12], [0, 0, 0]], np.float32)
        cp = cams.ctypes.data_as(C.POINTER(C.c_float))
        s8, s4 = w * h * 8, w * h * 4
        hdr2, hdr, ldr, ldr_only, ldr2 = dev.alloc(nv * s8), dev.alloc(nv * s8), dev.alloc(nv * s4), dev.alloc(nv * s4), dev.alloc(s4)
        dev.lib.lighting_blinnphong_batched(dev.ctx, C.byref(gb), s8, s4, nv, cp, hdr2.ptr, w * 8, s8, None)
        dev.lib.lighting_tonemap_fused_batched(dev.ctx, C.byref(gb), s8, nv, cp, hdr.ptr, w * 8, s8, ldr.ptr, w * 4, s4, None)
        dev.lib.lighting_tonemap_fused_batched(dev.ctx, C.byref(gb), s8, nv, cp, None, 0, 0, ldr_only.ptr, w * 4, s4, None)
        dev.sync()
        got_hdr, got_ldr = hdr.download((nv, h, w, 4), np.uint16), ldr.download((nv, h, w, 4), np.uint8)
        assert np.array_equal(got_hdr, hdr2.download((nv, h, w, 4), np.uint16))
        assert np.array_equal(ldr_only.download((nv, h, w, 4), np.uint8), got_ldr)
        for v in range(nv):
            dev.lib.tonemap_aces_gamma(dev.ctx, hdr2.ptr + v * s8, w * 8, w, h, pbr.WHOLE, ldr2.ptr, w * 4, None)
            dev.sync()
            assert np.array_equal(got_ldr[v], ldr2.download((h, w, 4), np.uint8)), f"view {v}"
            _, want = oracle.lighting(gs[v]["pos"], gs[v]["nrm"], gs[v]["alb"], cams[v])
            assert_hdr_parity(got_hdr[v], want, peak=2.0, what=f"fused batched view {v}")


def test_tonemap_bits_do_not_depend_on_the_path(pbr, dev):
    """K2 has a packed two-texel path (16-byte aligned images) and a single-texel path (row tails, images that are only 4-byte aligned).  They
    used to disagree by 1 LSB on exact rounding ties (5 bytes of a full-HD frame); both now run the same packed arithmetic.  2 M random HDR
    texels through the aligned and through a 4-byte-aligned LDR image must give the same bytes."""
    w, h = 1920, 1080
    rng = np.random.default_rng(11)
    hdr = (rng.random((h, w, 4), dtype=np.float32) ** 3 * 6.0).astype(np.float16)
    src = dev.upload(hdr.view(np.uint16))
    a, b = dev.alloc(w * h * 4), dev.alloc(w * h * 4 + 16)
    dev.lib.tonemap_aces_gamma(dev.ctx, src.ptr, w * 8, w, h, pbr.WHOLE, a.ptr, w * 4, None)
    dev.lib.tonemap_aces_gamma(dev.ctx, src.ptr, w * 8, w, h, pbr.WHOLE, b.ptr + 4, w * 4, None)
    dev.sync()
    va = a.download((h * w, 4), np.uint8)
    vb = b.download((h * w + 4, 4), np.uint8)[1:h * w + 1]
    assert va.any() and np.array_equal(va, vb), f"{np.count_nonzero(va != vb)} bytes differ between the packed and the single-texel path"


def test_frame_host(pbr, dev, oracle):
    w, h = 200, 50
    g = synth_gbuffer(w, h, seed=3)
    cam = (1.0, 2.0, 10.0)
    hdr = np.zeros((h, w, 4), np.uint16)
    ldr = np.zeros((h, w, 4), np.uint8)
    dev.lib.frame_host(dev.ctx, g["pos"].ctypes.data, g["nrm"].ctypes.data, g["alb"].ctypes.data, w, h, pbr.f3(*cam), hdr.ctypes.data, ldr.ctypes.data, None)
    _, want = oracle.lighting(g["pos"], g["nrm"], g["alb"], cam)
    assert_hdr_parity(hdr, want, peak=2.0, what="frame_host")
    assert np.abs(ldr.astype(int) - oracle.tonemap(want).astype(int)).max() <= 2


@pytest.mark.parametrize("w,h", [(1024, 520), (2000, 263), (640, 360)])
@pytest.mark.parametrize("outputs", ["hdr+ldr", "hdr", "ldr"])
def test_frame_host_banded_equals_the_device_path(pbr, dev, oracle, w, h, outputs):
    """pbrcu_frame_host pipelines frames of half a megapixel or more in four tapered row bands (copies-in on one stream,
    K1+K2 and the read-back on another); smaller frames go in one piece.  Both must return the bits of the device-side
    calls on the same planes, for every combination of requested outputs, and leave the rows after the frame alone."""
    g = synth_gbuffer(w, h, seed=w + 3 * h)
    cam = pbr.f3(1.0, 2.0, 10.0)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"])
    hdr_d, ldr_d = dev.alloc(w * h * 8), dev.alloc(w * h * 4)
    dev.lib.lighting_tonemap_fused(dev.ctx, C.byref(gb), cam, pbr.WHOLE, hdr_d.ptr, w * 8, ldr_d.ptr, w * 4, None)
    dev.sync()
    want_hdr, want_ldr = hdr_d.download((h, w, 4), np.uint16), ldr_d.download((h, w, 4), np.uint8)
    hdr = np.full((h + 1, w, 4), 0x7B7B, np.uint16)
    ldr = np.full((h + 1, w, 4), 0x7B, np.uint8)
    for _ in range(2):
        dev.lib.frame_host(dev.ctx, g["pos"].ctypes.data, g["nrm"].ctypes.data, g["alb"].ctypes.data, w, h, cam,
                           hdr.ctypes.data if "hdr" in outputs else None, ldr.ctypes.data if "ldr" in outputs else None, None)
        if "hdr" in outputs:
            assert np.array_equal(hdr[:h], want_hdr) and (hdr[h:] == 0x7B7B).all()
        if "ldr" in outputs:
            assert np.array_equal(ldr[:h], want_ldr) and (ldr[h:] == 0x7B).all()
    hdr_d.free(); ldr_d.free()


def test_full_size_properties(pbr, dev, oracle):
    """BASELINE shapes (4K, 8K) through size-independent properties: device-generated gbuf_spheres,
    cleared texels -> exactly 0, every value in [0, 2], row-band tiling == whole-frame result, and a
    random sample of 4096 texels vs the oracle run on the very bits the device generated."""
    for (w, h) in [(3840, 2160), (7680, 4320)]:
        gb = pbr.GBuffer()
        dev.lib.gbuffer_alloc(dev.ctx, w, h, 0, C.byref(gb))
        cam = (0.0, -1.0, -12.0)
        dev.lib.synth_gbuffer(dev.ctx, C.byref(gb), 1, 0x5EED0002, pbr.f3(*cam), pbr.f3(0, 0, 0), None)
        pitch = gb.pitch8
        out = dev.alloc(pitch * h); out2 = dev.alloc(pitch * h)
        dev.lib.lighting_blinnphong(dev.ctx, C.byref(gb), pbr.f3(*cam), pbr.WHOLE, out.ptr, pitch, None)
        bands = 8
        for r in range(bands):  # the multi-GPU row-band partition, all on one device
            y0, y1 = r * h // bands, (r + 1) * h // bands
            dev.lib.lighting_blinnphong(dev.ctx, C.byref(gb), pbr.f3(*cam), pbr.Rect(0, y0, w, y1 - y0), out2.ptr, pitch, None)
        dev.sync()
        a = out.download((h, w, 4), np.uint16); b = out2.download((h, w, 4), np.uint16)
        assert np.array_equal(a, b), "row-band result differs from whole-frame result"
        nrm = pbr.DeviceBuffer.__new__(pbr.DeviceBuffer); nrm.dev, nrm.nbytes, nrm.ptr = dev, pitch * h, gb.normals
        pos = pbr.DeviceBuffer.__new__(pbr.DeviceBuffer); pos.dev, pos.nbytes, pos.ptr = dev, pitch * h, gb.positions
        alb = pbr.DeviceBuffer.__new__(pbr.DeviceBuffer); alb.dev, alb.nbytes, alb.ptr = dev, pitch * h, gb.albedo
        N = nrm.download((h, w, 4), np.uint16); P = pos.download((h, w, 4), np.uint16); A = alb.download((h, w, 4), np.uint16)
        bg = (N[..., :3] == 0).all(-1)
        assert 0.02 < bg.mean() < 0.9
        assert (a[bg] == 0).all()
        v = f16(a)
        assert np.isfinite(v).all() and v.min() >= 0 and v.max() <= 2.05  # fp16 normals are unit to ~1e-3
        rng = np.random.default_rng(1)
        ys, xs = rng.integers(0, h, 4096), rng.integers(0, w, 4096)
        _, want = oracle.lighting(P[ys, xs][None].view(np.float16), N[ys, xs][None].view(np.float16), A[ys, xs][None], cam)
        assert_hdr_parity(a[ys, xs][None], want, peak=2.0, what=f"{w}x{h} sample")
        out.free(); out2.free()
        dev.lib.gbuffer_free(dev.ctx, C.byref(gb))
