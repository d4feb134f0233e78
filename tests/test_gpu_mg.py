"""GPU: the C multi-GPU host (pbrcu_mg_*, csrc/mg.cu) through the C ABI.  A group may list one device several times, so the
band cuts, the per-member clustering, the stores into the other members' frames / chains and the event barrier are covered on
a ONE-GPU box; with two or more GPUs the same tests also run over distinct devices, plus the NCCL gather arm.  Every result is
compared bit for bit with the single-device entry points on the same inputs (which the other tests pin to the oracle)."""
import ctypes as C

import numpy as np
import pytest
from oracle import synth_gbuffer, synth_lights
from test_gpu_tier_b import build_ibl

pytestmark = pytest.mark.gpu


class Member:
    """Allocations and copies in one member's memory (its device made current first)."""

    def __init__(self, pbr, lib, mg, i, ctx=None):
        self.pbr, self.lib, self.mg, self.i = pbr, lib, mg, i
        self.ctx = ctx if mg is None else lib.mg_ctx(mg, i)
        self.ptrs = []

    @classmethod
    def single(cls, pbr, lib, ctx):
        """The same helpers on a plain single-device context (device 0 current)."""
        return cls(pbr, lib, None, 0, ctx)

    def on(self):
        if self.mg is not None:
            self.lib.mg_set_device(self.mg, self.i)

    def alloc(self, nbytes):
        self.on()
        p = C.c_void_p()
        self.lib.malloc(self.ctx, max(int(nbytes), 16), C.byref(p))
        self.ptrs.append(p.value)
        return p.value

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        p = self.alloc(arr.nbytes)
        self.lib.memcpy_h2d(p, arr.ctypes.data, arr.nbytes, None)
        self.lib.stream_synchronize(None)
        return p

    def download(self, ptr, shape, dtype):
        self.on()
        out = np.empty(shape, dtype)
        self.lib.memcpy_d2h(out.ctypes.data, ptr, out.nbytes, None)
        self.lib.stream_synchronize(None)
        return out

    def gbuffer(self, g):
        h, w = g["pos"].shape[:2]
        gb = self.pbr.GBuffer()
        gb.width, gb.height, gb.pitch8, gb.pitch4 = w, h, w * 8, w * 4
        gb.positions, gb.normals, gb.albedo = self.upload(g["pos"].view(np.uint16)), self.upload(g["nrm"].view(np.uint16)), self.upload(g["alb"])
        gb.orm, gb.emissive = self.upload(g["orm"]), self.upload(g["emi"].view(np.uint16))
        return gb

    def ibl(self, host, accel=True):
        s = self.pbr.Ibl()
        s.irradiance, s.irr_res = self.upload(host["irr"]), host["irr"].shape[1]
        s.prefiltered, s.pref_res, s.pref_mips = self.upload(host["pref"]), host["pref_res"], host["pref_mips"]
        s.lut, s.lut_res = self.upload(host["lut"]), host["lut"].shape[0]
        if accel:
            a = C.c_void_p()
            self.lib.ibl_accel_build(self.ctx, C.byref(s), C.byref(a), None)
            self.lib.stream_synchronize(None)
            s.accel = a
        return s

    def free(self):
        self.on()
        for p in self.ptrs:
            self.lib.free(self.ctx, p)
        self.ptrs = []


def _device0_current(lib):
    """Leaves device 0 current for the tests that follow (the single-device fixtures assume it)."""
    d0 = C.c_void_p()
    lib.mg_create((C.c_int * 1)(0), 1, C.byref(d0))
    lib.mg_set_device(d0, 0)
    lib.mg_destroy(d0)


def groups(pbr):
    n = C.c_int()
    pbr.load().device_count(C.byref(n))
    out = [[0], [0, 0], [0, 0, 0]]
    if n.value >= 2:
        out.append(list(range(min(n.value, 8))))
    return out


@pytest.fixture(scope="module")
def ibl_host(dev, oracle):
    host, bufs = build_ibl(dev, oracle)
    return host


@pytest.mark.parametrize("tiled", [False, True])
def test_mg_frame_equals_the_single_device_frame(pbr, dev, ibl_host, tiled):
    lib = pbr.load()
    w, h = 96, 112
    nl = 40 if tiled else 4
    g = synth_gbuffer(w, h, seed=41 + nl, tier_b=True)
    lights = synth_lights(nl, seed=8, ranged=tiled)
    cam = pbr.f3(0.5, -1.0, 12.0)
    # the frame on one device: the comparison everything below must reproduce bit for bit
    gb0 = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    one = Member.single(pbr, lib, dev.ctx)
    ibl0 = one.ibl(ibl_host)
    out0 = dev.alloc(w * h * 8)
    counts = None
    if tiled:
        ld = dev.upload(lights)
        tiles0 = pbr.LightTiles()
        lib.light_tiles_alloc(dev.ctx, w, h, nl, C.byref(tiles0))
        lib.cluster_lights(dev.ctx, C.byref(gb0), ld.ptr, nl, C.byref(tiles0), None)
        lib.shade_ggx_ibl(dev.ctx, C.byref(gb0), cam, ld.ptr, nl, C.byref(tiles0), C.byref(ibl0), pbr.WHOLE, out0.ptr, w * 8, None)
        dev.sync()
        counts = np.empty(tiles0.tiles_x * tiles0.tiles_y, np.uint32)
        lib.memcpy_d2h(counts.ctypes.data, tiles0.counts, counts.nbytes, None)
        dev.sync()
        lib.light_tiles_free(dev.ctx, C.byref(tiles0))
    else:
        lib.shade_ggx_ibl(dev.ctx, C.byref(gb0), cam, lights.ctypes.data, nl, None, C.byref(ibl0), pbr.WHOLE, out0.ptr, w * 8, None)
        dev.sync()
    want = out0.download((h, w, 4), np.uint16)
    ldr0 = dev.alloc(w * h * 4)
    lib.tonemap_aces_gamma(dev.ctx, out0.ptr, w * 8, w, h, pbr.WHOLE, ldr0.ptr, w * 4, None)
    dev.sync()
    want_ldr = ldr0.download((h, w, 4), np.uint8)
    lib.ibl_accel_free(dev.ctx, ibl0.accel)

    for devices in groups(pbr):
        n = len(devices)
        mg = C.c_void_p()
        lib.mg_create((C.c_int * n)(*devices), n, C.byref(mg))
        assert lib.mg_size(mg) == n
        members = [Member(pbr, lib, mg, i) for i in range(n)]
        gbs = (pbr.GBuffer * n)(*[m.gbuffer(g) for m in members])
        ibls = (pbr.Ibl * n)(*[m.ibl(ibl_host) for m in members])
        tiles = None
        if tiled:
            tiles = (pbr.LightTiles * n)()
            for i, m in enumerate(members):
                m.on()
                lib.light_tiles_alloc(m.ctx, w, h, nl, C.byref(tiles[i]))
        bands = (C.c_uint32 * (2 * n))()
        costs = None
        if tiled:  # bands of equal cost from the frame's own light lists ("the previous frame")
            costs = (9.0 + counts.reshape(-1, (w + 15) // 16)).sum(axis=1).astype(np.float32)
        lib.mg_row_bands(h, n, costs.ctypes.data_as(C.POINTER(C.c_float)) if costs is not None else None, bands)
        modes = [0, 1] + ([2] if n > 1 and len(set(devices)) == n else [])
        for gather in modes:
            frames = [m.alloc(w * h * 8) for m in members]
            for m, f in zip(members, frames):
                m.on()
                lib.memset(f, 0x5A, w * h * 8, None)
                lib.stream_synchronize(None)
            lib.mg_shade_frame(mg, gbs, cam, lights.ctypes.data, nl, tiles, ibls, bands, (C.c_void_p * n)(*frames), w * 8, gather)
            lib.mg_synchronize(mg)
            for i, (m, f) in enumerate(zip(members, frames)):
                got = m.download(f, (h, w, 4), np.uint16)
                y0, y1 = bands[2 * i], bands[2 * i + 1]
                if gather == 0:
                    assert np.array_equal(got[y0:y1], want[y0:y1]), f"{devices} member {i}: band differs"
                    assert (got[:y0] == 0x5A5A).all() and (got[y1:] == 0x5A5A).all(), f"{devices} member {i}: wrote outside its band"
                else:
                    assert np.array_equal(got, want), f"{devices} gather {gather} member {i}: {np.count_nonzero(got != want)} values differ from the single-device frame"
        # the same frame delivered display-ready (tonemap fused into the stores, 4-byte texels on the links)
        ldr_frames = [m.alloc(w * h * 4) for m in members]
        lib.mg_shade_frame_ldr(mg, gbs, cam, lights.ctypes.data, nl, tiles, ibls, bands, (C.c_void_p * n)(*ldr_frames), w * 4, 1)
        lib.mg_synchronize(mg)
        for i, (m, f) in enumerate(zip(members, ldr_frames)):
            assert np.array_equal(m.download(f, (h, w, 4), np.uint8), want_ldr), f"{devices} member {i}: LDR frame differs from tonemap(single-device frame)"
        for i, m in enumerate(members):
            m.on()
            if tiled:
                lib.light_tiles_free(m.ctx, C.byref(tiles[i]))
            lib.ibl_accel_free(m.ctx, ibls[i].accel)
            m.free()
        lib.mg_destroy(mg)
    _device0_current(lib)


def test_mg_views_and_prefilter(pbr, dev, oracle, ibl_host):
    lib = pbr.load()
    w, h, nv, nl = 48, 32, 7, 3
    gs = [synth_gbuffer(w, h, seed=600 + v, tier_b=True) for v in range(nv)]
    lights = synth_lights(nl, seed=12)
    cams = np.random.default_rng(2).uniform(-12, 12, (nv, 3)).astype(np.float32)
    stack = lambda k, dt, sel: np.ascontiguousarray(np.stack([np.asarray(gs[v][k]).view(dt) for v in sel]))
    s8, s4 = w * h * 8, w * h * 4

    def block(m, sel):
        g0 = pbr.GBuffer()
        g0.width, g0.height, g0.pitch8, g0.pitch4 = w, h, w * 8, w * 4
        if len(sel):
            g0.positions, g0.normals, g0.albedo = m.upload(stack("pos", np.uint16, sel)), m.upload(stack("nrm", np.uint16, sel)), m.upload(stack("alb", np.uint16, sel))
            g0.orm, g0.emissive = m.upload(stack("orm", np.uint8, sel)), m.upload(stack("emi", np.uint16, sel))
        return g0

    one = Member.single(pbr, lib, dev.ctx)
    ibl0 = one.ibl(ibl_host)
    g_all = block(one, range(nv))
    out0 = dev.alloc(nv * s8)
    lib.shade_ggx_ibl_batched(dev.ctx, C.byref(g_all), s8, s4, nv, cams.ctypes.data_as(C.POINTER(C.c_float)), lights.ctypes.data, nl, C.byref(ibl0), out0.ptr, w * 8, s8, None)
    dev.sync()
    want = out0.download((nv, h, w, 4), np.uint16)
    # prefilter: the unsharded chain from a prepared source
    res, mips, ns = 32, 6, 64
    env_chain = ibl_host["pref"]  # any RGBA16F chain of that shape serves as the source
    src0 = one.upload(env_chain)
    source0 = C.c_void_p()
    lib.ibl_source_create(dev.ctx, src0, res, mips, None, C.byref(source0))
    dst0 = dev.alloc(env_chain.nbytes)
    lib.ibl_prefilter_ggx_prepared(dev.ctx, source0, ns, dst0.ptr, 0, mips - 1, 0, 1, None, 0, None)
    dev.sync()
    want_chain = dst0.download(env_chain.shape, np.uint16)
    lib.ibl_source_free(dev.ctx, source0)
    lib.ibl_accel_free(dev.ctx, ibl0.accel)

    for devices in groups(pbr):
        n = len(devices)
        mg = C.c_void_p()
        lib.mg_create((C.c_int * n)(*devices), n, C.byref(mg))
        members = [Member(pbr, lib, mg, i) for i in range(n)]
        ibls = (pbr.Ibl * n)(*[m.ibl(ibl_host) for m in members])
        cuts = [nv * k // n for k in range(n + 1)]
        sels = [list(range(cuts[k], cuts[k + 1])) for k in range(n)]
        g0s = (pbr.GBuffer * n)(*[block(m, sel) for m, sel in zip(members, sels)])
        outs = [m.alloc(max(len(sel), 1) * s8) for m, sel in zip(members, sels)]
        cam_blocks = [np.ascontiguousarray(cams[sel]) if len(sel) else np.zeros((1, 3), np.float32) for sel in sels]
        cam_ptrs = (C.POINTER(C.c_float) * n)(*[c.ctypes.data_as(C.POINTER(C.c_float)) for c in cam_blocks])
        lib.mg_shade_views(mg, g0s, s8, s4, (C.c_uint32 * n)(*[len(s) for s in sels]), cam_ptrs, lights.ctypes.data, nl, ibls, (C.c_void_p * n)(*outs), w * 8, s8)
        lib.mg_synchronize(mg)
        for m, sel, o in zip(members, sels, outs):
            if len(sel):
                assert np.array_equal(m.download(o, (len(sel), h, w, 4), np.uint16), want[sel]), f"{devices}: views {sel} differ"
        # sharded prefilter, every finished texel stored into every member's chain
        sources = []
        for m in members:
            m.on()
            s = C.c_void_p()
            lib.ibl_source_create(m.ctx, m.upload(env_chain), res, mips, None, C.byref(s))
            lib.stream_synchronize(None)
            sources.append(s)
        chains = [m.alloc(env_chain.nbytes) for m in members]
        for m, c in zip(members, chains):
            m.on()
            lib.memset(c, 0, env_chain.nbytes, None)
            lib.stream_synchronize(None)
        lib.mg_prefilter(mg, (C.c_void_p * n)(*[s.value for s in sources]), mips, ns, (C.c_void_p * n)(*chains), 1)
        lib.mg_synchronize(mg)
        for i, (m, c) in enumerate(zip(members, chains)):
            assert np.array_equal(m.download(c, env_chain.shape, np.uint16), want_chain), f"{devices}: chain of member {i} differs from the unsharded one"
        for i, m in enumerate(members):
            m.on()
            lib.ibl_source_free(m.ctx, sources[i])
            lib.ibl_accel_free(m.ctx, ibls[i].accel)
            m.free()
        lib.mg_destroy(mg)
    _device0_current(lib)


def test_mg_geometry_views_then_shading(pbr, dev, oracle, ibl_host):
    """BASELINE configs[4] as written, through the C multi-GPU host: every member rasterises its block of camera views of one scene
    (pbrcu_mg_geometry_views) and shades them (pbrcu_mg_shade_views).  G-buffer planes and HDR frames of every view must equal the
    single-device pbrcu_geometry_pass + pbrcu_shade_ggx_ibl_batched results bit for bit (which test_gpu_raster pins to the oracle)."""
    from oracle import raster_scene
    from test_gpu_raster import camera, upload_draws
    lib = pbr.load()
    W, H, nv, nl = 96, 64, 5, 3
    scene = raster_scene(1)
    cams_eye = [(8 * np.cos(a), -1.0 - 0.5 * a, 8 * np.sin(a)) for a in np.linspace(0.2, 5.0, nv)]
    cams = [camera(e, (0, 0, 0), aspect=W / H) for e in cams_eye]
    views = np.ascontiguousarray(np.stack([np.asarray(v, np.float32).reshape(16) for v, _ in cams]))
    projs = np.ascontiguousarray(np.stack([np.asarray(p, np.float32).reshape(16) for _, p in cams]))
    eyes = np.ascontiguousarray(np.array(cams_eye, np.float32))
    lights = synth_lights(nl, seed=21)
    s8, s4 = W * H * 8, W * H * 4
    names8, names4 = ("positions", "normals", "albedo", "emissive"), ("orm", "depth")

    def gbuffers(m, n):
        g0 = pbr.GBuffer()
        g0.width, g0.height, g0.pitch8, g0.pitch4 = W, H, W * 8, W * 4
        for k in names8:
            setattr(g0, k, m.alloc(max(n, 1) * s8))
        for k in names4:
            setattr(g0, k, m.alloc(max(n, 1) * s4))
        return g0

    def planes(m, g0, n):
        return {k: m.download(getattr(g0, k), (n, H, W, 4) if k != "depth" else (n, H, W), np.uint16 if k in names8 else (np.uint8 if k == "orm" else np.float32))
                for k in names8 + names4}

    # single device: the reference for every view
    one = Member.single(pbr, lib, dev.ctx)
    arr0, keep0 = upload_draws(pbr, dev, scene)
    g_all = gbuffers(one, nv)
    lib.geometry_pass_batched(dev.ctx, C.byref(g_all), s8, s4, nv, views.ctypes.data, projs.ctypes.data, arr0, len(scene), None)
    ibl0 = one.ibl(ibl_host)
    out0 = dev.alloc(nv * s8)
    lib.shade_ggx_ibl_batched(dev.ctx, C.byref(g_all), s8, s4, nv, eyes.ctypes.data_as(C.POINTER(C.c_float)), lights.ctypes.data, nl, C.byref(ibl0), out0.ptr, W * 8, s8, None)
    dev.sync()
    want_planes, want_hdr = planes(one, g_all, nv), out0.download((nv, H, W, 4), np.uint16)
    assert (want_planes["normals"] != 0).any()
    lib.ibl_accel_free(dev.ctx, ibl0.accel)

    class MemberDev:  # what upload_draws needs of a device: upload() into this member's memory
        def __init__(self, m):
            self.m = m

        def upload(self, a):
            class B:
                pass
            b = B()
            b.ptr = self.m.upload(np.ascontiguousarray(a))
            return b

    for devices in groups(pbr):
        n = len(devices)
        mg = C.c_void_p()
        lib.mg_create((C.c_int * n)(*devices), n, C.byref(mg))
        members = [Member(pbr, lib, mg, i) for i in range(n)]
        cuts = [nv * k // n for k in range(n + 1)]
        counts = [cuts[k + 1] - cuts[k] for k in range(n)]
        g0s = (pbr.GBuffer * n)(*[gbuffers(m, c) for m, c in zip(members, counts)])
        draw_arrays = [upload_draws(pbr, MemberDev(m), scene) for m in members]
        draws = (C.POINTER(pbr.Draw) * n)(*[C.cast(a, C.POINTER(pbr.Draw)) for a, _ in draw_arrays])
        vblocks = [np.ascontiguousarray(views[cuts[k]:cuts[k + 1]]) if counts[k] else np.zeros((1, 16), np.float32) for k in range(n)]
        pblocks = [np.ascontiguousarray(projs[cuts[k]:cuts[k + 1]]) if counts[k] else np.zeros((1, 16), np.float32) for k in range(n)]
        eblocks = [np.ascontiguousarray(eyes[cuts[k]:cuts[k + 1]]) if counts[k] else np.zeros((1, 3), np.float32) for k in range(n)]
        fptr = lambda blocks: (C.POINTER(C.c_float) * n)(*[b.ctypes.data_as(C.POINTER(C.c_float)) for b in blocks])
        lib.mg_geometry_views(mg, g0s, s8, s4, (C.c_uint32 * n)(*counts), fptr(vblocks), fptr(pblocks), draws, (C.c_uint32 * n)(*[len(scene)] * n))
        ibls = (pbr.Ibl * n)(*[m.ibl(ibl_host) for m in members])
        outs = [m.alloc(max(c, 1) * s8) for m, c in zip(members, counts)]
        lib.mg_shade_views(mg, g0s, s8, s4, (C.c_uint32 * n)(*counts), fptr(eblocks), lights.ctypes.data, nl, ibls, (C.c_void_p * n)(*outs), W * 8, s8)
        lib.mg_synchronize(mg)
        for k, (m, c) in enumerate(zip(members, counts)):
            if not c:
                continue
            got = planes(m, g0s[k], c)
            for name in got:
                a, b = got[name], want_planes[name][cuts[k]:cuts[k + 1]]
                assert np.array_equal(a.view(np.uint32) if name == "depth" else a, b.view(np.uint32) if name == "depth" else b), f"{devices} member {k}: plane {name} differs"
            assert np.array_equal(m.download(outs[k], (c, H, W, 4), np.uint16), want_hdr[cuts[k]:cuts[k + 1]]), f"{devices} member {k}: shaded views differ"
        for i, m in enumerate(members):
            m.on()
            lib.ibl_accel_free(m.ctx, ibls[i].accel)
            m.free()
        lib.mg_destroy(mg)
    _device0_current(lib)
