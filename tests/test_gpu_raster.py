"""GPU parity of the geometry pass (K10, csrc/raster.cu) vs the CPU oracle (oracle/pbr_oracle_raster.c), through the
C ABI (pbrcu_geometry_pass).  Coverage (winning triangle per pixel), depth and every plane must be BIT-EXACT: the
oracle's vertex / fragment stages are pinned to the reference GLSL (tests/test_oracle_raster.py), and the device
states the same correctly-rounded fp32 operations in the same order."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest
from oracle import (checker_texture, col_major, look_at_rh, make_draw, mesh_vertex, model_push_constant, perspective_rh_no,
                    raster_scene, uv_sphere)

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).resolve().parent / "golden" / "raster_golden.npz"
EMPTY = 0xFFFFFFFF


CACHE = True  # state vertex_count (the post-transform vertex cache); test_uncached_path_is_identical flips it


def upload_draws(pbr, dev, draws):
    """Device copies of the draws' buffers + the pbrcu_draw array; returns (array, keepalive)."""
    arr = (pbr.Draw * max(1, len(draws)))()
    keep = []
    for i, d in enumerate(draws):
        bufs = {k: dev.upload(d[k]) for k in ("vertices", "indices", "color_tex", "normal_tex")}
        keep.append(bufs)
        a = arr[i]
        a.vertices, a.indices = bufs["vertices"].ptr, bufs["indices"].ptr
        a.first_index, a.index_count, a.vertex_offset = d["first_index"], d["index_count"], d["vertex_offset"]
        a.vertex_count = len(d["vertices"]) - d["vertex_offset"] if CACHE else 0
        a.model[:] = np.asarray(d["model"], np.float32).tolist()
        a.normal_model[:] = np.asarray(d["normal_model"], np.float32).tolist()
        m = a.material
        m.color[:] = [float(c) for c in d["color"]]
        for name in ("color_tex", "normal_tex"):
            t, shp = getattr(m, name), d[name + "_shape"]
            t.texels, t.width, t.height, t.pitch = bufs[name].ptr, shp[1], shp[0], shp[1] * 4
        m.occlusion, m.roughness, m.metallic = d.get("occlusion", 1.0), d.get("roughness", 1.0), d.get("metallic", 1.0)
        m.emissive[:] = [float(c) for c in d.get("emissive", (0.0, 0.0, 0.0))]
    return arr, keep


def device_pass(pbr, dev, draws, view, proj, W, H, extended=True):
    z8, z4 = np.full((H, W, 4), 0x1234, np.uint16), np.full((H, W, 4), 0x55, np.uint8)  # stale contents: the pass must clear
    gb = dev.gbuffer(z8, z8, z8, orm=z4 if extended else None, emissive=z8 if extended else None, depth=np.full((H, W), 0.5, np.float32))
    prim = dev.alloc(W * H * 4)
    arr, keep = upload_draws(pbr, dev, draws)
    v, p = np.ascontiguousarray(view, np.float32), np.ascontiguousarray(proj, np.float32)
    dev.lib.geometry_pass(dev.ctx, C.byref(gb), v.ctypes.data, p.ctypes.data, arr, len(draws), prim.ptr, None)
    dev.sync()

    def plane(ptr, shape, dt):
        b = pbr.DeviceBuffer.__new__(pbr.DeviceBuffer)
        b.dev, b.nbytes, b.ptr = dev, int(np.prod(shape)) * np.dtype(dt).itemsize, ptr
        return b.download(shape, dt)

    out = dict(pos=plane(gb.positions, (H, W, 4), np.uint16), nrm=plane(gb.normals, (H, W, 4), np.uint16), alb=plane(gb.albedo, (H, W, 4), np.uint16),
               depth=plane(gb.depth, (H, W), np.float32), prim=prim.download((H, W), np.uint32))
    if extended:
        out.update(orm=plane(gb.orm, (H, W, 4), np.uint8), emi=plane(gb.emissive, (H, W, 4), np.uint16))
    return out


def assert_same(got, want, what):
    for k in want:
        if k not in got:
            continue
        bad = np.count_nonzero(got[k] != want[k]) if got[k].dtype != np.float32 else np.count_nonzero(got[k].view(np.uint32) != want[k].view(np.uint32))
        assert bad == 0, f"{what}: plane {k} differs in {bad} values"


def camera(eye=(0, -1, -10), target=(0, 0, 0), fov=np.pi / 2, aspect=16 / 9):
    return col_major(look_at_rh(eye, target)), col_major(perspective_rh_no(fov, aspect))


@pytest.mark.parametrize("W,H", [(160, 90), (333, 187), (1, 1)])
def test_scene_bit_exact(pbr, dev, oracle, W, H):
    view, proj = camera(aspect=W / H)
    draws = raster_scene(1)
    want = oracle.geometry_pass(draws, view, proj, W, H, extended=True)
    got = device_pass(pbr, dev, draws, view, proj, W, H)
    assert_same(got, want, f"scene {W}x{H}")
    if (W, H) == (160, 90):
        gold = np.load(GOLDEN)
        assert_same(got, {k: gold[k] for k in gold.files}, "golden frame")


def test_empty_scene_clears(pbr, dev, oracle):
    view, proj = camera()
    got = device_pass(pbr, dev, [], view, proj, 37, 19)
    want = oracle.geometry_pass([], view, proj, 37, 19, extended=True)
    assert_same(got, want, "empty scene")
    assert (got["depth"] == 1.0).all() and (got["prim"] == EMPTY).all()
    # plain (reference-format) G-buffer without the Tier-B planes
    got2 = device_pass(pbr, dev, raster_scene(2, 3), view, proj, 64, 36, extended=False)
    assert_same(got2, oracle.geometry_pass(raster_scene(2, 3), view, proj, 64, 36), "plain G-buffer")


def test_watertight_fan_top_left_rule_and_culling(pbr, dev, oracle):
    view = proj = np.eye(4, dtype=np.float32).reshape(-1)
    n = 97
    ang = np.linspace(0, 2 * np.pi, n + 1)
    v = [mesh_vertex((0.013, -0.021, 0.5))] + [mesh_vertex((3 * np.cos(a), 3 * np.sin(a), 0.5)) for a in ang]
    front, back = [], []
    for k in range(n):
        front += [0, 1 + k, 2 + k]
        back += [0, 2 + k, 1 + k]
    for idx in (front, back, back + front):
        d = [make_draw(v, idx)]
        assert_same(device_pass(pbr, dev, d, view, proj, 64, 48), oracle.geometry_pass(d, view, proj, 64, 48, extended=True), "fan")
    got = device_pass(pbr, dev, [make_draw(v, front)], view, proj, 64, 48)
    assert (got["prim"] != EMPTY).all()


def test_depth_order_and_ties(pbr, dev, oracle):
    view = proj = np.eye(4, dtype=np.float32).reshape(-1)

    def quad(z, size):
        return np.array([mesh_vertex((-size, -size, z)), mesh_vertex((size, -size, z)), mesh_vertex((size, size, z)), mesh_vertex((-size, size, z))], np.float32), [0, 1, 2, 0, 2, 3]

    a, b, c = make_draw(*quad(0.25, 0.5), color=(1, 0, 0, 1)), make_draw(*quad(0.75, 1.0), color=(0, 1, 0, 1)), make_draw(*quad(0.25, 0.5), color=(0, 0, 1, 1))
    for order in ([a, b], [b, a], [a, c], [c, a], [a, b, c], [c, b, a]):
        assert_same(device_pass(pbr, dev, order, view, proj, 32, 32), oracle.geometry_pass(order, view, proj, 32, 32, extended=True), "depth order")
    for z in (1.0, 1.5, -0.25):
        assert (device_pass(pbr, dev, [make_draw(*quad(z, 0.5))], view, proj, 16, 16)["prim"] == EMPTY).all()


def test_clipping_near_plane_and_guard_band(pbr, dev, oracle):
    view, proj = camera((0, -1, -10), (0, -1, 0), aspect=1.0)
    g = 50.0
    v = [mesh_vertex((-g, 1.0, -g), (0, -1, 0), (1, 0, 0, 1), (0, 0)), mesh_vertex((g, 1.0, -g), (0, -1, 0), (1, 0, 0, 1), (4, 0)),
         mesh_vertex((g, 1.0, g), (0, -1, 0), (1, 0, 0, 1), (4, 4)), mesh_vertex((-g, 1.0, g), (0, -1, 0), (1, 0, 0, 1), (0, 4))]
    d = [make_draw(v, [0, 1, 2, 0, 2, 3, 0, 2, 1, 0, 3, 2], color_tex=checker_texture(8, 8, 5))]
    want = oracle.geometry_pass(d, view, proj, 200, 200, extended=True)
    assert 0.3 < (want["prim"] != EMPTY).mean() < 0.7
    assert_same(device_pass(pbr, dev, d, view, proj, 200, 200), want, "near-plane clipping")
    eye = np.eye(4, dtype=np.float32).reshape(-1)
    huge = [mesh_vertex((-5000.0, -4000.0, 0.5)), mesh_vertex((6000.0, -3000.0, 0.5)), mesh_vertex((100.0, 7000.0, 0.5))]
    d2 = [make_draw(huge, [0, 1, 2, 0, 2, 1])]
    want2 = oracle.geometry_pass(d2, eye, eye, 40, 30, extended=True)
    assert (want2["prim"] != EMPTY).all()
    assert_same(device_pass(pbr, dev, d2, eye, eye, 40, 30), want2, "guard-band clipping")


def test_consecutive_passes_reuse_the_visibility_buffer(pbr, dev, oracle):
    """The resolve kernel leaves the visibility buffer empty for the next pass over the same scratch (no fill of its own): passes of
    changing extent and content through ONE context must each equal the oracle -- larger after smaller, smaller after larger, an empty
    scene in between."""
    view, proj = camera(aspect=1.0)
    sv, si = uv_sphere(1.0, 12, 24)
    ball = make_draw(sv, si, model_push_constant((0.3, -0.2, -3.0), (2, 2, 2), (0, 1, 0), 0.3), color_tex=checker_texture(8, 8, 3))
    far = make_draw(sv, si, model_push_constant((-1.0, 0.5, 2.0), (4, 4, 4), (1, 0, 0), 1.1), color=(0.2, 0.9, 0.4, 1.0))
    for draws, W, H in (([ball], 96, 96), ([ball, far], 200, 160), ([far], 64, 48), ([], 200, 160), ([ball, far], 200, 160), ([ball], 31, 17), ([far, ball], 96, 96)):
        assert_same(device_pass(pbr, dev, draws, view, proj, W, H), oracle.geometry_pass(draws, view, proj, W, H, extended=True), f"{len(draws)} draws at {W}x{H}")


def test_edge_function_width_switch(pbr, dev, oracle):
    """Pieces below 64 pixels of extent evaluate their edge functions in 32-bit integers, larger ones in 64-bit (csrc/raster.cu,
    narrow_extent): triangles on either side of the switch, slivers that leave the frame with a small on-screen box, and a fan of
    thin wedges whose extents straddle it, all against the oracle's 64-bit arithmetic."""
    eye = np.eye(4, dtype=np.float32).reshape(-1)
    W = H = 400  # NDC half-extent 1 = 200 pixels: an extent of e pixels is 2 e / 400 in NDC
    cases = []
    for e in (31.7, 62.5, 63.9, 63.998, 64.0, 64.004, 65.3, 126.5, 127.998, 128.004):
        a = e / 200.0
        cases.append([mesh_vertex((-0.4, -0.45, 0.3)), mesh_vertex((-0.4 + a, -0.45 + 0.11, 0.6)), mesh_vertex((-0.4 + 0.07, -0.45 + a, 0.9))])
    # sliver through the corner: tiny box on screen, edges far longer than 128 pixels
    cases.append([mesh_vertex((-1.004, -1.004, 0.5)), mesh_vertex((3.0, -1.3, 0.5)), mesh_vertex((-1.3, 3.0, 0.5))])
    cases.append([mesh_vertex((0.99, 0.99, 0.2)), mesh_vertex((0.9975, 40.0, 0.2)), mesh_vertex((40.0, 0.9975, 0.2))])
    for tri in cases:
        d = [make_draw(tri, [0, 1, 2, 0, 2, 1])]
        want = oracle.geometry_pass(d, eye, eye, W, H, extended=True)
        assert (want["prim"] != EMPTY).any()
        assert_same(device_pass(pbr, dev, d, eye, eye, W, H), want, "edge width switch")
    n = 181
    ang = np.linspace(0, 2 * np.pi, n + 1)
    rad = 0.3 + 0.04 * np.cos(7 * ang)  # 52 .. 68 pixels from the hub
    v = [mesh_vertex((0.003, -0.002, 0.5))] + [mesh_vertex((r * np.cos(a), r * np.sin(a), 0.25 + 0.2 * np.sin(a))) for a, r in zip(ang, rad)]
    idx = []
    for k in range(n):
        idx += [0, 1 + k, 2 + k]
    d = [make_draw(v, idx)]
    assert_same(device_pass(pbr, dev, d, eye, eye, W, H), oracle.geometry_pass(d, eye, eye, W, H, extended=True), "wedge fan")


def test_large_and_small_triangles_full_hd(pbr, dev, oracle):
    """1920x1080: screen-filling triangles go through the chunk queue, a dense sphere through the per-triangle warps;
    first_index / vertex_offset (MeshBuilder-style concatenated primitives) are honoured."""
    W, H = 1920, 1080
    view, proj = camera(aspect=W / H)
    sv, si = uv_sphere(1.0, 48, 96)
    g = 40.0
    quad = np.array([mesh_vertex((-g, -g, 6.0), (0, 0, -1), (1, 0, 0, 1), (0, 0)), mesh_vertex((g, -g, 6.0), (0, 0, -1), (1, 0, 0, 1), (16, 0)),
                     mesh_vertex((g, g, 6.0), (0, 0, -1), (1, 0, 0, 1), (16, 16)), mesh_vertex((-g, g, 6.0), (0, 0, -1), (1, 0, 0, 1), (0, 16))], np.float32)
    verts = np.concatenate([quad, sv])
    idx = np.concatenate([np.array([0, 1, 2, 0, 2, 3, 0, 2, 1, 0, 3, 2], np.uint16), si])
    wall = make_draw(verts, idx, color_tex=checker_texture(64, 64, 9), first_index=0, index_count=12, vertex_offset=0)
    ball = make_draw(verts, idx, model_push_constant((0.5, -0.5, -4.0), (3, 3, 3), (1, 1, 0), 0.7), color=(0.9, 0.7, 0.5, 1.0), color_tex=checker_texture(32, 16, 10),
                     normal_tex=checker_texture(16, 16, 11, normal_map=True), first_index=12, index_count=len(si), vertex_offset=4, roughness=0.3, metallic=0.0)
    draws = [wall, ball]
    want = oracle.geometry_pass(draws, view, proj, W, H, extended=True)
    assert (want["prim"] != EMPTY).mean() > 0.9
    assert_same(device_pass(pbr, dev, draws, view, proj, W, H), want, "full HD")


def test_geometry_pass_argument_errors(pbr, dev):
    view = proj = np.eye(4, dtype=np.float32).reshape(-1)
    gb = dev.gbuffer(np.zeros((4, 4, 4), np.uint16), np.zeros((4, 4, 4), np.uint16), np.zeros((4, 4, 4), np.uint16))
    arr = (pbr.Draw * 1)()
    with pytest.raises(pbr.PbrcuError):  # NULL vertex buffer
        dev.lib.geometry_pass(dev.ctx, C.byref(gb), view.ctypes.data, proj.ctypes.data, arr, 1, None, None)
    v = dev.upload(np.zeros((3, 12), np.float32)); i = dev.upload(np.array([0, 1, 2], np.uint16))
    arr[0].vertices, arr[0].indices, arr[0].index_count = v.ptr, i.ptr, 3
    with pytest.raises(pbr.PbrcuError):  # the reference's materials always carry both textures
        dev.lib.geometry_pass(dev.ctx, C.byref(gb), view.ctypes.data, proj.ctypes.data, arr, 1, None, None)


def test_uncached_path_is_identical(pbr, dev, oracle, monkeypatch):
    """vertex_count = 0 (the caller does not know how many vertices a draw addresses): the vertex stage runs per
    triangle and per pixel instead of once per vertex -- same bits; also a mix of cached and uncached draws."""
    import sys
    me = sys.modules[__name__]
    view, proj = camera((0.5, -1.5, -6.0), (0, 0, 0), aspect=16 / 9)  # close enough that the ground plane is clipped
    draws = raster_scene(1)
    want = oracle.geometry_pass(draws, view, proj, 200, 112, extended=True)
    cached = device_pass(pbr, dev, draws, view, proj, 200, 112)
    assert_same(cached, want, "cached")
    monkeypatch.setattr(me, "CACHE", False)
    assert_same(device_pass(pbr, dev, draws, view, proj, 200, 112), want, "uncached")
    monkeypatch.setattr(me, "CACHE", True)
    orig = me.upload_draws

    def mixed(pbr_, dev_, ds):
        arr, keep = orig(pbr_, dev_, ds)
        for i in range(0, len(ds), 2):
            arr[i].vertex_count = 0
        return arr, keep

    monkeypatch.setattr(me, "upload_draws", mixed)
    assert_same(device_pass(pbr, dev, draws, view, proj, 200, 112), want, "mixed")


def test_clip_cache_overflow_falls_back(pbr, dev, oracle, monkeypatch):
    """More clipped triangles than parked-polygon slots (forced: capacity 1): the rest re-clip per pixel, same bits."""
    view, proj = camera((0.5, -1.5, -6.0), (0, 0, 0), aspect=16 / 9)
    draws = raster_scene(1)
    want = oracle.geometry_pass(draws, view, proj, 200, 112, extended=True)
    for cap in ("1", "0"):
        monkeypatch.setenv("PBRCU_RASTER_CLIP_CAP", cap)
        assert_same(device_pass(pbr, dev, draws, view, proj, 200, 112), want, f"clip cap {cap}")


def test_chunk_queue_overflow_falls_back(pbr, dev, oracle, monkeypatch):
    """Screen-sized triangles hand their 64 x 32 chunks to the chunk queue.  With the queue forced to overflow (capacity 0,
    one slot, part of a reservation, exactly one triangle's chunks) the producers that do not fit keep their piece with the
    warp loop, their reserved slots carry the skip sentinel, and the consumer never reads a slot nobody wrote: same bits."""
    W, H = 320, 200
    view, proj = camera(aspect=W / H)
    g = 40.0
    quad = np.array([mesh_vertex((-g, -g, 6.0), (0, 0, -1), (1, 0, 0, 1), (0, 0)), mesh_vertex((g, -g, 6.0), (0, 0, -1), (1, 0, 0, 1), (16, 0)),
                     mesh_vertex((g, g, 6.0), (0, 0, -1), (1, 0, 0, 1), (16, 16)), mesh_vertex((-g, g, 6.0), (0, 0, -1), (1, 0, 0, 1), (0, 16))], np.float32)
    idx = np.array([0, 1, 2, 0, 2, 3, 0, 2, 1, 0, 3, 2], np.uint16)
    walls = [make_draw(quad, idx, model_push_constant((0.3 * k, -0.2 * k, float(k)), (1, 1, 1), (0, 0, 1), 0.1 * k), color=(0.9, 0.5 + 0.1 * k, 0.4, 1.0),
                       color_tex=checker_texture(64, 64, 9 + k)) for k in range(3)]
    want = oracle.geometry_pass(walls, view, proj, W, H, extended=True)
    assert (want["prim"] != EMPTY).mean() > 0.9
    assert_same(device_pass(pbr, dev, walls, view, proj, W, H), want, "full queue")
    for cap in ("0", "1", "17", "35", "36", "71"):
        monkeypatch.setenv("PBRCU_RASTER_QUEUE_CAP", cap)
        assert_same(device_pass(pbr, dev, walls, view, proj, W, H), want, f"chunk queue cap {cap}")


def test_batched_views_equal_single_passes(pbr, dev, oracle):
    """pbrcu_geometry_pass_batched: six cameras (more than the four internal streams / scratch sets, so sets are re-used) over one
    scene, one of them close enough that the ground plane is clipped -- every view's planes equal a pbrcu_geometry_pass call with
    that camera bit for bit, and the first view the oracle's; run twice (the scratch sets and tables of the first call are re-used)."""
    W, H, nv = 200, 112, 6
    draws = raster_scene(1)
    cams = [camera((0.5, -1.5, -6.0), (0, 0, 0), aspect=W / H)] + [camera((8 * np.cos(a), -1.0 - a, 8 * np.sin(a)), (0, 0, 0), aspect=W / H) for a in np.linspace(0.3, 5.5, nv - 1)]
    singles = [device_pass(pbr, dev, draws, v, p, W, H) for v, p in cams]
    assert_same(singles[0], oracle.geometry_pass(draws, cams[0][0], cams[0][1], W, H, extended=True), "view 0 vs oracle")
    s8, s4 = W * H * 8, W * H * 4
    stale8, stale4 = np.full((nv, H, W, 4), 0x1234, np.uint16), np.full((nv, H, W, 4), 0x55, np.uint8)
    bufs = dict(pos=dev.upload(stale8), nrm=dev.upload(stale8), alb=dev.upload(stale8), emi=dev.upload(stale8), orm=dev.upload(stale4),
                depth=dev.upload(np.full((nv, H, W), 0.5, np.float32)))
    g0 = pbr.GBuffer()
    g0.width, g0.height, g0.pitch8, g0.pitch4 = W, H, W * 8, W * 4
    g0.positions, g0.normals, g0.albedo, g0.emissive, g0.orm, g0.depth = (bufs[k].ptr for k in ("pos", "nrm", "alb", "emi", "orm", "depth"))
    arr, keep = upload_draws(pbr, dev, draws)
    views = np.ascontiguousarray(np.stack([np.asarray(v, np.float32).reshape(16) for v, _ in cams]))
    projs = np.ascontiguousarray(np.stack([np.asarray(p, np.float32).reshape(16) for _, p in cams]))
    for attempt in range(2):
        dev.lib.geometry_pass_batched(dev.ctx, C.byref(g0), s8, s4, nv, views.ctypes.data, projs.ctypes.data, arr, len(draws), None)
        dev.sync()
        got = dict(pos=bufs["pos"].download((nv, H, W, 4), np.uint16), nrm=bufs["nrm"].download((nv, H, W, 4), np.uint16), alb=bufs["alb"].download((nv, H, W, 4), np.uint16),
                   emi=bufs["emi"].download((nv, H, W, 4), np.uint16), orm=bufs["orm"].download((nv, H, W, 4), np.uint8), depth=bufs["depth"].download((nv, H, W), np.float32))
        for v in range(nv):
            assert_same({k: a[v] for k, a in got.items()}, {k: a for k, a in singles[v].items() if k != "prim"}, f"attempt {attempt}, view {v}")
