"""CPU: the geometry-pass oracle (oracle/pbr_oracle_raster.c).
Vertex and fragment stages are PINNED bit-for-bit to the reference's own GLSL (geometry_pass/vertex.glsl,
geometry_pass/fragment.glsl compiled for the CPU, oracle/_ref); the fixed-function rasteriser between them
lives in the Vulkan driver, not in the reference, and is checked through its defining properties (watertight
top-left coverage, back-face culling with front = clockwise, LESS depth test, near clipping) and a committed
golden frame (tests/golden/raster_golden.npz, made by tests/golden/make_raster_golden.py)."""
from pathlib import Path

import numpy as np
import pytest
from oracle import (checker_texture, col_major, look_at_rh, make_draw, mesh_vertex, model_push_constant, perspective_rh_no,
                    raster_scene)

GOLDEN = Path(__file__).resolve().parent / "golden" / "raster_golden.npz"
EMPTY = 0xFFFFFFFF


def camera(eye=(0, -1, -10), target=(0, 0, 0), fov=np.pi / 2, aspect=16 / 9):
    return col_major(look_at_rh(eye, target)), col_major(perspective_rh_no(fov, aspect))


def test_vertex_stage_bit_exact_vs_reference_glsl(oracle, ref_shaders):
    rng = np.random.default_rng(3)
    view, proj = camera((1.5, -2.0, -7.0), (0.2, 0.1, 0.0))
    for k in range(200):
        model, nm = model_push_constant(rng.uniform(-3, 3, 3), rng.uniform(0.3, 2.0, 3), rng.standard_normal(3), rng.uniform(0, 6.28))
        n = rng.standard_normal(3); n /= np.linalg.norm(n)
        t = np.cross(n, rng.standard_normal(3)); t /= np.linalg.norm(t)
        vtx = np.array(mesh_vertex(rng.uniform(-2, 2, 3), n, (*t, rng.choice([-1.0, 1.0])), rng.uniform(-2, 2, 2)), np.float32)
        got, want = oracle.geom_vertex(model, nm, view, proj, vtx), ref_shaders.geom_vertex(model, nm, view, proj, vtx)
        for g, w in zip(got, want):
            assert np.array_equal(g.view(np.uint32), w.view(np.uint32)), (k, g, w)


def test_fragment_stage_bit_exact_vs_reference_glsl(oracle, ref_shaders):
    rng = np.random.default_rng(4)
    ctex, ntex = checker_texture(16, 8, 1), checker_texture(8, 4, 2, normal_map=True)
    for k in range(200):
        attrs = rng.uniform(-1, 1, 14).astype(np.float32)
        attrs[12:] = rng.uniform(-3, 3, 2)  # uv outside [0,1): repeat addressing
        color = rng.uniform(0, 1.2, 4).astype(np.float32)
        got, want = oracle.geom_fragment(attrs, color, ctex, ntex), ref_shaders.geom_fragment(attrs, color, ctex, ntex)
        for g, w in zip(got, want):
            assert np.array_equal(g.view(np.uint32), w.view(np.uint32)), (k, g, w)


def quad(z=0.0, size=1.0, flip=False):
    v = np.array([mesh_vertex((-size, -size, z)), mesh_vertex((size, -size, z)), mesh_vertex((size, size, z)), mesh_vertex((-size, size, z))], np.float32)
    idx = [0, 2, 1, 0, 3, 2] if flip else [0, 1, 2, 0, 2, 3]
    return v, idx


def ortho_like():
    """view = identity, proj maps z in [0,1] straight through with w = 1: clip = (x, y, z, 1)."""
    return np.eye(4, dtype=np.float32).reshape(-1), np.eye(4, dtype=np.float32).reshape(-1)


def test_empty_scene_is_the_clear_values(oracle):
    view, proj = camera()
    out = oracle.geometry_pass([], view, proj, 33, 17, extended=True)
    assert not out["pos"].any() and not out["nrm"].any() and not out["alb"].any() and not out["orm"].any() and not out["emi"].any()
    assert (out["depth"] == 1.0).all() and (out["prim"] == EMPTY).all()


def test_watertight_no_gaps_no_double_hits(oracle):
    """A screen-covering fan of thin triangles sharing one apex: every pixel is covered by exactly the triangle
    whose interior holds its centre (top-left rule on the shared edges) -- no pixel is left empty."""
    view, proj = ortho_like()
    n = 97
    ang = np.linspace(0, 2 * np.pi, n + 1)
    v = [mesh_vertex((0.013, -0.021, 0.5))] + [mesh_vertex((3 * np.cos(a), 3 * np.sin(a), 0.5)) for a in ang]
    idx = []
    for k in range(n):
        idx += [0, 1 + k, 2 + k]
    out = oracle.geometry_pass([make_draw(v, idx)], view, proj, 64, 48)
    assert (out["prim"] != EMPTY).all(), "gap between triangles sharing an edge"
    assert np.abs(out["depth"] - 0.5).max() < 1e-6  # l0 + l1 + l2 is 1 only up to rounding
    # the opposite winding is back-facing (front = clockwise with y down): nothing is drawn
    idx2 = []
    for k in range(n):
        idx2 += [0, 2 + k, 1 + k]
    out2 = oracle.geometry_pass([make_draw(v, idx2)], view, proj, 64, 48)
    assert (out2["prim"] == EMPTY).all()
    # both windings in one draw: each pixel still belongs to exactly one front-facing triangle, the same as before
    out3 = oracle.geometry_pass([make_draw(v, idx2 + idx)], view, proj, 64, 48)
    assert np.array_equal(out3["prim"], out["prim"] + n)


def test_pixel_centre_sampling_and_top_left_rule(oracle):
    """An axis-aligned quad whose edges pass exactly through pixel centres: left/top edges are in, right/bottom out."""
    view, proj = ortho_like()
    W = H = 8
    # framebuffer x = (ndc+1)*W/2: corners at pixel centres 1.5 and 5.5
    lo, hi = 1.5 / (W / 2) - 1, 5.5 / (W / 2) - 1
    v = [mesh_vertex((lo, lo, 0.25)), mesh_vertex((hi, lo, 0.25)), mesh_vertex((hi, hi, 0.25)), mesh_vertex((lo, hi, 0.25))]
    out = oracle.geometry_pass([make_draw(v, [0, 1, 2, 0, 2, 3])], view, proj, W, H)
    cov = out["prim"] != EMPTY
    want = np.zeros((H, W), bool)
    want[1:5, 1:5] = True  # centres 1.5..4.5 inside; 5.5 lies on the right / bottom edge -> excluded
    assert np.array_equal(cov, want)


def test_depth_less_and_draw_order_ties(oracle):
    view, proj = ortho_like()
    near_q, far_q = quad(0.25, 0.5), quad(0.75, 1.0)
    a = make_draw(*near_q, color=(1, 0, 0, 1))
    b = make_draw(*far_q, color=(0, 1, 0, 1))
    o1, o2 = oracle.geometry_pass([a, b], view, proj, 32, 32), oracle.geometry_pass([b, a], view, proj, 32, 32)
    assert np.array_equal(o1["depth"], o2["depth"]) and np.array_equal(o1["alb"], o2["alb"])
    assert np.abs(o1["depth"][12:20, 12:20] - 0.25).max() < 1e-6 and np.abs(o1["depth"][0, :] - 0.75).max() < 1e-6
    # equal depth: LESS keeps the first triangle drawn
    c = make_draw(*quad(0.25, 0.5), color=(0, 0, 1, 1))
    o3, o4 = oracle.geometry_pass([a, c], view, proj, 32, 32), oracle.geometry_pass([c, a], view, proj, 32, 32)
    assert (o3["alb"][16, 16] == [65535, 0, 0, 65535]).all() and (o4["alb"][16, 16] == [0, 0, 65535, 65535]).all()
    # z_fb = 1.0 fails LESS against the clear value; z outside [0, 1) is clipped
    for z in (1.0, 1.5, -0.25):
        assert (oracle.geometry_pass([make_draw(*quad(z, 0.5))], view, proj, 16, 16)["prim"] == EMPTY).all()


def test_near_plane_clipping(oracle):
    """A ground quad that runs from far in front of the camera to behind it: only the part in front of the near plane
    is drawn, nothing is NaN, and depth is monotone towards the camera."""
    view, proj = camera((0, -1, -10), (0, -1, 0), aspect=1.0)
    g = 50.0
    v = [mesh_vertex((-g, 1.0, -g), (0, -1, 0), (1, 0, 0, 1), (0, 0)), mesh_vertex((g, 1.0, -g), (0, -1, 0), (1, 0, 0, 1), (4, 0)),
         mesh_vertex((g, 1.0, g), (0, -1, 0), (1, 0, 0, 1), (4, 4)), mesh_vertex((-g, 1.0, g), (0, -1, 0), (1, 0, 0, 1), (0, 4))]
    out = None
    for idx in ([0, 1, 2, 0, 2, 3], [0, 2, 1, 0, 3, 2]):
        o = oracle.geometry_pass([make_draw(v, idx, color_tex=checker_texture(8, 8, 5))], view, proj, 64, 64)
        if (o["prim"] != EMPTY).any():
            out = o
    assert out is not None
    cov = out["prim"] != EMPTY
    assert 0.3 < cov.mean() < 0.7 and np.isfinite(out["depth"]).all()
    ys = np.where(cov.any(axis=1))[0]
    assert len(ys) in (31, 32, 33)  # the plane y = +1 fills the half of the screen on one side of the horizon
    col = out["depth"][ys, 32]
    assert (np.diff(col) != 0).all() and (np.all(np.diff(col) > 0) or np.all(np.diff(col) < 0))
    pos = out["pos"].view(np.float16).astype(np.float32)[cov]
    assert np.allclose(pos[:, 1], 1.0, atol=2e-3)  # interpolated world positions stay on the plane y = 1


def test_guard_band_clipping_of_huge_triangles(oracle):
    view, proj = ortho_like()
    v = [mesh_vertex((-5000.0, -4000.0, 0.5)), mesh_vertex((6000.0, -3000.0, 0.5)), mesh_vertex((100.0, 7000.0, 0.5))]
    out = oracle.geometry_pass([make_draw(v, [0, 1, 2])], view, proj, 40, 30)
    out2 = oracle.geometry_pass([make_draw(v, [0, 2, 1])], view, proj, 40, 30)
    cov = (out["prim"] != EMPTY) | (out2["prim"] != EMPTY)
    assert cov.all()


def test_golden_frame(oracle):
    """The committed golden frame pins the oracle itself against silent changes."""
    view, proj = camera()
    out = oracle.geometry_pass(raster_scene(1), view, proj, 160, 90, extended=True)
    if not GOLDEN.exists():
        pytest.fail("tests/golden/raster_golden.npz is missing: run tests/golden/make_raster_golden.py")
    gold = np.load(GOLDEN)
    for k in ("pos", "nrm", "alb", "depth", "prim", "orm", "emi"):
        assert np.array_equal(out[k], gold[k]), k
