"""GPU: the glTF import path and the headless viewer end to end.
A GLB written by tests/gltfwriter.py (PNG textures embedded, interleaved + separate vertex buffers, u8 / u16 indices,
a node hierarchy) goes through tools/gltf_viewer: gltf::Loader -> Asset::loadScene -> TransferStager -> per frame
CameraController / node spin / PbrRenderSystem::render / TonemapperSystem::run.  The last frame's G-buffer must equal
the oracle's geometry pass BIT FOR BIT on the matrices the viewer computed, the HDR / LDR frames the oracle's lighting and
tonemap, and the PNG it writes must decode to that LDR frame."""
import json
import subprocess
from pathlib import Path

import numpy as np
import pytest
from conftest import f16
from gltfwriter import UBYTE, Builder
from oracle import checker_texture, make_draw, mesh_vertex, uv_sphere

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
VIEWER = ROOT / "tools" / "gltf_viewer"


@pytest.fixture(scope="module")
def viewer():
    subprocess.run(["make", "-C", str(VIEWER)], check=True, capture_output=True)
    subprocess.run(["make", "-s", "-C", str(ROOT / "tests" / "cpp"), "cpu"], check=True)
    return str(VIEWER / "build" / "gltf_viewer")


def asset(tmp_path):
    """Returns the path of the GLB and the draws (vertex data, textures, colours) in the order render() issues them:
    Scene::iterateAllNodes visits children before their parent (Scene.hpp), one draw per primitive."""
    b = Builder()
    sv, si = uv_sphere(1.0, 10, 20)
    sv = sv.reshape(-1, 12).astype(np.float32)
    g = 6.0
    quad = np.array([mesh_vertex((-g, -g, 3.0), (0, 0, -1), (1, 0, 0, 1), (0, 0)), mesh_vertex((g, -g, 3.0), (0, 0, -1), (1, 0, 0, 1), (3, 0)),
                     mesh_vertex((g, g, 3.0), (0, 0, -1), (1, 0, 0, 1), (3, 3)), mesh_vertex((-g, g, 3.0), (0, 0, -1), (1, 0, 0, 1), (0, 3))], np.float32)
    quad_i = np.array([0, 1, 2, 0, 2, 3, 0, 2, 1, 0, 3, 2], np.uint16)
    tex = dict(c0=checker_texture(16, 8, 21), n0=checker_texture(8, 8, 22, normal_map=True), c1=checker_texture(32, 32, 23),
               n1=np.array([[[128, 128, 255, 255]]], np.uint8), c2=checker_texture(8, 4, 24), n2=checker_texture(4, 4, 25, normal_map=True))
    # c2 travels as a progressive 4:2:0 JPEG: what the material sees is whatever the import path's decoder makes of it
    # (that decoder is pinned bit-exact to the reference's stb_image in tests/test_image_decode_cpu.py)
    import io
    from PIL import Image
    buf = io.BytesIO()
    Image.fromarray(np.kron(tex["c2"][..., :3], np.ones((4, 4, 1), np.uint8))).save(buf, "JPEG", quality=80, progressive=True, subsampling=2)
    (tmp_path / "c2.jpg").write_bytes(buf.getvalue())
    dec = subprocess.run([str(ROOT / "tests" / "cpp" / "build" / "image_decode"), str(tmp_path / "c2.jpg"), str(tmp_path / "c2.rgba")], capture_output=True, text=True)
    assert dec.returncode == 0, dec.stderr
    jw, jh = map(int, dec.stdout.split())
    tex["c2"] = np.fromfile(tmp_path / "c2.rgba", np.uint8).reshape(jh, jw, 4)
    t = {k: (b.image(None, k, encoded=buf.getvalue(), mime="image/jpeg") if k == "c2" else b.image(v, k, how="view", filters=(4, 2), level=6, interlace=(k == "c1")))
         for k, v in tex.items()}
    m0 = b.material("m0", (0.9, 0.6, 0.3, 1.0), t["c0"], t["n0"])
    m1 = b.material("m1", (0.5, 0.7, 0.9, 1.0), t["c1"], t["n1"])
    m2 = b.material("m2", (0.2, 0.9, 0.4, 0.5), t["c2"], t["n2"])
    two = b.mesh("two", [b.primitive(sv, si, m0, layout="interleaved"), b.primitive(quad, quad_i, m1, index_type=UBYTE)])
    one = b.mesh("one", [b.primitive(sv, si, m2)])
    child = b.node("child", mesh=one, translation=(-2.0, -1.0, -1.5), rotation_xyzw=(0.0, 0.3826834, 0.0, 0.9238795), scale=(1.5, 0.7, 1.0))
    parent = b.node("parent", mesh=two, translation=(0.5, 0.2, 0.0), children=[child])
    b.scene([parent])
    path = tmp_path / "scene.glb"
    b.write_glb(path)
    both_v, both_i = np.concatenate([sv, quad]), np.concatenate([np.asarray(si, np.uint16), quad_i])
    draws = [dict(v=sv, i=np.asarray(si, np.uint16), fi=0, n=len(si), vo=0, color=(0.2, 0.9, 0.4, 0.5), ct=tex["c2"], nt=tex["n2"]),     # child first
             dict(v=both_v, i=both_i, fi=0, n=len(si), vo=0, color=(0.9, 0.6, 0.3, 1.0), ct=tex["c0"], nt=tex["n0"]),
             dict(v=both_v, i=both_i, fi=len(si), n=len(quad_i), vo=len(sv), color=(0.5, 0.7, 0.9, 1.0), ct=tex["c1"], nt=tex["n1"])]
    return path, draws


def check_frame(oracle, d, draws, W, H, prefix="", ldr_frame=True):
    class _Prefixed:  # d / "name" -> d / (prefix + "name")
        def __truediv__(self, name):
            return d_plain / (prefix + name)
    d_plain, d = d, _Prefixed()
    mats = np.fromfile(d / "matrices.bin", np.float32)
    view, proj, eye, per_draw = mats[:16], mats[16:32], mats[32:35], mats[36:].reshape(-1, 28)
    assert len(per_draw) == len(draws)
    ds = [make_draw(x["v"], x["i"], (per_draw[k, :16], per_draw[k, 16:]), color=x["color"], color_tex=x["ct"], normal_tex=x["nt"], first_index=x["fi"], index_count=x["n"],
                    vertex_offset=x["vo"]) for k, x in enumerate(draws)]
    want = oracle.geometry_pass(ds, view, proj, W, H)
    got = dict(pos=np.fromfile(d / "pos.bin", np.uint16).reshape(H, W, 4), nrm=np.fromfile(d / "nrm.bin", np.uint16).reshape(H, W, 4),
               alb=np.fromfile(d / "alb.bin", np.uint16).reshape(H, W, 4), depth=np.fromfile(d / "depth.bin", np.float32).reshape(H, W))
    covered = (want["prim"] != 0xFFFFFFFF).mean()
    for name in ("pos", "nrm", "alb"):
        assert np.array_equal(got[name], want[name]), f"G-buffer plane {name} differs in {np.count_nonzero(got[name] != want[name])} values"
    assert np.array_equal(got["depth"].view(np.uint32), want["depth"].view(np.uint32))
    hdr = np.fromfile(d / "hdr.bin", np.uint16).reshape(H, W, 4)
    _, want_hdr = oracle.lighting(want["pos"].view(np.float16), want["nrm"].view(np.float16), want["alb"], tuple(eye))
    assert np.abs(f16(hdr) - f16(want_hdr)).max() <= 1e-3
    if not ldr_frame:
        return covered, eye, per_draw, None
    ldr = np.fromfile(d / "ldr.bin", np.uint8).reshape(H, W, 4)
    assert np.abs(ldr.astype(int) - oracle.tonemap(hdr).astype(int)).max() <= 1
    return covered, eye, per_draw, ldr


@pytest.mark.parametrize("W,H", [(200, 120), (1024, 1024)], ids=["200x120", "baseline-config0-1024x1024"])
def test_viewer_single_frame_matches_oracle(viewer, oracle, tmp_path, W, H):
    """One headless frame of a glTF through the viewer; 1024x1024 is BASELINE configs[0] (the lavapipe run it names is
    unavailable on this box: the checker is the oracle's geometry pass + the reference-pinned lighting / tonemap)."""
    path, draws = asset(tmp_path)
    r = subprocess.run([viewer, str(path), "-vulkan-validation", "--size", f"{W}x{H}", "--frames", "1", "--dt", "0", "--no-spin", "--position", "1,-2,-9", "--yaw", "1.6",
                        "--pitch", "0.2", "--fov", "1.2", "--dump", str(tmp_path), "--out", str(tmp_path / "frame.png"), "--hdr", str(tmp_path / "frame.pfm")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Frame time" in r.stdout
    covered, eye, _, ldr = check_frame(oracle, tmp_path, draws, W, H)
    assert 0.3 < covered < 1.0 and np.allclose(eye, (1, -2, -9))
    # the PNG the viewer wrote decodes (with the import path's own decoder) to the LDR frame
    dec = subprocess.run([str(ROOT / "tests" / "cpp" / "build" / "image_decode"), str(tmp_path / "frame.png"), str(tmp_path / "frame.rgba")], capture_output=True, text=True)
    assert dec.returncode == 0 and dec.stdout.split() == [str(W), str(H)]
    assert np.array_equal(np.fromfile(tmp_path / "frame.rgba", np.uint8).reshape(H, W, 4), ldr)
    pfm = (tmp_path / "frame.pfm").read_bytes()
    head = f"PF\n{W} {H}\n-1.0\n".encode()
    assert pfm.startswith(head) and len(pfm) == len(head) + W * H * 12


def test_viewer_frame_loop_moves_camera_and_spins_nodes(viewer, oracle, tmp_path):
    """Five frames with W held, a mouse drag and the default node spin (App.cpp:285-318): the last frame still
    matches the oracle on the matrices of that frame, and those matrices moved the way CameraController says."""
    path, draws = asset(tmp_path)
    W, H = 160, 90
    r = subprocess.run([viewer, str(path), "--size", f"{W}x{H}", "--frames", "5", "--dt", "0.05", "--keys", "W", "--look", "3,1", "--position", "1,-2,-9", "--yaw", "1.6",
                        "--dump", str(tmp_path), "--times", str(tmp_path / "times.json")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    _, eye, per_draw, _ = check_frame(oracle, tmp_path, draws, W, H)
    assert np.linalg.norm(eye - np.array([1, -2, -9], np.float32)) > 1e-3  # W held: the camera moved forward
    # spin: rotation about +Y by dt per frame accumulates in every mesh node's model matrix (parent had identity rotation)
    parent_model = per_draw[1, :16].reshape(4, 4).T
    ang = 5 * 0.05
    assert np.allclose(parent_model[:3, :3], [[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]], atol=1e-5)
    times = json.loads((tmp_path / "times.json").read_text())
    assert times["frames"] == 5 and len(times["frame_time_ms"]) == 5 and all(t > 0 for t in times["frame_time_ms"])


def test_viewer_views_mode_matches_oracle(viewer, oracle, tmp_path):
    """--views: a block of orbit cameras of the glTF scene through PbrRenderSystem::renderViews (one batched geometry pass, one batched
    lighting pass; BASELINE configs[4]).  First, middle and last view: G-buffer planes BIT FOR BIT the oracle's geometry pass on the
    matrices the viewer computed, frames within the lighting tolerance; the views differ from each other; a shard of the same orbit
    (--view-first / --view-total) reproduces the corresponding views of the whole one byte for byte."""
    path, draws = asset(tmp_path)
    W, H, N = 200, 120, 7
    d = tmp_path / "views"; d.mkdir()
    r = subprocess.run([viewer, str(path), "--size", f"{W}x{H}", "--views", str(N), "--frames", "1", "--orbit", "9,-2", "--dump", str(d), "--times", str(tmp_path / "v.json")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    t = json.loads((tmp_path / "v.json").read_text())
    assert t["views"] == N and t["draws"] == len(draws) and t["triangles_per_view"] == sum(x["n"] // 3 for x in draws)
    assert 0 < t["ms_geometry"] <= t["ms_geometry_plus_shading"] * 1.5
    eyes = []
    for v in (0, N // 2, N - 1):
        covered, eye, _, _ = check_frame(oracle, d, draws, W, H, prefix=f"view{v}_", ldr_frame=False)
        assert covered > 0.02
        eyes.append(tuple(eye))
    assert len(set(eyes)) == 3
    # cameras 3..6 of the same orbit as a shard of their own
    d2 = tmp_path / "shard"; d2.mkdir()
    r = subprocess.run([viewer, str(path), "--size", f"{W}x{H}", "--views", "4", "--view-first", "3", "--view-total", str(N), "--frames", "1", "--orbit", "9,-2",
                        "--dump", str(d2)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    for name in ("pos", "nrm", "alb", "depth", "hdr", "matrices"):
        assert (d2 / f"view0_{name}.bin").read_bytes() == (d / f"view{N // 2}_{name}.bin").read_bytes(), name
        assert (d2 / f"view3_{name}.bin").read_bytes() == (d / f"view{N - 1}_{name}.bin").read_bytes(), name


def test_viewer_ggx_mode_runs(viewer, tmp_path):
    path, _ = asset(tmp_path)
    r = subprocess.run([viewer, str(path), "--size", "128x72", "--frames", "2", "--dt", "0.016", "--ggx", "--position", "1,-2,-9", "--yaw", "1.6", "--out", str(tmp_path / "ggx.png")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    dec = subprocess.run([str(ROOT / "tests" / "cpp" / "build" / "image_decode"), str(tmp_path / "ggx.png"), str(tmp_path / "ggx.rgba")], capture_output=True, text=True)
    assert dec.returncode == 0
    img = np.fromfile(tmp_path / "ggx.rgba", np.uint8).reshape(72, 128, 4)
    assert img[..., :3].std() > 5  # a lit, textured frame rather than a constant


def test_viewer_without_arguments_and_with_a_bad_asset(viewer, tmp_path):
    assert subprocess.run([viewer]).returncode == 0  # Main.cpp:18: no arguments, nothing to do
    (tmp_path / "bad.glb").write_bytes(b"glTF\x02\0\0\0\x10\0\0\0abcd")
    r = subprocess.run([viewer, str(tmp_path / "bad.glb")], capture_output=True, text=True)
    assert r.returncode == 1 and "Cant load gltf asset" in r.stderr
