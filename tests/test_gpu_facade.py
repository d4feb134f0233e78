"""GPU: one frame through the C++ facade (pbr::PbrRenderSystem::render -> HdrImage::updateOutputTexture ->
pbr::TonemapperSystem::run, the reference's App.cpp:387-392 sequence) vs the oracle."""
import subprocess
from pathlib import Path

import numpy as np
import pytest
from conftest import assert_hdr_parity, f16
from oracle import synth_environment, synth_gbuffer, synth_lights

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
CPP = ROOT / "tests" / "cpp"


@pytest.fixture(scope="module")
def exe():
    subprocess.run(["make", "-C", str(CPP)], check=True, capture_output=True)
    return str(CPP / "build" / "facade_frame")


def dump(g, d, tier_b):
    g["pos"].tofile(d / "pos.bin"); g["nrm"].tofile(d / "nrm.bin"); g["alb"].tofile(d / "alb.bin")
    if tier_b:
        g["orm"].tofile(d / "orm.bin"); g["emi"].tofile(d / "emi.bin")


@pytest.mark.parametrize("w,h", [(64, 48), (333, 17)])
def test_facade_blinnphong_frame(exe, oracle, tmp_path, w, h):
    g = synth_gbuffer(w, h, seed=5)
    dump(g, tmp_path, False)
    cam = (0.5, -1.0, 12.0)
    r = subprocess.run([exe, "blinnphong", str(w), str(h), *map(str, cam), str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    hdr = np.fromfile(tmp_path / "hdr.bin", np.uint16).reshape(h, w, 4)
    ldr = np.fromfile(tmp_path / "ldr.bin", np.uint8).reshape(h, w, 4)
    _, want = oracle.lighting(g["pos"], g["nrm"], g["alb"], cam)
    assert np.abs(f16(hdr) - f16(want)).max() <= 1e-3
    assert np.abs(ldr.astype(int) - oracle.tonemap(hdr).astype(int)).max() <= 1  # tonemap of the device's own HDR


def test_facade_ggx_frame_with_environment(exe, oracle, tmp_path):
    w, h, nl = 96, 64, 4
    g = synth_gbuffer(w, h, seed=11, tier_b=True)
    dump(g, tmp_path, True)
    lights = synth_lights(nl, seed=4)
    lights.tofile(tmp_path / "lights.bin")
    env = synth_environment(256, 128, 1)
    env.tofile(tmp_path / "env.bin")
    res, mips, ns, ires, lres = 32, 6, 256, 16, 32
    cam = (0.5, -1.0, 12.0)
    r = subprocess.run([exe, "ggx", str(w), str(h), *map(str, cam), str(tmp_path), "256", "128", str(res), str(mips), str(ns), str(ires), str(lres)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    hdr = np.fromfile(tmp_path / "hdr.bin", np.uint16).reshape(h, w, 4)
    chain = oracle.make_chain(oracle.equirect_to_cube(env, res), mips)
    ibl = dict(irr=oracle.irradiance(chain, res, mips, ires, ns), pref=oracle.prefilter(chain, res, mips, ns), pref_res=res, pref_mips=mips,
               lut=oracle.brdf_lut(lres, ns))
    _, want = oracle.shade(g["pos"], g["nrm"], g["alb"], g["orm"], g["emi"], cam, lights, ibl)
    # the whole chain (equirect -> cube -> prefilter -> shade) ran on the device: fp16 roundings of the
    # intermediate textures can differ by an ulp from the oracle's, so allow two fp16 ulps on top
    assert_hdr_parity(hdr, want, peak=float(max(f16(want).max(), 1.0)), rel_tol=2.0 ** -8, what="facade ggx frame")


def test_facade_mesh_scene_geometry_and_lighting(oracle, tmp_path):
    """A mesh scene built and drawn the way the reference's client does (MeshBuilder -> TransferStager -> Mesh /
    Material -> Node -> Scene, then PbrRenderSystem::render = geometry pass + lighting pass, TonemapperSystem::run):
    the G-buffer must equal the oracle's geometry pass bit for bit on the matrices the facade computed, and the HDR /
    LDR frames the oracle's lighting and tonemap of that G-buffer."""
    from oracle import checker_texture, make_draw, mesh_vertex, uv_sphere
    subprocess.run(["make", "-C", str(CPP)], check=True, capture_output=True)
    exe = str(CPP / "build" / "facade_scene")
    W, H = 200, 120
    eye, target, fov = (1.0, -2.0, -9.0), (0.0, 0.0, 0.0), 1.2
    sv, si = uv_sphere(1.0, 10, 20)
    g = 6.0
    quad_v = np.array([mesh_vertex((-g, -g, 3.0), (0, 0, -1), (1, 0, 0, 1), (0, 0)), mesh_vertex((g, -g, 3.0), (0, 0, -1), (1, 0, 0, 1), (3, 0)),
                       mesh_vertex((g, g, 3.0), (0, 0, -1), (1, 0, 0, 1), (3, 3)), mesh_vertex((-g, g, 3.0), (0, 0, -1), (1, 0, 0, 1), (0, 3))], np.float32)
    quad_i = np.array([0, 1, 2, 0, 2, 3, 0, 2, 1, 0, 3, 2], np.uint16)
    # mesh 0: two primitives (sphere + quad) under one transform; mesh 1: a scaled, rotated sphere
    meshes = [
        dict(pos=(0.5, 0.2, 0.0), quat=(1.0, 0.0, 0.0, 0.0), scale=(1.0, 1.0, 1.0),
             prims=[dict(v=sv, i=si, color=(0.9, 0.6, 0.3, 1.0), ctex=checker_texture(16, 8, 21), ntex=checker_texture(8, 8, 22, normal_map=True)),
                    dict(v=quad_v, i=quad_i, color=(0.5, 0.7, 0.9, 1.0), ctex=checker_texture(32, 32, 23), ntex=np.array([[[128, 128, 255, 255]]], np.uint8))]),
        dict(pos=(-2.0, -1.0, -1.5), quat=(0.9238795, 0.0, 0.3826834, 0.0), scale=(1.5, 0.7, 1.0),
             prims=[dict(v=sv, i=si, color=(0.2, 0.9, 0.4, 0.5), ctex=checker_texture(8, 4, 24), ntex=checker_texture(4, 4, 25, normal_map=True))]),
    ]
    lines = [f"{W} {H} {' '.join(map(str, eye))} {' '.join(map(str, target))} {fov} {len(meshes)}"]
    for mi, m in enumerate(meshes):
        lines.append(f"{len(m['prims'])} {' '.join(map(str, m['pos']))} {' '.join(map(str, m['quat']))} {' '.join(map(str, m['scale']))}")
        for pi, p in enumerate(m["prims"]):
            prefix = f"m{mi}p{pi}"
            p["v"].astype(np.float32).tofile(tmp_path / f"{prefix}_v.bin"); p["i"].astype(np.uint16).tofile(tmp_path / f"{prefix}_i.bin")
            p["ctex"].tofile(tmp_path / f"{prefix}_c.bin"); p["ntex"].tofile(tmp_path / f"{prefix}_n.bin")
            lines.append(f"{prefix} {' '.join(map(str, p['color']))} {p['ctex'].shape[1]} {p['ctex'].shape[0]} {p['ntex'].shape[1]} {p['ntex'].shape[0]}")
    (tmp_path / "scene.txt").write_text("\n".join(lines) + "\n")
    r = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    mats = np.fromfile(tmp_path / "matrices.bin", np.float32)
    view, proj, per_draw = mats[:16], mats[16:32], mats[32:].reshape(-1, 28)
    # the draws in the order render() issues them: MeshBuilder concatenates the primitives of a mesh (firstVertex / firstIndex)
    draws, k = [], 0
    for m in meshes:
        verts = np.concatenate([p["v"].reshape(-1, 12) for p in m["prims"]])
        idx = np.concatenate([p["i"] for p in m["prims"]])
        fv = fi = 0
        for p in m["prims"]:
            draws.append(make_draw(verts, idx, (per_draw[k, :16], per_draw[k, 16:]), color=p["color"], color_tex=p["ctex"], normal_tex=p["ntex"],
                                   first_index=fi, index_count=len(p["i"]), vertex_offset=fv))
            fv += len(p["v"].reshape(-1, 12)); fi += len(p["i"]); k += 1
    want = oracle.geometry_pass(draws, view, proj, W, H)
    assert 0.3 < (want["prim"] != 0xFFFFFFFF).mean() < 1.0
    got = dict(pos=np.fromfile(tmp_path / "pos.bin", np.uint16).reshape(H, W, 4), nrm=np.fromfile(tmp_path / "nrm.bin", np.uint16).reshape(H, W, 4),
               alb=np.fromfile(tmp_path / "alb.bin", np.uint16).reshape(H, W, 4), depth=np.fromfile(tmp_path / "depth.bin", np.float32).reshape(H, W))
    for name in ("pos", "nrm", "alb"):
        assert np.array_equal(got[name], want[name]), f"G-buffer plane {name} differs in {np.count_nonzero(got[name] != want[name])} values"
    assert np.array_equal(got["depth"].view(np.uint32), want["depth"].view(np.uint32))
    hdr = np.fromfile(tmp_path / "hdr.bin", np.uint16).reshape(H, W, 4)
    ldr = np.fromfile(tmp_path / "ldr.bin", np.uint8).reshape(H, W, 4)
    _, want_hdr = oracle.lighting(want["pos"].view(np.float16), want["nrm"].view(np.float16), want["alb"], eye)
    assert np.abs(f16(hdr) - f16(want_hdr)).max() <= 1e-3
    assert np.abs(ldr.astype(int) - oracle.tonemap(hdr).astype(int)).max() <= 1


def test_device_group_frame_from_cpp():
    """tests/cpp/mg_frame.cpp: a frame over pbr::DeviceGroup (the C multi-GPU host behind the facade) -- two and three members on GPU 0,
    and all GPUs with the NCCL gather when the box has several -- equals the single-device frame bit for bit."""
    import os
    subprocess.run(["make", "-C", str(CPP)], check=True, capture_output=True)
    env = dict(os.environ)
    r = subprocess.run([str(CPP / "build" / "mg_frame")], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0 and "mg_frame ok" in r.stdout, r.stdout + r.stderr

