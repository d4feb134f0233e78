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
