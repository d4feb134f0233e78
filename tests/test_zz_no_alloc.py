"""GPU: pbrcu_shade_ggx_ibl_batched_static, the no-allocation variant of the batched K3 launch (its per-view camera table
lives in a file-scope device array instead of context scratch grown with cudaMalloc).  No counterpart in the reference.
Checked through the C ABI against the existing entry point
(pbrcu_shade_ggx_ibl_batched) on two calls in a row -- the second with other cameras and fewer views, so stale table
entries of the first call would show -- and against the limits it documents.  Named test_zz_* so pytest collects it last."""
import ctypes as C

import numpy as np
import pytest
from oracle import synth_gbuffer, synth_lights

pytestmark = pytest.mark.gpu


def _views(dev, pbr, w, h, n, seed):
    gs = [synth_gbuffer(w, h, seed=seed + v, tier_b=True) for v in range(n)]
    stack = lambda k, dt: np.ascontiguousarray(np.stack([np.asarray(g[k]).view(dt) for g in gs]))
    planes = dict(pos=stack("pos", np.uint16), nrm=stack("nrm", np.uint16), alb=stack("alb", np.uint16), orm=stack("orm", np.uint8),
                  emi=stack("emi", np.uint16))
    bufs = {k: dev.upload(v) for k, v in planes.items()}
    g0 = pbr.GBuffer()
    g0.width, g0.height, g0.pitch8, g0.pitch4 = w, h, w * 8, w * 4
    g0.positions, g0.normals, g0.albedo, g0.orm, g0.emissive = bufs["pos"].ptr, bufs["nrm"].ptr, bufs["alb"].ptr, bufs["orm"].ptr, bufs["emi"].ptr
    return g0, bufs


def test_static_scratch_entry_point_equals_the_allocating_one(pbr, dev):
    w, h, n = 80, 45, 5
    g0, _keep = _views(dev, pbr, w, h, n, seed=900)
    lights = synth_lights(5, seed=9)
    ld = dev.upload(lights)
    stride8, stride4 = w * h * 8, w * h * 4
    out_a, out_b = dev.alloc(n * stride8), dev.alloc(n * stride8)
    rng = np.random.default_rng(3)
    for call, nv in enumerate((n, n - 2)):  # two calls in a row; the second re-uses the shared table with other contents
        cams = np.ascontiguousarray(rng.uniform(-12, 12, (nv, 3)).astype(np.float32))
        cam_ptr = cams.ctypes.data_as(C.POINTER(C.c_float))
        for buf in (out_a, out_b):
            dev.lib.memset(buf.ptr, 0x5A, n * stride8, None)
        dev.lib.shade_ggx_ibl_batched(dev.ctx, C.byref(g0), stride8, stride4, nv, cam_ptr, ld.ptr, 5, None, out_a.ptr, w * 8, stride8, None)
        dev.lib.shade_ggx_ibl_batched_static(dev.ctx, C.byref(g0), stride8, stride4, nv, cam_ptr, ld.ptr, 5, None, out_b.ptr, w * 8, stride8, None)
        dev.sync()
        want = out_a.download((n, h, w, 4), np.uint16)
        got = out_b.download((n, h, w, 4), np.uint16)
        assert np.isfinite(got[:nv].view(np.float16).astype(np.float32)).all()
        assert (want[:nv] != 0x5A5A).any(), "the allocating entry point wrote nothing"
        # same kernel, same inputs: the tolerance of the batched tests is bit equality
        assert np.array_equal(got, want), f"call {call}: {np.count_nonzero(got != want)} values differ"


def test_static_scratch_entry_point_states_its_limits(pbr, dev):
    w, h = 16, 8
    g0, _keep = _views(dev, pbr, w, h, 1, seed=950)
    out = dev.alloc(w * h * 8)
    lights = synth_lights(2, seed=1)
    ld = dev.upload(lights)
    too_many = 4096 + 1  # PBRCU_STATIC_MAX_VIEWS + 1
    cams = np.zeros((too_many, 3), np.float32)
    with pytest.raises(pbr.PbrcuError) as e:
        dev.lib.shade_ggx_ibl_batched_static(dev.ctx, C.byref(g0), w * h * 8, w * h * 4, too_many, cams.ctypes.data_as(C.POINTER(C.c_float)), ld.ptr, 2, None,
                                             out.ptr, w * 8, w * h * 8, None)
    assert "PBRCU_STATIC_MAX_VIEWS" in str(e.value)
    with pytest.raises(pbr.PbrcuError) as e:  # a host light list would need an upload buffer
        dev.lib.shade_ggx_ibl_batched_static(dev.ctx, C.byref(g0), w * h * 8, w * h * 4, 1, cams.ctypes.data_as(C.POINTER(C.c_float)), lights.ctypes.data, 2, None,
                                             out.ptr, w * 8, w * h * 8, None)
    assert "DEVICE" in str(e.value)
