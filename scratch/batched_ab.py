"""256 views x 1080p through pbrcu_shade_ggx_ibl_batched: host light list (shade_ggx_cl_kernel<BATCHED>) vs PBRCU_K3_CL=0 (shade_ggx_one_kernel<BATCHED>)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import bench, pbr_b200 as pbr
dev = pbr.Device(0)
scene = bench.Scene(pbr, dev, torch)
lib, ctx = dev.lib, dev.ctx
WV, HV, NV = 1920, 1080, 256
s8, s4 = WV * HV * 8, WV * HV * 4
planes = {k: scene.buf(NV * (s4 if k == "orm" else s8)) for k in ("positions", "normals", "albedo", "orm", "emissive")}
outs = scene.buf(NV * s8)
g0 = pbr.GBuffer(); g0.width, g0.height, g0.pitch8, g0.pitch4 = WV, HV, WV * 8, WV * 4
for k, t in planes.items(): setattr(g0, k, t.data_ptr())
cams = np.zeros((NV, 3), np.float32)
for v in range(NV):
    ring, k = divmod(v, NV // 2); ang = 2 * np.pi * k / (NV // 2)
    cams[v] = (12 * np.cos(ang), -1 - 4 * ring, 12 * np.sin(ang))
    gv = pbr.GBuffer(); gv.width, gv.height, gv.pitch8, gv.pitch4 = WV, HV, WV * 8, WV * 4
    for k2, t in planes.items(): setattr(gv, k2, t.data_ptr() + v * (s4 if k2 == "orm" else s8))
    lib.synth_gbuffer(ctx, C.byref(gv), 1, bench.SEED + 50, pbr.f3(*cams[v]), pbr.f3(0, 0, 0), None)
torch.cuda.synchronize()
cp = cams.ctypes.data_as(C.POINTER(C.c_float))
def run(ptr):
    f = lambda: lib.shade_ggx_ibl_batched(ctx, C.byref(g0), s8, s4, NV, cp, ptr, 4, C.byref(scene.ibl), outs.data_ptr(), WV * 8, s8, scene.stream())
    for _ in range(2): f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(4): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / 4
for rep in range(2):
    os.environ.pop("PBRCU_K3_CL", None)
    os.environ["PBRCU_K3_BRANCH"] = "1"
    print("host list, cl kernel BRANCH:", round(run(scene.lights_host.ctypes.data), 3), "ms", flush=True)
    os.environ["PBRCU_K3_BRANCH"] = "0"
    print("host list, cl kernel      :", round(run(scene.lights_host.ctypes.data), 3), "ms", flush=True)
    print("device list, one_kernel   :", round(run(scene.lights.data_ptr()), 3), "ms", flush=True)
    os.environ["PBRCU_K3_CL"] = "0"
    print("host list, CL=0 (uploaded):", round(run(scene.lights_host.ctypes.data), 3), "ms", flush=True)
