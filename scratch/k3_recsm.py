"""K3 on the bench's 4K frame: the committed kernel against the variants that send the two prefiltered-cube records through shared memory
(cp.async) instead of registers, at 5 / 6 / 7 CTAs per SM (PBRCU_K3_EXP = 11 / 12 / 13; 14 = 6 CTAs with the late filter)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import bench, pbr_b200 as pbr
dev = pbr.Device(0)
scene = bench.Scene(pbr, dev, torch)
lib, ctx = dev.lib, dev.ctx
def shade(i):
    g, _ = scene.gbufs[i % len(scene.gbufs)]
    lib.shade_ggx_ibl(ctx, C.byref(g), pbr.f3(*bench.CAM), scene.lights_host.ctypes.data, bench.N_LIGHTS, None, C.byref(scene.ibl), pbr.WHOLE,
                      scene.outs[i % len(scene.outs)].data_ptr(), bench.W4K * 8, scene.stream())
def run(reps=200):
    shade(0); torch.cuda.synchronize()
    out = scene.outs[0].clone()
    for i in range(20): shade(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps): shade(i)
    b.record(); torch.cuda.synchronize()
    return out, a.elapsed_time(b) / reps * 1e3
ref = None
for rnd in range(2):
    for v in (None, 11, 12, 13, 14):
        if v is None: os.environ.pop("PBRCU_K3_EXP", None)
        else: os.environ["PBRCU_K3_EXP"] = str(v)
        out, us = run()
        if ref is None: ref = out
        print(f"variant {v}: {us:.1f} us per 4K frame, {bench.W4K * bench.H4K / us / 1e3:.2f} Gpix/s, equal to the committed kernel: {bool(torch.equal(out, ref))}", flush=True)
