"""K3 on the bench's 4K frame with the per-pixel-material G-buffer (kind 1, the headline input) and with per-object materials (kind 2)."""
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
    for i in range(20): shade(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps): shade(i)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3
for kind in (1, 2, 1, 2):
    for i, (g, _) in enumerate(scene.gbufs):
        lib.synth_gbuffer(ctx, C.byref(g), kind, bench.SEED + i, pbr.f3(*bench.CAM), pbr.f3(0, 0, 0), None)
    torch.cuda.synchronize()
    us = run()
    out = scene.outs[0].view(torch.float16).float()
    print(f"kind {kind}: {us:.1f} us per 4K frame, {bench.W4K * bench.H4K / us / 1e3:.2f} Gpix/s, frac of {bench.measured_peak_gbs()[0]:.0f} GB/s at 44 B/px: {44 * bench.W4K * bench.H4K / (us * 1e-6) / 1e9 / bench.measured_peak_gbs()[0]:.3f}; finite {bool(out.isfinite().all())} mean {float(out.mean()):.4f}", flush=True)
