"""e2e through pbrcu_shade_frame_host vs the number of pipelined row bands, beside the raw pinned H2D rate : python scratch/e2e_bands.py"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench, pbr_b200 as pbr
dev = pbr.Device(0); scene = bench.Scene(pbr, dev, torch)
host, out = bench.host_frame(scene, torch)
lib, ctx = dev.lib, dev.ctx
cam = pbr.f3(*bench.CAM)
def ev(fn, n=10):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
big = torch.empty(299 * 1000 * 1000, dtype=torch.uint8, pin_memory=True); dbig = torch.empty_like(big, device="cuda")
ms = ev(lambda: dbig.copy_(big, non_blocking=True)); print(f"raw H2D 299 MB: {ms:.3f} ms = {0.299 / ms * 1e3:.1f} GB/s")
small = torch.empty(66 * 1000 * 1000, dtype=torch.uint8, pin_memory=True); dsmall = torch.empty_like(small, device="cuda")
ms = ev(lambda: small.copy_(dsmall, non_blocking=True)); print(f"raw D2H 66 MB: {ms:.3f} ms = {0.066 / ms * 1e3:.1f} GB/s")
for bands in (1, 2, 4, 8, 16, 32, 64):
    def step():
        lib.shade_frame_host(ctx, host["positions"].data_ptr(), host["normals"].data_ptr(), host["albedo"].data_ptr(),
                             host["orm"].data_ptr(), host["emissive"].data_ptr(), bench.W4K, bench.H4K, cam, scene.lights_host.ctypes.data,
                             bench.N_LIGHTS, C.byref(scene.ibl), out.data_ptr(), bands, scene.stream())
    ms = ev(step)
    print(f"bands {bands:3d}: {ms:.3f} ms/frame = {bench.W4K * bench.H4K / ms / 1e6:.3f} Gpix/s", flush=True)
