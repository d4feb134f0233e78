"""Times K3c (light clustering) alone on the 8K / 256-light frame of the bench : python scratch/k3c_time.py"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch, pbr_b200 as pbr
from oracle import synth_lights
dev = pbr.Device(0); lib, ctx = dev.lib, dev.ctx
W8, H8, NL = 7680, 4320, 256
g = pbr.GBuffer(); lib.gbuffer_alloc(ctx, W8, H8, 1, C.byref(g))
lib.synth_gbuffer(ctx, C.byref(g), 1, 0x5EED0000 + 40, pbr.f3(0.0, -1.0, -12.0), pbr.f3(0, 0, 0), None)
lights = dev.upload(synth_lights(NL, seed=4, ranged=True))
tiles = pbr.LightTiles(); lib.light_tiles_alloc(ctx, W8, H8, NL, C.byref(tiles))
for _ in range(3): lib.cluster_lights(ctx, C.byref(g), lights.ptr, NL, C.byref(tiles), None)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20): lib.cluster_lights(ctx, C.byref(g), lights.ptr, NL, C.byref(tiles), None)
b.record(); torch.cuda.synchronize()
print(f"cluster_lights 8K/256: {a.elapsed_time(b) / 20 * 1e3:.1f} us (positions + normals = {2 * W8 * H8 * 8 / 1e6:.0f} MB)")
