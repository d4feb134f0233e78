"""Where do pbrcu_tonemap_aces_gamma's vector path, its scalar path (LDR pointer only 4-byte aligned) and K3's fused LDR store differ?"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import pbr_b200 as pbr
from pbr_b200.synth import synth_lights
dev = pbr.Device(0); lib = dev.lib
w, h = 1920, 1080
gb = pbr.GBuffer(); lib.gbuffer_alloc(dev.ctx, w, h, 1, C.byref(gb))
lib.synth_gbuffer(dev.ctx, C.byref(gb), 1, 0x5EED0099, pbr.f3(0.0, -1.0, -12.0), pbr.f3(0, 0, 0), None)
lights = synth_lights(4, seed=4)
hdr, vec, sca, fus = dev.alloc(w * h * 8), dev.alloc(w * h * 4), dev.alloc(w * h * 4 + 16), dev.alloc(w * h * 4)
cam = pbr.f3(0.0, -1.0, -12.0)
lib.shade_ggx_ibl(dev.ctx, C.byref(gb), cam, lights.ctypes.data, 4, None, None, pbr.WHOLE, hdr.ptr, w * 8, None)
lib.tonemap_aces_gamma(dev.ctx, hdr.ptr, w * 8, w, h, pbr.WHOLE, vec.ptr, w * 4, None)
lib.tonemap_aces_gamma(dev.ctx, hdr.ptr, w * 8, w, h, pbr.WHOLE, sca.ptr + 4, w * 4, None)
lib.shade_ggx_ibl_ldr(dev.ctx, C.byref(gb), cam, lights.ctypes.data, 4, None, None, pbr.WHOLE, fus.ptr, w * 4, None, 0, None)
dev.sync()
H = hdr.download((h, w, 4), np.uint16).view(np.float16).astype(np.float32)
V = vec.download((h, w, 4), np.uint8)
S = sca.download((h * w + 4, 4), np.uint8)[1:h * w + 1].reshape(h, w, 4)
F = fus.download((h, w, 4), np.uint8)
print("vector vs scalar K2:", np.count_nonzero(V != S), " vector vs fused:", np.count_nonzero(V != F), " scalar vs fused:", np.count_nonzero(S != F))
ys, xs, cs = np.nonzero(V != F)
for y, x, c in list(zip(ys, xs, cs))[:8]:
    print(f"  ({x},{y}) ch {c}: hdr {H[y, x]}  vector {V[y, x]}  scalar {S[y, x]}  fused {F[y, x]}  x parity {x & 1}")
