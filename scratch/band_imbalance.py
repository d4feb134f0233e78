"""1 GPU: time each of the 8 row bands of the 8K / 256-light frame separately (K3 only) to see the N=8 imbalance."""
import sys, ctypes as C
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import torch, numpy as np
import pbr_b200 as pbr
from pbr_b200.partition import row_band
from oracle import synth_lights
dev = pbr.Device(0); lib, ctx = dev.lib, dev.ctx
buf = lambda n: torch.empty(int(n), dtype=torch.uint8, device="cuda")
W8, H8, NL = 7680, 4320, 256
g = pbr.GBuffer(); g.width, g.height, g.pitch8, g.pitch4 = W8, H8, W8 * 8, W8 * 4
planes = {k: buf(W8 * H8 * (4 if k in ("depth", "orm") else 8)) for k in ("positions", "normals", "albedo", "depth", "orm", "emissive")}
for k, t in planes.items(): setattr(g, k, t.data_ptr())
lib.synth_gbuffer(ctx, C.byref(g), 1, 0x5EED0000 + 40, pbr.f3(0.5, -1.0, 12.0), pbr.f3(0, 0, 0), None)
hdr = buf(W8 * H8 * 8)
lh = synth_lights(NL, seed=4, ranged=True); lights = buf(lh.nbytes); lights.copy_(torch.from_numpy(lh.view(np.uint8).reshape(-1)))
tiles = pbr.LightTiles(); lib.light_tiles_alloc(ctx, W8, H8, NL, C.byref(tiles))
s = torch.cuda.current_stream().cuda_stream
def t(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / n
print("cluster ms", t(lambda: lib.cluster_lights(ctx, C.byref(g), lights.data_ptr(), NL, C.byref(tiles), s)))
cam = pbr.f3(0.5, -1.0, 12.0)
for world in (2, 8):
    ts = []
    for r in range(world):
        y0, y1 = row_band(H8, world, r, align=16)
        band = pbr.Rect(0, y0, W8, y1 - y0)
        ts.append(t(lambda: lib.shade_ggx_ibl(ctx, C.byref(g), cam, lights.data_ptr(), NL, C.byref(tiles), None, band, hdr.data_ptr(), W8 * 8, s)))
    print(world, "bands ms", [round(x, 3) for x in ts], "sum", round(sum(ts), 3), "max", round(max(ts), 3))
