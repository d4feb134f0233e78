"""A/B of the tiled K3 kernel on the 8K / 256-light frame (BASELINE configs[3]): PBRCU_K3_PRE = 0 / 1, alternating."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import pbr_b200 as pbr
from pbr_b200.synth import synth_lights
import bench
dev = pbr.Device(0); lib, ctx = dev.lib, dev.ctx
W, H, NL = 7680, 4320, 256
gb = pbr.GBuffer(); lib.gbuffer_alloc(ctx, W, H, 1, C.byref(gb))
lib.synth_gbuffer(ctx, C.byref(gb), 1, 1234, pbr.f3(0.0, -1.0, -6.0), pbr.f3(0, 0, 0), None)
lights = synth_lights(NL, seed=4, ranged=True)
ld = torch.from_numpy(lights.view(np.uint8).reshape(-1)).cuda()
tiles = pbr.LightTiles(); lib.light_tiles_alloc(ctx, W, H, NL, C.byref(tiles))
out = torch.empty(W * H * 8, dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
lib.cluster_lights(ctx, C.byref(gb), ld.data_ptr(), NL, C.byref(tiles), st)
def frame():
    lib.shade_ggx_ibl(ctx, C.byref(gb), pbr.f3(0.0, -1.0, -6.0), ld.data_ptr(), NL, C.byref(tiles), None, pbr.WHOLE, out.data_ptr(), W * 8, st)
ref = None
for rep in range(3):
    for pre in ("0", "1"):
        os.environ["PBRCU_K3_PRE"] = pre
        for _ in range(3): frame()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20): frame()
        b.record(); torch.cuda.synchronize()
        chk = int(out.view(torch.int32).sum().item())
        if ref is None: ref = chk
        print(f"pre={pre}: {a.elapsed_time(b) / 20 * 1e3:.1f} us (K3 only, no IBL), same {chk == ref}", flush=True)
