"""Runs the bench's 590K-triangle geometry pass a few times (for ncu captures of the K10 kernels)."""
import sys, ctypes as C, json
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import torch, numpy as np
import pbr_b200 as pbr
import bench
class S: pass
s = S(); s.pbr = pbr; s.dev = pbr.Device(0); s.torch = torch; s.keep = []
def buf(n):
    t = torch.empty(int(n), dtype=torch.uint8, device="cuda"); s.keep.append(t); return t
s.buf = buf
s.stream = lambda: torch.cuda.current_stream().cuda_stream
from oracle import synth_lights
lh = synth_lights(4, seed=2); s.lights = buf(lh.nbytes); s.lights.copy_(torch.from_numpy(lh.view(np.uint8).reshape(-1)))
# no IBL here: shade without ibl
class FakeIbl(C.Structure): _fields_ = pbr.Ibl._fields_
orig = s.dev.lib.shade_ggx_ibl
s.ibl = pbr.Ibl()
import types
def shade(ctx, g, cam, lights, n, tiles, ibl, rect, out, pitch, st):
    return orig(ctx, g, cam, lights, n, tiles, None, rect, out, pitch, st)
s.dev.lib.__dict__["shade_ggx_ibl"] = shade
print(json.dumps(bench.geometry_bench(s, torch, None, 1, 0)))
