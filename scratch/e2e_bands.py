"""e2e (pbrcu_shade_frame_host) at different band counts."""
import sys, ctypes as C
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import torch, numpy as np
import pbr_b200 as pbr
import bench
dev = pbr.Device(0)
scene = bench.Scene(pbr, dev, torch, n_sets=1)
host, out = bench.host_frame(scene, torch)
lib, ctx = dev.lib, dev.ctx
cam = pbr.f3(*bench.CAM)
for nb in (1, 2, 4, 8, 16, 32, 64):
    def step(_):
        lib.shade_frame_host(ctx, host["positions"].data_ptr(), host["normals"].data_ptr(), host["albedo"].data_ptr(), host["orm"].data_ptr(), host["emissive"].data_ptr(),
                             bench.W4K, bench.H4K, cam, scene.lights_host.ctypes.data, bench.N_LIGHTS, C.byref(scene.ibl), out.data_ptr(), nb, scene.stream())
    ms = bench.timed(torch, None, step, 10, 3) / 10
    print(nb, "bands:", round(ms, 3), "ms", round(bench.W4K * bench.H4K / ms / 1e3, 1), "Mpix/s", flush=True)
