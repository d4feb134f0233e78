"""Runs the config-3 prefilter once (for ncu captures of conv_kernel)."""
import sys, ctypes as C
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import torch
import pbr_b200 as pbr
dev = pbr.Device(0)
lib, ctx = dev.lib, dev.ctx
cuda = torch.device("cuda", 0)
buf = lambda n: torch.empty(int(n), dtype=torch.uint8, device=cuda)
ENV_W, ENV_H, RES, MIPS, NS = 4096, 2048, 2048, 12, 1024
env = buf(ENV_W * ENV_H * 16)
lib.synth_environment(ctx, env.data_ptr(), ENV_W, ENV_H, 0x5EED0002, None)
nb = lib.cube_chain_texels(RES, MIPS) * 8
src, dst = buf(nb), buf(nb)
lib.ibl_equirect_to_cube(ctx, env.data_ptr(), ENV_W, ENV_H, RES, src.data_ptr(), None)
lib.ibl_cube_mipgen(ctx, src.data_ptr(), RES, MIPS, None)
for it in range(2):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    lib.ibl_prefilter_ggx(ctx, src.data_ptr(), RES, MIPS, NS, dst.data_ptr(), 0, MIPS - 1, 0, 1, None)
    b.record(); torch.cuda.synchronize()
    print("prefilter ms", a.elapsed_time(b))
