"""torchrun -n 2: sharded prefilter + gather_chain vs unsharded, per level mismatch report."""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import torch, torch.distributed as dist
import pbr_b200 as pbr
from pbr_b200.partition import gather_chain, chain_blocks
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
dev = pbr.Device(int(os.environ["LOCAL_RANK"]))
lib, ctx = dev.lib, dev.ctx
buf = lambda n: torch.empty(int(n), dtype=torch.uint8, device="cuda")
ENV_W, ENV_H, RES, MIPS, NS = 1024, 512, 256, 9, 256
env = buf(ENV_W * ENV_H * 16)
lib.synth_environment(ctx, env.data_ptr(), ENV_W, ENV_H, 0x5EED0002, None)
nb = lib.cube_chain_texels(RES, MIPS) * 8
src, full, dst = buf(nb), buf(nb), buf(nb)
lib.ibl_equirect_to_cube(ctx, env.data_ptr(), ENV_W, ENV_H, RES, src.data_ptr(), None)
lib.ibl_cube_mipgen(ctx, src.data_ptr(), RES, MIPS, None)
s = torch.cuda.current_stream().cuda_stream
lib.ibl_prefilter_ggx(ctx, src.data_ptr(), RES, MIPS, NS, full.data_ptr(), 0, MIPS - 1, 0, 1, s)
dst.zero_()
lib.ibl_prefilter_ggx(ctx, src.data_ptr(), RES, MIPS, NS, dst.data_ptr(), 0, MIPS - 1, rank, world, s)
torch.cuda.synchronize()
own = chain_blocks(RES, MIPS, world, rank)
for m, (o, n) in enumerate(own):
    bad = int((dst[o:o + n] != full[o:o + n]).sum())
    print(f"rank {rank} level {m} own block [{o},{o+n}) mismatches before gather: {bad}", flush=True)
gather_chain(dst, RES, MIPS, world, rank, dist)
torch.cuda.synchronize()
off = 0
for m in range(MIPS):
    r = RES >> m; n = 6 * r * r * 8
    bad = int((dst[off:off + n] != full[off:off + n]).sum())
    print(f"rank {rank} level {m} after gather mismatches: {bad} of {n}", flush=True)
    off += n
dist.destroy_process_group()
