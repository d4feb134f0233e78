"""(Ran on a build that carried the L2 cache-policy plumbing -- PBRCU_K3_L2 -- which was reverted after this measurement: profiles/r02_k3_l2_hints.txt.)
K3 on the bench's 4K frame with L2 cache-policy hints (PBRCU_K3_L2: bit 0 = evict-first on the staged planes, bit 1 = evict-last on the
prefiltered records), both filter modes; results must stay bit-identical (hints do not change values)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import bench, pbr_b200 as pbr
dev = pbr.Device(0)
scene = bench.Scene(pbr, dev, torch)
lib, ctx = dev.lib, dev.ctx

def run(reps=300):
    scene.shade(0); torch.cuda.synchronize()
    out = scene.outs[0].clone()
    for i in range(30): scene.shade(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps): scene.shade(i)
    b.record(); torch.cuda.synchronize()
    return out, a.elapsed_time(b) / reps * 1e3

for exact in (0, 1):
    lib.ibl_set_exact_filtering(ctx, exact)
    ref = None
    for mode in (0, 1, 2, 3, 0):
        os.environ["PBRCU_K3_L2"] = str(mode)
        out, us = run()
        if ref is None: ref = out
        print(f"exact={exact} l2_mode={mode}: {us:.1f} us/frame {bench.W4K*bench.H4K/us/1e3:.2f} Gpix/s same_bits={bool(torch.equal(out, ref))}", flush=True)
lib.ibl_set_exact_filtering(ctx, 0)
