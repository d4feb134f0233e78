"""(Ran on a build that carried scratch/shade_ggx_cl_pipe.cu.txt -- PBRCU_K3_PIPE -- which was removed after this measurement: profiles/r02_k3_pipe.txt.)
K3 on the bench's 4K frame: default shade_ggx_cl_kernel vs the L2-prefetch pipelined form (PBRCU_K3_PIPE=1: 5 CTAs/SM, =2: 4 CTAs/SM), both filter modes."""
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
    for mode in ("", "1", "3", ""):
        if mode: os.environ["PBRCU_K3_PIPE"] = mode
        else: os.environ.pop("PBRCU_K3_PIPE", None)
        out, us = run()
        if ref is None: ref = out
        d = (out.view(torch.int16).int() - ref.view(torch.int16).int()).abs()
        print(f"exact={exact} pipe={mode or '0'}: {us:.1f} us/frame {bench.W4K*bench.H4K/us/1e3:.2f} Gpix/s  differing halves {int((d != 0).sum())} max ulps {int(d.max())}", flush=True)
lib.ibl_set_exact_filtering(ctx, 0)
