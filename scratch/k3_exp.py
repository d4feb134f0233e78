"""A/B of K3 on the bench's 4K frame: python scratch/k3_exp.py 1 0 ...  (1 = fp32 weights everywhere via pbrcu_ibl_set_exact_filtering,
0 = irradiance and LUT through the texture unit); the first run is the reference the later ones are compared with."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import bench, pbr_b200 as pbr
dev = pbr.Device(0)
scene = bench.Scene(pbr, dev, torch)
variants = [int(v) for v in sys.argv[1:]] or [0, 1]
ref = None
for v in variants:
    dev.lib.ibl_set_exact_filtering(dev.ctx, v)
    scene.shade(0); torch.cuda.synchronize()
    out = scene.outs[0].clone()
    for i in range(20): scene.shade(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(200): scene.shade(i)
    b.record(); torch.cuda.synchronize()
    us = a.elapsed_time(b) / 200 * 1e3
    msg = ""
    if ref is None: ref = out
    else:
        d = (out.view(torch.int16).int() - ref.view(torch.int16).int())
        f0 = ref.view(torch.float16).float(); f1 = out.view(torch.float16).float()
        err = (f1 - f0).abs().nan_to_num(1e9)
        tol = 1e-3 + 2.0 ** -9 * f0.abs()
        msg = (f" differing halves {int((d != 0).sum())} max|diff| {float(err.max()):.4g} max err/tol {float((err / tol).max()):.3f}"
               f" psnr {float(10 * torch.log10(f0.max() ** 2 / (err.double() ** 2).mean())):.1f} dB nan {int(f1.isnan().sum())}/{int(f0.isnan().sum())}")
    print(f"exact={v}: {us:.1f} us/frame {bench.W4K*bench.H4K/us/1e3:.2f} Gpix/s{msg}", flush=True)
