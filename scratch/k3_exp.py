"""A/B of K3 kernel variants on the bench's 4K frame: python scratch/k3_exp.py 0 1 ...  (PBRCU_K3_VARIANT values)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import bench, pbr_b200 as pbr
dev = pbr.Device(0)
scene = bench.Scene(pbr, dev, torch)
variants = [int(v) for v in sys.argv[1:]] or [0, 1]
ref = None
for v in variants:
    os.environ["PBRCU_K3_VARIANT"] = str(v)
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
        msg = f" differing halves {int((d != 0).sum())} max|diff| {float((f1 - f0).abs().nan_to_num(1e9).max()):.4g} nan {int(f1.isnan().sum())}/{int(f0.isnan().sum())}"
    print(f"variant {v}: {us:.1f} us/frame {bench.W4K*bench.H4K/us/1e3:.2f} Gpix/s{msg}", flush=True)
