import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import bench, pbr_b200 as pbr
dev = pbr.Device(0)
scene = bench.Scene(pbr, dev, torch)
outs = []
for v in (1, 0):
    dev.lib.ibl_set_exact_filtering(dev.ctx, v)
    scene.shade(0); torch.cuda.synchronize()
    outs.append(scene.outs[0].clone().view(torch.float16).float().view(bench.H4K, bench.W4K, 4))
ref, out = outs
err = (out - ref).abs()[..., :3].amax(-1)
tol = 1e-3 + 2.0 ** -9 * ref[..., :3].abs().amax(-1)
bad = (err > tol)
print("pixels over tolerance:", int(bad.sum()), "of", bad.numel())
g, planes = scene.gbufs[0]
nrm = planes["normals"].view(torch.float16).float().view(bench.H4K, bench.W4K, 4)
idx = torch.topk((err / tol).flatten(), 12).indices
for i in idx.tolist():
    y, x = divmod(i, bench.W4K)
    n = nrm[y, x, :3]; n = n / n.norm()
    print(y, x, "N", [round(float(c), 4) for c in n], "ref", [round(float(c), 4) for c in ref[y, x, :3]], "out", [round(float(c), 4) for c in out[y, x, :3]])
# how are the bad pixels distributed over |N| components (face edges have two equal-magnitude components)
n = nrm[..., :3]; n = n / n.norm(dim=-1, keepdim=True).clamp_min(1e-6)
a = n.abs().sort(dim=-1, descending=True).values
ratio = (a[..., 1] / a[..., 0].clamp_min(1e-6))[bad]
print("second/first |N| component of bad pixels: min %.4f median %.4f" % (float(ratio.min()), float(ratio.median())))
