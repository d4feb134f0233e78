"""PCIe rates of the box: pinned H2D alone, D2H alone, both at once (two streams) -- the floor of the e2e number."""
import torch
n_in, n_out = 298598400, 66355200
h_in = torch.empty(n_in, dtype=torch.uint8, pin_memory=True); d_in = torch.empty(n_in, dtype=torch.uint8, device="cuda")
h_out = torch.empty(n_out, dtype=torch.uint8, pin_memory=True); d_out = torch.empty(n_out, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    torch.cuda.synchronize(); b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / n
ms = t(lambda: d_in.copy_(h_in, non_blocking=True)); print(f"H2D 299 MB: {ms:.3f} ms = {n_in/ms/1e6:.1f} GB/s")
ms = t(lambda: h_out.copy_(d_out, non_blocking=True)); print(f"D2H 66 MB: {ms:.3f} ms = {n_out/ms/1e6:.1f} GB/s")
def both():
    with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
ms = t(both); print(f"both: {ms:.3f} ms -> frame floor {3840*2160/ms/1e3:.1f} Mpix/s")
