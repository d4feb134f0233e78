"""Where the e2e frame's 0.7 ms over the raw H2D time goes: copy granularity vs. H2D/D2H interference : python scratch/pcie_overlap.py"""
import torch
W, H = 3840, 2160
planes = [W * H * 8] * 4 + [W * H * 4]
host = [torch.empty(n, dtype=torch.uint8, pin_memory=True) for n in planes]
dev = [torch.empty(n, dtype=torch.uint8, device="cuda") for n in planes]
hout = torch.empty(W * H * 8, dtype=torch.uint8, pin_memory=True); dout = torch.empty(W * H * 8, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def ev(fn, n=10):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
        torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
def h2d(bands, stream):
    with torch.cuda.stream(stream):
        for k in range(bands):
            for h, d in zip(host, dev):
                n = h.numel() // bands
                d[k * n:(k + 1) * n].copy_(h[k * n:(k + 1) * n], non_blocking=True)
def d2h(bands, stream):
    with torch.cuda.stream(stream):
        n = hout.numel() // bands
        for k in range(bands):
            hout[k * n:(k + 1) * n].copy_(dout[k * n:(k + 1) * n], non_blocking=True)
def start():
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
for bands in (1, 8, 32):
    print(f"H2D alone, {bands} bands x 5 planes: {ev(lambda: (start(), h2d(bands, s1))):.3f} ms")
print(f"D2H alone, 8 bands: {ev(lambda: (start(), d2h(8, s2))):.3f} ms")
for bands in (1, 8):
    print(f"H2D ({bands} bands) with concurrent D2H (8 bands): {ev(lambda: (start(), h2d(bands, s1), d2h(8, s2))):.3f} ms")
