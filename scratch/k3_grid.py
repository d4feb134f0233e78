import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench, pbr_b200 as pbr
dev = pbr.Device(0); scene = bench.Scene(pbr, dev, torch)
os.environ["PBRCU_K3_VARIANT"] = sys.argv[1] if len(sys.argv) > 1 else "1"
for g in (1, 2, 3, 4, 5, 6):
    os.environ["PBRCU_K3_GRID"] = str(g)
    for i in range(10): scene.shade(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(100): scene.shade(i)
    b.record(); torch.cuda.synchronize()
    print(f"ctas/SM {g}: {a.elapsed_time(b)*10:.1f} us", flush=True)
