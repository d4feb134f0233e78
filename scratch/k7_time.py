"""Times the 2048^2 12-mip GGX prefilter (bench Scene's own measurement) : python scratch/k7_time.py"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench, pbr_b200 as pbr
dev = pbr.Device(0); scene = bench.Scene(pbr, dev, torch)
print(json.dumps(scene.ibl_ms)); print(json.dumps(scene.prefilter_error))
lib, ctx = dev.lib, dev.ctx
for m in range(1, bench.CUBE_MIPS):
    lib.ibl_prefilter_ggx(ctx, scene.src_chain.data_ptr(), bench.CUBE_RES, bench.CUBE_MIPS, bench.IBL_SAMPLES, scene.pref.data_ptr(), m, m, 0, 1, scene.stream())
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    lib.ibl_prefilter_ggx(ctx, scene.src_chain.data_ptr(), bench.CUBE_RES, bench.CUBE_MIPS, bench.IBL_SAMPLES, scene.pref.data_ptr(), m, m, 0, 1, scene.stream())
    b.record(); torch.cuda.synchronize()
    print(f"level {m} (res {bench.CUBE_RES >> m}): {a.elapsed_time(b):.3f} ms (incl. the texture upload of the source chain)", flush=True)
