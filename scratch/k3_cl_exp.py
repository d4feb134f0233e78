"""(Ran on the build of commit 3f889c4+: PBRCU_K3_CL then selected 1 = gathers / 2 = records; today the gather mode is pbrcu_ibl.flags &
PBRCU_IBL_GATHER.  Results: profiles/r02_k3_variants_a.txt, _b.txt.)
A/B of the K3 variants on the bench's 4K frame (one GPU):  python scratch/k3_cl_exp.py
old = lights in device memory (shade_ggx_one_kernel: footprint records), new = lights on the host (shade_ggx_cl_kernel: gathers),
PBRCU_K3_EXP selects the experimental instantiations of shade_ggx_cl.cu's dispatch.  Prints time, and the differences
to the old kernel's frame of the same filtering mode."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import bench, pbr_b200 as pbr
dev = pbr.Device(0)
scene = bench.Scene(pbr, dev, torch)
lib, ctx = dev.lib, dev.ctx
host_lights = scene.lights_host

def shade(i, lights_ptr):
    g, _ = scene.gbufs[i % len(scene.gbufs)]
    lib.shade_ggx_ibl(ctx, C.byref(g), pbr.f3(*bench.CAM), lights_ptr, bench.N_LIGHTS, None, C.byref(scene.ibl), pbr.WHOLE,
                      scene.outs[i % len(scene.outs)].data_ptr(), bench.W4K * 8, scene.stream())

def run(lights_ptr, reps=200):
    shade(0, lights_ptr); torch.cuda.synchronize()
    out = scene.outs[0].clone()
    for i in range(20): shade(i, lights_ptr)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps): shade(i, lights_ptr)
    b.record(); torch.cuda.synchronize()
    return out, a.elapsed_time(b) / reps * 1e3

def diff(out, ref):
    d = (out.view(torch.int16).int() - ref.view(torch.int16).int())
    f0 = ref.view(torch.float16).float(); f1 = out.view(torch.float16).float()
    err = (f1 - f0).abs().nan_to_num(1e9)
    tol = 1e-3 + 2.0 ** -10 * f0.abs()
    return (f"differing halves {int((d != 0).sum())} max|diff| {float(err.max()):.4g} max err/(1e-3+2^-10|ref|) {float((err / tol).max()):.3f} "
            f"nan {int(f1.isnan().sum())}/{int(f0.isnan().sum())}")

refs = {}
names = {0: "branch ctas5", 1: "nobranch ctas5", 2: "branch ctas4", 3: "branch ctas6", 4: "branch ctas4 late", 5: "branch ctas5 late", 6: "nobranch ctas6", 7: "nobranch ctas5 late", 8: "nobranch ctas6 late", 9: "nobranch ctas4"}
for exact in (0, 1):
    lib.ibl_set_exact_filtering(ctx, exact)
    os.environ.pop("PBRCU_K3_EXP", None); os.environ.pop("PBRCU_K3_CL", None)
    refs[exact], us = run(scene.lights.data_ptr())
    print(f"exact={exact} OLD (device lights, records, shade_ggx_one_kernel): {us:.1f} us/frame {bench.W4K*bench.H4K/us/1e3:.2f} Gpix/s", flush=True)
    for cl, what in ((2, "records"),):
        os.environ["PBRCU_K3_CL"] = str(cl)
        for v in (0, 1, 6, 7, 8, 9):
            os.environ["PBRCU_K3_EXP"] = str(v)
            out, us = run(host_lights.ctypes.data)
            print(f"exact={exact} CL {what} v{v} ({names[v]}): {us:.1f} us/frame {bench.W4K*bench.H4K/us/1e3:.2f} Gpix/s  vs old: {diff(out, refs[exact])}", flush=True)
os.environ.pop("PBRCU_K3_EXP", None)
os.environ["PBRCU_K3_CL"] = "2"
lib.ibl_set_exact_filtering(ctx, 0)
# no IBL, and the run-time light count form
for nl in (0, 1, 3, 4, 7, 16):
    lights = bench_lights = None
    from oracle import synth_lights
    hl = synth_lights(max(nl, 1), seed=2)
    dl = scene.buf(hl.nbytes); dl.copy_(torch.from_numpy(hl.view(np.uint8).reshape(-1)))
    g, _ = scene.gbufs[0]
    outs = []
    for ptr in (dl.data_ptr(), hl.ctypes.data):
        lib.shade_ggx_ibl(ctx, C.byref(g), pbr.f3(*bench.CAM), ptr, nl, None, C.byref(scene.ibl), pbr.WHOLE, scene.outs[0].data_ptr(), bench.W4K * 8, scene.stream())
        torch.cuda.synchronize()
        outs.append(scene.outs[0].clone())
    print(f"nl={nl}: new vs old: {diff(outs[1], outs[0])}", flush=True)
