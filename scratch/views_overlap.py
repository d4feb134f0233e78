"""Rasterise + shade 128 views of the bench scene: geometry of all views then shading of all views (as bench.py does), against chunks of
views with the shading of chunk c on a second stream while chunk c + 1 is rasterised."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import pbr_b200 as pbr
from pbr_b200.synth import checker_texture, col_major, look_at_rh, mesh_vertex, model_push_constant, perspective_rh_no, uv_sphere
dev = pbr.Device(0); lib, ctx = dev.lib, dev.ctx
W, H, NV = 1920, 1080, 128
rng = np.random.default_rng(7)
keep = []
def dev_of(a):
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda(); keep.append(t); return t
sv, si = uv_sphere(1.0, 48, 96)
g = 40.0
quad = np.array([mesh_vertex((-g, 3.0, -g), (0, -1, 0), (1, 0, 0, 1), (0, 0)), mesh_vertex((g, 3.0, -g), (0, -1, 0), (1, 0, 0, 1), (16, 0)),
                 mesh_vertex((g, 3.0, g), (0, -1, 0), (1, 0, 0, 1), (16, 16)), mesh_vertex((-g, 3.0, g), (0, -1, 0), (1, 0, 0, 1), (0, 16))], np.float32)
qi = np.array([0, 1, 2, 0, 2, 3, 0, 2, 1, 0, 3, 2], np.uint16)
vb, ib, qvb, qib = dev_of(sv), dev_of(si), dev_of(quad), dev_of(qi)
ctex, ntex, flat = dev_of(checker_texture(64, 64, 1)), dev_of(checker_texture(32, 32, 2, normal_map=True)), dev_of(np.array([[[128, 128, 255, 255]]], np.uint8))
draws = (pbr.Draw * 65)()
def fill(d, vbuf, ibuf, nidx, model, color, ct, cshape, nt, nshape):
    d.vertices, d.indices, d.first_index, d.index_count, d.vertex_offset = vbuf.data_ptr(), ibuf.data_ptr(), 0, nidx, 0
    d.vertex_count = vbuf.numel() // 48
    d.model[:] = model[0].tolist(); d.normal_model[:] = model[1].tolist()
    m = d.material; m.color[:] = [float(c) for c in color]
    m.color_tex.texels, m.color_tex.width, m.color_tex.height, m.color_tex.pitch = ct.data_ptr(), cshape[1], cshape[0], cshape[1] * 4
    m.normal_tex.texels, m.normal_tex.width, m.normal_tex.height, m.normal_tex.pitch = nt.data_ptr(), nshape[1], nshape[0], nshape[1] * 4
    m.occlusion, m.roughness, m.metallic = 1.0, 0.5, 0.0
for k in range(64):
    fill(draws[k], vb, ib, len(si), model_push_constant(rng.uniform(-8, 8, 3), rng.uniform(0.5, 1.5, 3), rng.standard_normal(3), rng.uniform(0, 6.28)), rng.uniform(0.3, 1.0, 4), ctex, (64, 64), ntex, (32, 32))
fill(draws[64], qvb, qib, len(qi), model_push_constant(), (0.8, 0.8, 0.8, 1.0), ctex, (64, 64), flat, (1, 1))
s8, s4 = W * H * 8, W * H * 4
planes = {k: torch.empty(NV * (s4 if k in ("depth", "orm") else s8), dtype=torch.uint8, device="cuda") for k in ("positions", "normals", "albedo", "depth", "orm", "emissive")}
gv = pbr.GBuffer(); gv.width, gv.height, gv.pitch8, gv.pitch4 = W, H, W * 8, W * 4
for k, t in planes.items(): setattr(gv, k, t.data_ptr())
proj = col_major(perspective_rh_no(np.pi / 2, W / H))
views = np.ascontiguousarray(np.stack([col_major(look_at_rh((12 * np.cos(a), -2.0, 12 * np.sin(a)), (0, 0, 0))) for a in np.linspace(0, 2 * np.pi, NV, endpoint=False)]).astype(np.float32))
projs = np.ascontiguousarray(np.tile(np.asarray(proj, np.float32).reshape(1, 16), (NV, 1)))
st = torch.cuda.current_stream().cuda_stream
from pbr_b200.synth import synth_lights
lights = synth_lights(4, seed=2)
out = torch.empty(NV * s8, dtype=torch.uint8, device="cuda")
cams = np.ascontiguousarray(np.array([(12 * np.cos(a), -2.0, 12 * np.sin(a)) for a in np.linspace(0, 2 * np.pi, NV, endpoint=False)], np.float32))
sG, sS = torch.cuda.Stream(), torch.cuda.Stream()
def gview(v0):
    g = pbr.GBuffer(); g.width, g.height, g.pitch8, g.pitch4 = W, H, W * 8, W * 4
    for k, t in planes.items(): setattr(g, k, t.data_ptr() + v0 * (s4 if k in ("depth", "orm") else s8))
    return g
def geometry(v0, n, stream):
    g = gview(v0)
    rc = lib.geometry_pass_batched(ctx, C.byref(g), s8, s4, n, views[v0:].ctypes.data, projs[v0:].ctypes.data, draws, 65, stream.cuda_stream); assert rc == 0
def shade(v0, n, stream):
    g = gview(v0)
    rc = lib.shade_ggx_ibl_batched(ctx, C.byref(g), s8, s4, n, cams[v0:].ctypes.data_as(C.POINTER(C.c_float)), lights.ctypes.data, 4, None, out.data_ptr() + v0 * s8, W * 8, s8, stream.cuda_stream); assert rc == 0
def sequential():
    geometry(0, NV, sG); shade(0, NV, sG)
def chunked(cs):
    def run():
        for v0 in range(0, NV, cs):
            n = min(cs, NV - v0)
            geometry(v0, n, sG)
            e = torch.cuda.Event(); e.record(sG); sS.wait_event(e)
            shade(v0, n, sS)
        e = torch.cuda.Event(); e.record(sS); sG.wait_event(e)
    return run
def timeit(fn, tag):
    global refsum
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(sG); fn(); fn(); b.record(sG); torch.cuda.synchronize()
    chk = int(out.view(torch.int32).sum().item())
    print(f"{tag}: {a.elapsed_time(b) / 2:.2f} ms for {NV} views  checksum {chk}", flush=True)
timeit(sequential, "all geometry, then all shading")
for cs in (64, 32, 16, 8):
    timeit(chunked(cs), f"chunks of {cs}, shading on a second stream")
timeit(sequential, "all geometry, then all shading")
