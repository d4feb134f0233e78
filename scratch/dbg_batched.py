import sys, ctypes as C, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import pbr_b200 as pbr
from oracle import Oracle, synth_gbuffer
dev=pbr.Device(0); o=Oracle()
w,h,nv=64,32,3
g2=[synth_gbuffer(w,h,seed=100+v) for v in range(3)]
pos=np.stack([g["pos"] for g in g2]); nrm=np.stack([g["nrm"] for g in g2]); alb=np.stack([g["alb"] for g in g2])
print(pos.shape,pos.dtype,pos.flags['C_CONTIGUOUS'], alb.dtype, alb.shape)
gb=pbr.GBuffer(); gb.width,gb.height,gb.pitch8,gb.pitch4=w,h,w*8,w*4
bp=dev.upload(pos.view(np.uint16)); bn=dev.upload(nrm.view(np.uint16)); ba=dev.upload(alb)
gb.positions=bp.ptr; gb.normals=bn.ptr; gb.albedo=ba.ptr
print('roundtrip', np.array_equal(bp.download(pos.shape,np.uint16), pos.view(np.uint16)), np.array_equal(ba.download(alb.shape,np.uint16), alb))
cams=np.array([[0,0,12]]*3,np.float32)
out=dev.alloc(nv*w*h*8)
dev.lib.lighting_blinnphong_batched(dev.ctx,C.byref(gb),w*h*8,w*h*4,nv,cams.ctypes.data_as(C.POINTER(C.c_float)),out.ptr,w*8,w*h*8,None)
dev.sync()
r=out.download((nv,h,w,4),np.uint16).view(np.float16).astype(np.float32)
for v in range(3):
    # single-view call on view v's planes
    g1=pbr.GBuffer(); g1.width,g1.height,g1.pitch8,g1.pitch4=w,h,w*8,w*4
    g1.positions=bp.ptr+v*w*h*8; g1.normals=bn.ptr+v*w*h*8; g1.albedo=ba.ptr+v*w*h*8
    o1=dev.alloc(w*h*8)
    dev.lib.lighting_blinnphong(dev.ctx,C.byref(g1),pbr.f3(0,0,12),pbr.WHOLE,o1.ptr,w*8,None); dev.sync()
    s=o1.download((h,w,4),np.uint16).view(np.float16).astype(np.float32)
    want=o.lighting(g2[v]["pos"],g2[v]["nrm"],g2[v]["alb"],cams[v])[1].view(np.float16).astype(np.float32)
    print(v,'batched-vs-single',np.abs(r[v]-s).max(),'single-vs-oracle',np.abs(s-want).max())
