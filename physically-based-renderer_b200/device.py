"""numpy-level conveniences over the C ABI (tests / bench only; no compute happens here)."""
from __future__ import annotations

import ctypes as C

import numpy as np


class DeviceBuffer:
    def __init__(self, dev: "Device", nbytes: int):
        self.dev = dev
        self.nbytes = int(nbytes)
        p = C.c_void_p()
        dev.lib.malloc(dev.ctx, self.nbytes, C.byref(p))
        self.ptr = p.value or 0

    def free(self):
        if self.ptr:
            self.dev.lib.free(self.dev.ctx, self.ptr)
            self.ptr = 0

    def upload(self, arr: np.ndarray, stream=None):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes <= self.nbytes
        self.dev.lib.memcpy_h2d(self.ptr, arr.ctypes.data, arr.nbytes, stream)
        self.dev.lib.stream_synchronize(stream)
        return self

    def download(self, shape, dtype, stream=None) -> np.ndarray:
        out = np.empty(shape, dtype)
        assert out.nbytes <= self.nbytes
        self.dev.lib.memcpy_d2h(out.ctypes.data, self.ptr, out.nbytes, stream)
        self.dev.lib.stream_synchronize(stream)
        return out

    def zero(self, stream=None):
        self.dev.lib.memset(self.ptr, 0, self.nbytes, stream)


class Device:
    """One pbrcu context on one GPU.  Raises PbrcuError (status 2) when there is no CUDA device."""

    def __init__(self, ordinal: int = 0):
        from . import load

        self.lib = load()
        ctx = C.c_void_p()
        self.lib.ctx_create(ordinal, C.byref(ctx))
        self.ctx = ctx
        self.ordinal = ordinal
        self._bufs: list[DeviceBuffer] = []

    @property
    def sm_count(self) -> int:
        return self.lib.ctx_sm_count(self.ctx)

    def alloc(self, nbytes: int) -> DeviceBuffer:
        b = DeviceBuffer(self, nbytes)
        self._bufs.append(b)
        return b

    def upload(self, arr: np.ndarray) -> DeviceBuffer:
        arr = np.ascontiguousarray(arr)
        return self.alloc(max(arr.nbytes, 16)).upload(arr)

    def gbuffer(self, positions, normals, albedo, orm=None, emissive=None, depth=None):
        """Upload tightly packed host planes ((H,W,4) arrays) and describe them as a pbrcu_gbuffer."""
        from . import GBuffer

        h, w = positions.shape[:2]
        g = GBuffer()
        g.width, g.height = w, h
        g.pitch8, g.pitch4 = w * 8, w * 4
        g.positions = self.upload(positions.view(np.uint16)).ptr
        g.normals = self.upload(normals.view(np.uint16)).ptr
        g.albedo = self.upload(albedo).ptr
        if depth is not None:
            g.depth = self.upload(depth.astype(np.float32)).ptr
        if orm is not None:
            g.orm = self.upload(orm).ptr
        if emissive is not None:
            g.emissive = self.upload(emissive.view(np.uint16)).ptr
        return g

    def sync(self, stream=None):
        self.lib.stream_synchronize(stream)

    def close(self):
        for b in self._bufs:
            b.free()
        self._bufs.clear()
        if self.ctx:
            self.lib.ctx_destroy(self.ctx)
            self.ctx = None
