"""Synthetic inputs for tests and benchmarks: seeded G-buffers, light lists, environments, cameras and small meshes, as
plain numpy.  Host-side helper of the package: no compute, no checker -- bench.py's GPU arm imports its
inputs from here, tests/oracle.py re-exports the same functions for the tests."""
from __future__ import annotations

import numpy as np


def look_at_rh(eye, center, up=(0.0, -1.0, 0.0)):
    eye, center, up = (np.asarray(v, np.float64) for v in (eye, center, up))
    f = center - eye; f /= np.linalg.norm(f)
    s = np.cross(f, up); s /= np.linalg.norm(s)
    u = np.cross(s, f)
    m = np.eye(4)
    m[0, :3], m[1, :3], m[2, :3] = s, u, -f
    m[0, 3], m[1, 3], m[2, 3] = -s @ eye, -u @ eye, f @ eye
    return m  # row-major maths matrix; use .T.reshape(-1) for column-major storage


def perspective_rh_no(fovy, aspect, near=0.01, far=1024.0):
    t = np.tan(fovy / 2)
    m = np.zeros((4, 4))
    m[0, 0], m[1, 1] = 1 / (aspect * t), 1 / t
    m[2, 2], m[3, 2], m[2, 3] = -(far + near) / (far - near), -1.0, -(2 * far * near) / (far - near)
    return m


def col_major(m):
    return np.ascontiguousarray(np.asarray(m, np.float64).T.reshape(-1), np.float32)


def model_push_constant(translation=(0, 0, 0), scale=(1, 1, 1), rot_axis=(0, 1, 0), rot_angle=0.0):
    """(model[16], normalModel[12]) like makeModelPushConstant: T * R * S and mat3(transpose(inverse(model)))
    padded to three vec4 columns."""
    a = np.asarray(rot_axis, np.float64); a /= np.linalg.norm(a)
    c, s = np.cos(rot_angle), np.sin(rot_angle)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    R = np.eye(3) + s * K + (1 - c) * K @ K
    M = np.eye(4)
    M[:3, :3] = R @ np.diag(np.asarray(scale, np.float64))
    M[:3, 3] = translation
    nm = np.linalg.inv(M).T[:3, :3]
    nm4 = np.zeros((3, 4)); nm4[:, :3] = nm.T  # three columns, each padded to a vec4
    return col_major(M), np.ascontiguousarray(nm4.reshape(-1), np.float32)


def mesh_vertex(pos, nrm=(0, 0, 1), tan=(1, 0, 0, 1), uv=(0, 0)):
    return [*pos, *nrm, *tan, *uv]


def uv_sphere(radius=1.0, stacks=16, slices=32):
    """UV sphere as MeshVertex rows + uint16 triangle indices, outward normals, clockwise front faces in the
    reference's framebuffer convention (y down), tangents along +longitude."""
    v = []
    for i in range(stacks + 1):
        th = np.pi * i / stacks
        for j in range(slices + 1):
            ph = 2 * np.pi * j / slices
            n = np.array([np.sin(th) * np.cos(ph), np.cos(th), np.sin(th) * np.sin(ph)])
            t = np.array([-np.sin(ph), 0.0, np.cos(ph)])
            v.append(mesh_vertex(radius * n, n, (*t, 1.0), (j / slices * 4, i / stacks * 2)))
    idx = []
    for i in range(stacks):
        for j in range(slices):
            a, b = i * (slices + 1) + j, (i + 1) * (slices + 1) + j
            idx += [a, a + 1, b, a + 1, b + 1, b]
    return np.array(v, np.float32), np.array(idx, np.uint16)


def checker_texture(w, h, seed, normal_map=False):
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    if normal_map:  # perturbed +Z normals: xy in [96,160], z high
        img[..., 0:2] = rng.integers(96, 161, (h, w, 2))
        img[..., 2] = rng.integers(200, 256, (h, w))
    img[..., 3] = 255
    return img


def synth_gbuffer(w, h, seed, background=0.2, tier_b=False, cam_at_origin_bg=True):
    """A seeded G-buffer in the reference's formats: unit normals, positions in [-10,10]^3, random
    UNORM16 albedo, `background` fraction of cleared (all-zero) texels."""
    rng = np.random.default_rng(seed)
    pos = np.zeros((h, w, 4), np.float16)
    pos[..., :3] = rng.uniform(-10, 10, (h, w, 3))
    n = rng.standard_normal((h, w, 3))
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    nrm = np.zeros((h, w, 4), np.float16)
    nrm[..., :3] = n
    alb = rng.integers(0, 65536, (h, w, 4), dtype=np.uint16)
    bg = rng.random((h, w)) < background
    pos[bg] = 0
    nrm[bg] = 0
    alb[bg] = 0
    out = dict(pos=pos, nrm=nrm, alb=alb, bg=bg)
    if tier_b:
        orm = np.zeros((h, w, 4), np.uint8)
        orm[..., 0] = rng.integers(128, 256, (h, w))
        orm[..., 1] = rng.integers(0, 256, (h, w))
        m = rng.random((h, w))
        orm[..., 2] = np.where(m < 0.45, 0, np.where(m < 0.9, 255, rng.integers(0, 256, (h, w))))
        orm[..., 3] = 255
        emi = np.zeros((h, w, 4), np.float16)
        e = rng.random((h, w)) < 0.05
        emi[e, :3] = rng.uniform(0, 4, (int(e.sum()), 3))
        orm[bg] = 0
        emi[bg] = 0
        out.update(orm=orm, emi=emi)
    return out


def synth_environment(sw, sh, seed, sun=50.0):
    """Small procedural lat-long HDR sky: gradient + bright sun disc + noise (RGBA32F)."""
    rng = np.random.default_rng(seed)
    v = (np.arange(sh) + 0.5) / sh
    u = (np.arange(sw) + 0.5) / sw
    theta = v[:, None] * np.pi
    phi = (u[None, :] - 0.5) * 2 * np.pi
    d = np.stack([np.sin(theta) * np.cos(phi), np.cos(theta) * np.ones_like(phi), np.sin(theta) * np.sin(phi)], -1)
    t = 0.5 * (d[..., 1] + 1)
    env = np.ones((sh, sw, 4), np.float32)
    env[..., 0] = 0.15 + 0.3 * t
    env[..., 1] = 0.2 + 0.5 * t
    env[..., 2] = 0.25 + 1.4 * t
    env[..., :3] *= (1 + 0.1 * (rng.random((sh, sw, 1)) - 0.5)).astype(np.float32)
    s = np.array([0.35, 0.8, 0.45])
    s /= np.linalg.norm(s)
    env[(d @ s) > np.cos(np.radians(4.0)), :3] = sun
    return env


def synth_lights(n, seed, ranged=False):
    rng = np.random.default_rng(seed)
    L = np.zeros((n, 8), np.float32)
    L[:, :3] = rng.uniform(-8, 8, (n, 3))
    L[:, 3] = rng.uniform(2, 6, n) if ranged else 0.0
    L[:, 4:7] = rng.uniform(0.2, 1.0, (n, 3))
    L[:, 7] = rng.uniform(50, 200, n)
    return L
