"""CPU: the glTF import path's parser (include/pbr/gltf/Document.hpp + Json.hpp) on assets written by
tests/gltfwriter.py: GLB and .gltf containers, external / data: / embedded buffers and images, interleaved and
separate vertex buffers, u8 / u16 / u32 / generated indices, normalised integer attributes, TRS and matrix nodes,
node hierarchies.  tests/cpp/gltf_dump prints what the reference's gltf::Asset reads (gltf/Asset.cpp:158-273), the
test prints the same lines from the data it wrote.  (fastgltf itself is not available: parity UNPINNED for this layer.)"""
import subprocess
from pathlib import Path

import numpy as np
import pytest
from gltfwriter import UBYTE, UINT, USHORT, Builder
from oracle import checker_texture, mesh_vertex, uv_sphere

ROOT = Path(__file__).resolve().parent.parent
TOOL = ROOT / "tests" / "cpp" / "build" / "gltf_dump"


@pytest.fixture(scope="module")
def tool():
    subprocess.run(["make", "-s", "-C", str(ROOT / "tests" / "cpp"), "cpu"], check=True)
    return TOOL


def fnv(data: bytes) -> str:
    h = 1469598103934665603
    for b in data:
        h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return f"{h:016x}"


def g9(x) -> str:
    return "%.9g" % float(np.float32(x))


def quat_from_matrix(R):
    """the 'Christian' method Document.hpp restates (rows/cols as in a column-vector rotation matrix)"""
    R = R.astype(np.float32)
    f = np.float32
    x = np.sqrt(max(f(0), f(1) + R[0, 0] - R[1, 1] - R[2, 2])) / f(2)
    y = np.sqrt(max(f(0), f(1) - R[0, 0] + R[1, 1] - R[2, 2])) / f(2)
    z = np.sqrt(max(f(0), f(1) - R[0, 0] - R[1, 1] + R[2, 2])) / f(2)
    w = np.sqrt(max(f(0), f(1) + R[0, 0] + R[1, 1] + R[2, 2])) / f(2)
    return (np.copysign(x, R[2, 1] - R[1, 2]), np.copysign(y, R[0, 2] - R[2, 0]), np.copysign(z, R[1, 0] - R[0, 1]), w)


def build_asset():
    """Returns (Builder, expected dump lines)."""
    rng = np.random.default_rng(5)
    b = Builder()
    sv, si = uv_sphere(1.0, 6, 8)
    sv = sv.reshape(-1, 12).astype(np.float32)
    quad = np.array([mesh_vertex((-1, -1, 0), (0, 0, -1), (1, 0, 0, 1), (0, 0)), mesh_vertex((1, -1, 0), (0, 0, -1), (1, 0, 0, 1), (1, 0)),
                     mesh_vertex((1, 1, 0), (0, 0, -1), (1, 0, 0, -1), (1, 1)), mesh_vertex((-1, 1, 0), (0, 0, -1), (1, 0, 0, -1), (0, 1))], np.float32)
    tri_soup = quad[[0, 1, 2, 0, 2, 3]]  # no indices: GenerateMeshIndices
    texs = {name: checker_texture(w, h, seed, normal_map=nm) for name, (w, h, seed, nm) in
            dict(colA=(16, 8, 1, False), nrmA=(8, 8, 2, True), colB=(5, 7, 3, False), nrmB=(4, 4, 4, True)).items()}
    tA = b.image(texs["colA"], "colA", how="view", filters=(4, 1), level=9)
    tnA = b.image(texs["nrmA"], "nrmA", how="data", interlace=True)
    tB = b.image(texs["colB"], "colB", how="file", level=0)
    tnB = b.image(texs["nrmB"], "nrmB", how="view")
    mA = b.material("matA", (0.9, 0.6, 0.3, 1.0), tA, tnA, metallic=0.25, roughness=0.5, emissive=(0.1, 0.0, 0.2))
    mB = b.material("matB", (0.2, 0.9, 0.4, 0.5), tB, tnB)
    mats = {mA: ("matA", (0.9, 0.6, 0.3, 1.0), 0.25, 0.5, (0.1, 0.0, 0.2), "colA", "nrmA"), mB: ("matB", (0.2, 0.9, 0.4, 0.5), 1.0, 1.0, (0, 0, 0), "colB", "nrmB")}
    quad_q = quad.copy()
    quad_q[:, 10:12] = (np.round(quad[:, 10:12] * 65535) / np.float32(65535.0)).astype(np.float32)  # what the quantised uv decode to
    prims = {
        "sphere": [(b.primitive(sv, si, mA, layout="interleaved"), sv, np.asarray(si, np.uint16), mA),
                   (b.primitive(quad, [0, 1, 2, 0, 2, 3], mB, index_type=UBYTE, uv_quantised=True), quad_q, np.array([0, 1, 2, 0, 2, 3], np.uint16), mB)],
        "soup": [(b.primitive(tri_soup, None, mB), tri_soup, np.arange(6, dtype=np.uint16), mB)],
        "wide": [(b.primitive(sv, si, mA, index_type=UINT), sv, np.asarray(si, np.uint16), mA)],
    }
    meshes = {name: b.mesh(name, [p[0] for p in plist]) for name, plist in prims.items()}
    # nodes: a TRS parent with a TRS child, a matrix node, a bare transform-less node
    ang = 0.7
    Rm = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    S = np.diag([1.5, 0.5, 2.0])
    M = np.eye(4); M[:3, :3] = Rm @ S; M[:3, 3] = (0.5, -1.0, 2.0)
    child = b.node("child", mesh=meshes["soup"], translation=(0, 1, 0), scale=(0.5, 0.5, 0.5))
    parent = b.node("parent", mesh=meshes["sphere"], translation=(0.5, 0.2, 0.0), rotation_xyzw=(0.0, 0.3826834, 0.0, 0.9238795), scale=(1.5, 0.7, 1.0), children=[child])
    mat_node = b.node("matrixNode", mesh=meshes["wide"], matrix=M)
    bare = b.node("bare")
    b.scene([parent, mat_node, bare], name="main")

    def prim_lines(name, depth):
        out = []
        for _p, v, i, m in prims[name]:
            mname, color, metal, rough, emis, ct, nt = mats[m]
            out.append("  " * depth + f"  prim vertices {len(v)} indices {len(i)} vhash {fnv(np.ascontiguousarray(v, np.float32).tobytes())} ihash {fnv(i.astype(np.uint16).tobytes())} material {m}")
            out.append("  " * depth + f"  material {mname} color {' '.join(g9(c) for c in color)} metallic {g9(metal)} roughness {g9(rough)} emissive {' '.join(g9(e) for e in emis)}"
                       f" ctex {texs[ct].shape[1]}x{texs[ct].shape[0]} {fnv(texs[ct].tobytes())} ntex {texs[nt].shape[1]}x{texs[nt].shape[0]} {fnv(texs[nt].tobytes())}")
        return out

    M32 = M.astype(np.float32)
    scale = [np.sqrt(np.float32(np.sum(M32[:3, c] * M32[:3, c], dtype=np.float32))) for c in range(3)]
    Rn = np.stack([M32[:3, c] / scale[c] for c in range(3)], axis=1)
    q = quat_from_matrix(Rn)
    lines = ["scene main nodes 3",
             f"node parent T {g9(0.5)} {g9(0.2)} {g9(0.0)} R {g9(0.0)} {g9(0.3826834)} {g9(0.0)} {g9(0.9238795)} S {g9(1.5)} {g9(0.7)} {g9(1.0)} mesh {meshes['sphere']} children 1",
             f"  node child T 0 1 0 R 0 0 0 1 S 0.5 0.5 0.5 mesh {meshes['soup']} children 0"] + prim_lines("soup", 1) + prim_lines("sphere", 0) + [
             f"node matrixNode T {g9(0.5)} {g9(-1.0)} {g9(2.0)} R {' '.join(g9(c) for c in q)} S {' '.join(g9(s) for s in scale)} mesh {meshes['wide']} children 0"] + prim_lines("wide", 0) + [
             "node bare T 0 0 0 R 0 0 0 1 S 1 1 1 mesh - children 0"]
    return b, lines


def run(tool, path):
    r = subprocess.run([str(tool), str(path)], capture_output=True, text=True)
    return r.returncode, r.stdout.splitlines(), r.stderr


@pytest.mark.parametrize("container", ["glb", "gltf_file", "gltf_data"])
def test_parse_matches_what_was_written(tool, tmp_path, container):
    b, want = build_asset()
    if container == "glb":
        path = tmp_path / "asset.glb"
        for fname, data in b.files.items():
            (tmp_path / fname).write_bytes(data)
        b.write_glb(path)
    else:
        path = tmp_path / "asset.gltf"
        b.write_gltf(path, buffer="file" if container == "gltf_file" else "data")
    rc, got, err = run(tool, path)
    assert rc == 0, err
    assert len(got) == len(want), "\n".join(got)
    for g, w in zip(got, want):
        assert g == w


def test_errors_are_reported(tool, tmp_path):
    b, _ = build_asset()
    path = tmp_path / "asset.glb"
    for fname, data in b.files.items():
        (tmp_path / fname).write_bytes(data)
    b.write_glb(path)
    raw = path.read_bytes()
    cases = {"not_gltf.glb": b"glTF" + raw[4:8].replace(b"\x02", b"\x07") + raw[8:], "truncated.glb": raw[:200], "garbage.gltf": b"{ \"asset\": ",
             "v1.gltf": b'{"asset": {"version": "1.0"}}'}
    for name, data in cases.items():
        (tmp_path / name).write_bytes(data)
        rc, _, err = run(tool, tmp_path / name)
        assert rc == 2 and "pbr::gltf" in err, name
    # a primitive without TANGENT: the reference's message (gltf/Asset.cpp:166)
    b2, _ = build_asset()
    del b2.doc["meshes"][0]["primitives"][0]["attributes"]["TANGENT"]
    for fname, data in b2.files.items():
        (tmp_path / fname).write_bytes(data)
    b2.write_glb(tmp_path / "no_tangent.glb")
    rc, _, err = run(tool, tmp_path / "no_tangent.glb")
    assert rc == 2 and "Primitive does not have TANGENT attribute" in err
    assert run(tool, tmp_path / "missing.glb")[0] == 2


def test_hostile_assets_are_rejected(tool, tmp_path):
    """Malformed files must fail in the parser, not in the rasteriser: an index past the primitive's vertices (K10 would read
    another draw's vertex-cache slots), an accessor whose count * stride wraps size_t, a node graph with a cycle."""
    import json as js
    import struct

    def variant(name, edit):
        b, _ = build_asset()
        edit(b)
        for fname, data in b.files.items():
            (tmp_path / fname).write_bytes(data)
        path = tmp_path / name
        b.write_gltf(path, buffer="file")
        return run(tool, path)

    def bad_index(b):  # patch the first index of the first indexed primitive to 65535
        doc = b.doc
        prim = next(p for m in doc["meshes"] for p in m["primitives"] if "indices" in p)
        acc = doc["accessors"][prim["indices"]]
        bv = doc["bufferViews"][acc["bufferView"]]
        off = bv.get("byteOffset", 0) + acc.get("byteOffset", 0)
        size = {5121: 1, 5123: 2, 5125: 4}[acc["componentType"]]
        b.bin[off:off + size] = b"\xff" * size

    rc, _, err = variant("bad_index.gltf", bad_index)
    assert rc == 2 and "addresses a vertex past" in err, err

    def huge_count(b):
        acc = b.doc["accessors"][0]
        acc["count"] = (1 << 64) // 12 + 2  # (count - 1) * stride wraps to a small number

    rc, _, err = variant("huge_count.gltf", huge_count)
    assert rc == 2 and "pbr::gltf" in err, err

    def cycle(b):
        nodes = b.doc["nodes"]
        parent = next(i for i, n in enumerate(nodes) if n.get("children"))
        child = nodes[parent]["children"][0]
        nodes[child].setdefault("children", []).append(parent)

    rc, _, err = variant("cycle.gltf", cycle)
    assert rc == 2 and ("cycle" in err or "more than one parent" in err), err


def test_bench_views_scene_imports(tool, tmp_path):
    """The .glb bench.py writes for `config4_256_views_gltf` (64 spheres sharing one set of accessors under 64 materials and node
    transforms + the ground plane) goes through the import path: 65 mesh nodes, one primitive each, 589,828 triangles."""
    import sys
    sys.path.insert(0, str(ROOT))
    import bench
    path = tmp_path / "views.glb"
    bench.write_views_glb(path)
    rc, got, err = run(tool, path)
    assert rc == 0, err
    assert got[0].split()[-1] == "65"
    prims = [line.split() for line in got if line.strip().startswith("prim ")]
    assert len(prims) == 65
    assert sum(int(p[p.index("indices") + 1]) // 3 for p in prims) == 589828
    assert len({p[p.index("vhash") + 1] for p in prims}) == 2  # the spheres share their vertex data, the ground has its own
    assert len({line.split()[1] for line in got if line.strip().startswith("material ")}) == 65
