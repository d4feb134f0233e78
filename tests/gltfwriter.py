"""A small glTF 2.0 writer for tests: builds .glb (or .gltf + external .bin / data: URIs) files from numpy mesh data,
with the layout choices the import path has to cope with -- interleaved or separate vertex buffers, u8 / u16 / u32
indices or none, normalised integer texture coordinates, node TRS or matrix, embedded (bufferView) or external images.
Written from the glTF 2.0 specification."""
import base64
import json
import struct
from pathlib import Path

import numpy as np
from pngwriter import encode_rgba8

FLOAT, UBYTE, USHORT, UINT, SHORT = 5126, 5121, 5123, 5125, 5122


class Builder:
    def __init__(self):
        self.bin = bytearray()
        self.doc = {"asset": {"version": "2.0", "generator": "tests/gltfwriter.py"}, "buffers": [], "bufferViews": [], "accessors": [],
                    "images": [], "samplers": [{"name": "default"}], "textures": [], "materials": [], "meshes": [], "nodes": [], "scenes": [], "scene": 0}
        self.files = {}  # external files to write next to a .gltf

    def _view(self, data: bytes, stride=None) -> int:
        while len(self.bin) % 4:
            self.bin.append(0)
        view = {"buffer": 0, "byteOffset": len(self.bin), "byteLength": len(data)}
        if stride:
            view["byteStride"] = stride
        self.bin += data
        self.doc["bufferViews"].append(view)
        return len(self.doc["bufferViews"]) - 1

    def accessor(self, array: np.ndarray, ctype: int, kind: str, normalized=False, view=None, offset=0, count=None) -> int:
        if view is None:
            view = self._view(np.ascontiguousarray(array).tobytes())
        acc = {"bufferView": view, "byteOffset": offset, "componentType": ctype, "type": kind, "count": int(len(array) if count is None else count)}
        if normalized:
            acc["normalized"] = True
        self.doc["accessors"].append(acc)
        return len(self.doc["accessors"]) - 1

    def image(self, rgba, name: str, how="view", encoded: bytes = None, mime="image/png", **png_kw) -> int:
        """rgba: pixels to store as PNG -- or pass `encoded` (+ mime), an already encoded PNG / JPEG file."""
        png = encoded if encoded is not None else encode_rgba8(rgba, **png_kw)
        if how == "view":
            entry = {"name": name, "bufferView": self._view(png), "mimeType": mime}
        elif how == "data":
            entry = {"name": name, "uri": "data:image/png;base64," + base64.b64encode(png).decode()}
        else:  # external file
            self.files[f"{name}.png"] = png
            entry = {"name": name, "uri": f"{name}.png"}
        self.doc["images"].append(entry)
        self.doc["textures"].append({"sampler": 0, "source": len(self.doc["images"]) - 1})
        return len(self.doc["textures"]) - 1

    def material(self, name, color, color_tex: int, normal_tex: int, metallic=1.0, roughness=1.0, emissive=(0, 0, 0)) -> int:
        self.doc["materials"].append({"name": name, "pbrMetallicRoughness": {"baseColorFactor": [float(c) for c in color], "baseColorTexture": {"index": color_tex},
                                                                            "metallicFactor": metallic, "roughnessFactor": roughness},
                                      "normalTexture": {"index": normal_tex}, "emissiveFactor": [float(e) for e in emissive]})
        return len(self.doc["materials"]) - 1

    def primitive(self, vertices: np.ndarray, indices, material: int, layout="separate", index_type=USHORT, uv_quantised=False) -> dict:
        """vertices: (n, 12) MeshVertex rows (position 3, normal 3, tangent 4, uv 2)."""
        v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 12)
        if layout == "interleaved":  # one bufferView with byteStride 48
            view = self._view(v.tobytes(), stride=48)
            attrs = {"POSITION": self.accessor(v, FLOAT, "VEC3", view=view, offset=0), "NORMAL": self.accessor(v, FLOAT, "VEC3", view=view, offset=12),
                     "TANGENT": self.accessor(v, FLOAT, "VEC4", view=view, offset=24), "TEXCOORD_0": self.accessor(v, FLOAT, "VEC2", view=view, offset=40)}
        else:
            attrs = {"POSITION": self.accessor(v[:, 0:3], FLOAT, "VEC3"), "NORMAL": self.accessor(v[:, 3:6], FLOAT, "VEC3"),
                     "TANGENT": self.accessor(v[:, 6:10], FLOAT, "VEC4")}
            if uv_quantised:  # normalised unsigned shorts (uv in [0, 1])
                attrs["TEXCOORD_0"] = self.accessor(np.round(v[:, 10:12] * 65535).astype(np.uint16), USHORT, "VEC2", normalized=True)
            else:
                attrs["TEXCOORD_0"] = self.accessor(v[:, 10:12], FLOAT, "VEC2")
        prim = {"attributes": attrs, "material": material}
        if indices is not None:
            dt = {UBYTE: np.uint8, USHORT: np.uint16, UINT: np.uint32}[index_type]
            prim["indices"] = self.accessor(np.asarray(indices).astype(dt), index_type, "SCALAR")
        return prim

    def mesh(self, name, primitives) -> int:
        self.doc["meshes"].append({"name": name, "primitives": primitives})
        return len(self.doc["meshes"]) - 1

    def node(self, name, mesh=None, translation=None, rotation_xyzw=None, scale=None, matrix=None, children=()) -> int:
        n = {"name": name}
        if mesh is not None:
            n["mesh"] = mesh
        if matrix is not None:
            n["matrix"] = [float(x) for x in np.asarray(matrix, np.float64).reshape(4, 4).T.reshape(-1)]  # column-major
        else:
            if translation is not None:
                n["translation"] = [float(x) for x in translation]
            if rotation_xyzw is not None:
                n["rotation"] = [float(x) for x in rotation_xyzw]
            if scale is not None:
                n["scale"] = [float(x) for x in scale]
        if children:
            n["children"] = list(children)
        self.doc["nodes"].append(n)
        return len(self.doc["nodes"]) - 1

    def scene(self, nodes, name="scene"):
        self.doc["scenes"].append({"name": name, "nodes": list(nodes)})

    def write_glb(self, path: Path):
        doc = dict(self.doc)
        doc["buffers"] = [{"byteLength": len(self.bin)}]
        js = json.dumps(doc).encode()
        js += b" " * ((-len(js)) % 4)
        binary = bytes(self.bin) + b"\0" * ((-len(self.bin)) % 4)
        total = 12 + 8 + len(js) + 8 + len(binary)
        Path(path).write_bytes(struct.pack("<III", 0x46546C67, 2, total) + struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(binary), 0x004E4942) + binary)

    def write_gltf(self, path: Path, buffer="file"):
        path = Path(path)
        doc = dict(self.doc)
        if buffer == "data":
            doc["buffers"] = [{"byteLength": len(self.bin), "uri": "data:application/octet-stream;base64," + base64.b64encode(bytes(self.bin)).decode()}]
        else:
            name = path.stem + " data.bin"  # a space: the uri is percent-encoded
            (path.parent / name).write_bytes(bytes(self.bin))
            doc["buffers"] = [{"byteLength": len(self.bin), "uri": name.replace(" ", "%20")}]
        for fname, data in self.files.items():
            (path.parent / fname).write_bytes(data)
        path.write_text(json.dumps(doc, indent=1))
