"""CPU: the import path's PNG decoder (include/pbr/image/Decode.hpp, the stand-in for the reference's
stbi_load_from_memory(..., 4) in image/LoadImage.cpp:58-73) against
  * the REFERENCE'S OWN stb_image build (oracle/_ref/libpbr_ref_stb.so, compiled from /root/reference where it lies), and
  * committed golden files (tests/golden/png/*.png + .rgba made by tests/golden/make_png_golden.py with that build),
on every colour type / bit depth / interlace / filter / deflate-block combination tests/pngwriter.py can write."""
import ctypes as C
import subprocess
import zlib
from pathlib import Path

import numpy as np
import pytest
from pngwriter import CHANNELS, encode

ROOT = Path(__file__).resolve().parent.parent
TOOL = ROOT / "tests" / "cpp" / "build" / "image_decode"
STB = ROOT / "oracle" / "_ref" / "libpbr_ref_stb.so"
GOLD = ROOT / "tests" / "golden" / "png"


@pytest.fixture(scope="module")
def tool():
    subprocess.run(["make", "-s", "-C", str(ROOT / "tests" / "cpp"), "cpu"], check=True)
    return TOOL


def ours(tool, png: bytes, tmp_path):
    src, dst = tmp_path / "in.png", tmp_path / "out.rgba"
    src.write_bytes(png)
    r = subprocess.run([str(tool), str(src), str(dst)], capture_output=True, text=True)
    if r.returncode != 0:
        return None, r.stderr.strip()
    w, h = map(int, r.stdout.split())
    return np.frombuffer(dst.read_bytes(), np.uint8).reshape(h, w, 4), ""


def stb_decode(png: bytes):
    lib = C.CDLL(str(STB))
    lib.stbi_load_from_memory.restype = C.POINTER(C.c_ubyte)
    lib.stbi_load_from_memory.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
    lib.stbi_image_free.argtypes = [C.c_void_p]
    w, h, n = C.c_int(), C.c_int(), C.c_int()
    p = lib.stbi_load_from_memory(png, len(png), C.byref(w), C.byref(h), C.byref(n), 4)
    if not p:
        return None
    out = np.ctypeslib.as_array(p, (h.value, w.value, 4)).copy()
    lib.stbi_image_free(p)
    return out


def cases():
    """(name, png bytes): a sweep over the PNG feature matrix with seeded content."""
    rng = np.random.default_rng(11)
    out = []
    for color, depths in ((0, (1, 2, 4, 8, 16)), (2, (8, 16)), (3, (1, 2, 4, 8)), (4, (8, 16)), (6, (8, 16))):
        for depth in depths:
            for interlace in (False, True):
                for (w, h) in ((1, 1), (5, 3), (13, 9), (32, 17)):
                    s = rng.integers(0, 1 << depth, (h, w, CHANNELS[color]))
                    kw = dict(interlace=interlace, filters=tuple(rng.integers(0, 5, 4)), level=int(rng.choice([0, 1, 6, 9])))
                    if color == 3:
                        kw["palette"] = rng.integers(0, 256, (1 << depth, 3))
                        if rng.random() < 0.5:
                            kw["trns"] = rng.integers(0, 256, int(rng.integers(1, (1 << depth) + 1))).astype(np.uint8).tobytes()
                    elif color in (0, 2) and rng.random() < 0.5:
                        key = s[0, 0]  # a colour key that really occurs
                        kw["trns"] = b"".join(int(v).to_bytes(2, "big") for v in key)
                    if rng.random() < 0.3:
                        kw["strategy"] = zlib.Z_FIXED
                    if rng.random() < 0.3:
                        kw["idat_split"] = 7
                    if rng.random() < 0.3:
                        kw["extra_chunks"] = [(b"gAMA", (45455).to_bytes(4, "big")), (b"tEXt", b"Comment\0synthetic")]
                    out.append((f"c{color}d{depth}{'i' if interlace else 'n'}_{w}x{h}", encode(s, color, depth, **kw)))
    # larger, compressible content: long matches, distances near 32 KiB, dynamic blocks with long codes
    y, x = np.mgrid[0:300, 0:257]
    tex = np.stack([(x * 7 + y) % 256, (x // 8 * 32) % 256, (y * 3) % 256, 255 - (x % 64) * 4], axis=-1)
    out.append(("big_rgba8", encode(tex, 6, 8, filters=(4, 1, 2, 3, 0), level=9)))
    out.append(("big_rgba8_stored", encode(tex, 6, 8, level=0)))
    out.append(("big_noise_rgb16", encode(rng.integers(0, 65536, (64, 200, 3)), 2, 16, filters=(4,), level=6)))
    return out


CASES = cases()


@pytest.mark.skipif(not STB.exists(), reason="oracle/_ref/libpbr_ref_stb.so not built (needs /root/reference)")
def test_matches_reference_stb_on_the_feature_matrix(tool, tmp_path):
    assert len(CASES) > 100
    for name, png in CASES:
        want = stb_decode(png)
        got, err = ours(tool, png, tmp_path)
        assert want is not None, name
        assert got is not None, f"{name}: {err}"
        assert got.shape == want.shape and np.array_equal(got, want), name


def test_matches_committed_golden_files(tool, tmp_path):
    files = sorted(GOLD.glob("*.png"))
    assert len(files) >= 10, "tests/golden/png is missing: run tests/golden/make_png_golden.py"
    for f in files:
        want = np.frombuffer(f.with_suffix(".rgba").read_bytes(), np.uint8)
        got, err = ours(tool, f.read_bytes(), tmp_path)
        assert got is not None, f"{f.name}: {err}"
        assert np.array_equal(got.reshape(-1), want), f.name


def test_rejects_what_it_does_not_decode(tool, tmp_path):
    good = CASES[0][1]
    for bad, what in ((b"\xff\xd8\xff\xe0" + b"\0" * 64, "not a PNG"), (good[:40], "truncated"), (good[:8] + good[8 + 25:], "IHDR")):
        got, err = ours(tool, bad, tmp_path)
        assert got is None and "pbr::image::decode" in err, what
