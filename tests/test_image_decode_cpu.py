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


def jpeg_cases():
    """(name, jpeg bytes) written by Pillow and OpenCV: baseline and progressive, grey / 4:4:4 / 4:2:2 / 4:2:0 / 4:1:1 /
    4:4:0 sampling, odd sizes (partial MCUs, 1-pixel images), restart intervals, optimised Huffman tables, 16-bit-free
    quality sweep, Adobe RGB / CMYK / YCCK files."""
    import io

    import cv2
    from PIL import Image
    rng = np.random.default_rng(21)
    out = []

    def content(w, h, kind):
        y, x = np.mgrid[0:h, 0:w]
        if kind == "smooth":
            img = np.stack([(x * 3 + y) % 256, (x + 2 * y) % 256, (x * y // 7) % 256], axis=-1)
        elif kind == "noise":
            img = rng.integers(0, 256, (h, w, 3))
        else:  # edges: hard colour steps stress chroma upsampling and clamping
            img = np.stack([((x // 5 + y // 3) % 2) * 255, ((x // 2) % 2) * 255, ((y // 4) % 2) * 200 + 30], axis=-1)
        return img.astype(np.uint8)

    def pil(img, **kw):
        buf = io.BytesIO()
        Image.fromarray(img).save(buf, "JPEG", **kw)
        return buf.getvalue()

    sizes = ((1, 1), (7, 5), (16, 16), (17, 33), (40, 23), (65, 49))
    for (w, h) in sizes:
        for kind in ("smooth", "noise", "edges"):
            img = content(w, h, kind)
            for sub, name in ((0, "444"), (1, "422"), (2, "420")):
                for prog in (False, True):
                    q = int(rng.choice([30, 75, 95, 100]))
                    out.append((f"pil_{kind}_{w}x{h}_{name}_{'p' if prog else 'b'}_q{q}", pil(img, quality=q, subsampling=sub, progressive=prog, optimize=bool(rng.integers(0, 2)))))
            out.append((f"pil_{kind}_{w}x{h}_grey", pil(img[..., 0], quality=85)))
            out.append((f"pil_{kind}_{w}x{h}_grey_p", pil(img[..., 0], quality=60, progressive=True)))
    big = content(200, 150, "smooth") // 2 + content(200, 150, "noise") // 2
    for prog in (0, 1):
        for rst in (0, 1, 7):
            for samp, name in ((cv2.IMWRITE_JPEG_SAMPLING_FACTOR_411, "411"), (cv2.IMWRITE_JPEG_SAMPLING_FACTOR_440, "440"), (cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420, "420")):
                ok, enc = cv2.imencode(".jpg", big, [cv2.IMWRITE_JPEG_QUALITY, 80, cv2.IMWRITE_JPEG_PROGRESSIVE, prog, cv2.IMWRITE_JPEG_RST_INTERVAL, rst,
                                                     cv2.IMWRITE_JPEG_SAMPLING_FACTOR, samp, cv2.IMWRITE_JPEG_OPTIMIZE, 1])
                assert ok
                out.append((f"cv_{name}_{'p' if prog else 'b'}_rst{rst}", enc.tobytes()))
    # Adobe-flavoured files: RGB stored without colour transform, CMYK, YCCK
    rgb = content(33, 21, "edges")
    out.append(("pil_rgb_keep", pil(rgb, quality=90, keep_rgb=True)))
    cmyk = Image.fromarray(rng.integers(0, 256, (19, 27, 4)).astype(np.uint8), "CMYK")
    buf = io.BytesIO(); cmyk.save(buf, "JPEG", quality=90); out.append(("pil_cmyk", buf.getvalue()))
    return out


def jpeg_cases_or_skip():
    try:
        return jpeg_cases()
    except ImportError as e:  # Pillow / OpenCV missing on this machine
        pytest.skip(f"no JPEG encoder: {e}")


@pytest.mark.skipif(not STB.exists(), reason="oracle/_ref/libpbr_ref_stb.so not built (needs /root/reference)")
def test_matches_reference_stb_on_the_feature_matrix(tool, tmp_path):
    assert len(CASES) > 100
    for name, png in CASES:
        want = stb_decode(png)
        got, err = ours(tool, png, tmp_path)
        assert want is not None, name
        assert got is not None, f"{name}: {err}"
        assert got.shape == want.shape and np.array_equal(got, want), name


@pytest.mark.skipif(not STB.exists(), reason="oracle/_ref/libpbr_ref_stb.so not built (needs /root/reference)")
def test_jpeg_matches_reference_stb(tool, tmp_path):
    cases_ = jpeg_cases_or_skip()
    assert len(cases_) > 150
    for name, jpg in cases_:
        want = stb_decode(jpg)
        got, err = ours(tool, jpg, tmp_path)
        assert want is not None, name
        assert got is not None, f"{name}: {err}"
        assert got.shape == want.shape, name
        assert np.array_equal(got, want), f"{name}: {np.count_nonzero(got != want)} of {got.size} bytes differ (max {np.abs(got.astype(int) - want).max()})"


def test_matches_committed_golden_files(tool, tmp_path):
    files = sorted(GOLD.glob("*.png")) + sorted(GOLD.glob("*.jpg"))
    assert len(files) >= 10, "tests/golden/png is missing: run tests/golden/make_png_golden.py"
    assert any(f.suffix == ".jpg" for f in files)
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
