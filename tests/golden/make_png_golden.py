"""Writes tests/golden/png/*.png and, beside each, the RGBA8 pixels the REFERENCE'S stb_image build
(oracle/_ref/libpbr_ref_stb.so, compiled from /root/reference/src/pbr_engine/image/stb) decodes it to.
Run in the build container (needs /root/reference for `make -C oracle ref`)."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT / "tests"))
from test_image_decode_cpu import CASES, jpeg_cases, stb_decode  # noqa: E402

out = ROOT / "tests" / "golden" / "png"
out.mkdir(parents=True, exist_ok=True)
keep = [c for i, c in enumerate(CASES) if c[0].endswith("13x9") or c[0] == "big_rgba8"]
for name, png in keep:
    pix = stb_decode(png)
    assert pix is not None, name
    (out / f"{name}.png").write_bytes(png)
    (out / f"{name}.rgba").write_bytes(pix.tobytes())
n_png = len(keep)
jk = [c for c in jpeg_cases() if "_40x23_" in c[0] or c[0] in ("cv_411_b_rst7", "cv_440_p_rst1") or c[0] in ("pil_rgb_keep", "pil_cmyk")]
for name, jpg in jk:
    pix = stb_decode(jpg)
    assert pix is not None, name
    (out / f"{name}.jpg").write_bytes(jpg)
    (out / f"{name}.rgba").write_bytes(pix.tobytes())
keep = keep + jk
print(len(keep), "golden files,", sum(f.stat().st_size for f in out.iterdir()), "bytes")
