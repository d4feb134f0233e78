"""CPU tests of the C-ABI boundary: the library loads, exports every symbol include/pbrcu.h declares,
and fails LOUDLY (status 2, no fallback) when asked to compute without a CUDA device."""
import ctypes as C
import re
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def test_library_exports_every_declared_symbol(pbr):
    lib = pbr.load()
    out = subprocess.check_output(["nm", "-D", "--defined-only", str(pbr.LIB_PATH)], text=True)
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    declared = pbr.declared_symbols()
    assert len(declared) >= 40
    missing = [s for s in declared if s not in exported]
    assert not missing, f"declared in pbrcu.h but not exported: {missing}"
    assert set(declared) == set(pbr._SIGNATURES), "ctypes signatures out of sync with the header"
    # nothing but the C ABI leaks out of the library
    leaked = [s for s in exported if not s.startswith("pbrcu_")]
    assert not leaked, f"non-ABI symbols exported: {leaked[:5]}"
    assert lib.version() == 100


def test_header_is_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "pbrcu.h"\nint main(void){ pbrcu_rect r = {0,0,0,0}; (void)r; return PBRCU_OK; }\n')
    subprocess.run(["/usr/bin/gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", str(ROOT / "include"), "-c", str(src), "-o", str(tmp_path / "t.o")], check=True)


def test_header_cites_reference_interfaces():
    text = (ROOT / "include" / "pbrcu.h").read_text()
    for cite in ["PbrRenderSystem.cpp:466-501", "TonemapperSystem.cpp:97-134", "lighting.glsl:19-29", "aces+gamma.glsl:30-40", "GBuffer.hpp:13-16"]:
        assert cite in text


def test_product_never_touches_the_oracle():
    """The product path (library sources, python binding, facade) must not include, import, link or call
    anything under oracle/ (comments may cite the oracle function that defines a kernel's semantics)."""
    pkg = ROOT / "physically-based-renderer_b200"
    offenders = []
    for p in list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")) + list(pkg.rglob("*.py")) + list(pkg.rglob("*.cpp")) + list(pkg.rglob("*.hpp")) + list((ROOT / "include").rglob("*")):
        if p.is_file() and re.search(r"#\s*include[^\n]*oracle|pbro_\w+\s*\(|(CDLL|dlopen)[^\n]*oracle|import oracle|from oracle", p.read_text(errors="ignore")):
            offenders.append(str(p.relative_to(ROOT)))
    assert not offenders, offenders
    assert "oracle" not in subprocess.check_output(["ldd", str(pkg / "lib" / "libpbrcu.so")], text=True)


def test_no_gpu_means_loud_failure(pbr):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pbr.PbrcuError) as e:
        pbr.Device(0)
    assert e.value.code == 2 and "CUDA error" in str(e.value)


def test_argument_validation_without_gpu(pbr):
    lib = pbr.load()
    # NULL arguments are rejected before any CUDA call: status 1 + a message, never a crash
    assert lib.cdll.pbrcu_lighting_blinnphong(None, None, None, pbr.WHOLE, None, 0, None) == 1
    assert "NULL" in lib.last_error()
    assert lib.cdll.pbrcu_ctx_create(0, None) == 1
    assert lib.cube_chain_texels(2048, 12) == 33554430  # texels of the full chain
    assert lib.cube_mip_offset(2048, 1) == 6 * 2048 * 2048
