import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def _bundled_nccl():
    """PyTorch's own libnccl (found without importing torch): the C multi-GPU host must use the copy a later `import torch` in the
    same process will need, not the system's older one under the same SONAME (include/pbrcu.h, PBRCU_MG_GATHER_NCCL)."""
    import importlib.util
    import os
    try:
        spec = importlib.util.find_spec("nvidia.nccl")
    except (ImportError, ValueError):
        spec = None
    for base in (spec.submodule_search_locations if spec and spec.submodule_search_locations else []):
        lib = Path(base) / "lib" / "libnccl.so.2"
        if lib.exists():
            os.environ.setdefault("PBRCU_NCCL_LIB", str(lib))
            return


def pytest_configure(config):
    _bundled_nccl()
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import Oracle

    return Oracle()


@pytest.fixture(scope="session")
def ref_shaders():
    from oracle import RefShaders

    try:
        return RefShaders()
    except FileNotFoundError as e:  # neither /root/reference nor a prebuilt oracle/_ref
        pytest.skip(str(e))


@pytest.fixture(scope="session")
def pbr():
    import pbr_b200

    return pbr_b200


@pytest.fixture(scope="session")
def dev(pbr):
    """One pbrcu context on cuda:0.  No fallback: without a GPU this raises and the test fails."""
    d = pbr.Device(0)
    yield d
    d.close()


def f16(bits: np.ndarray) -> np.ndarray:
    return bits.view(np.float16).astype(np.float32)


def psnr(a: np.ndarray, b: np.ndarray, peak: float) -> float:
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return float("inf") if mse == 0 else 10.0 * np.log10(peak * peak / mse)


def assert_hdr_parity(gpu_bits, oracle_bits, peak, abs_tol=1e-3, rel_tol=0.0, min_psnr=60.0, what=""):
    """north_star tolerance in linear HDR on the fp16-decoded values: max |err| <= abs_tol (+ rel_tol*|ref|
    where the value range exceeds fp16's resolution at abs_tol) and PSNR >= min_psnr dB."""
    g, o = f16(gpu_bits), f16(oracle_bits)
    assert np.isfinite(g).all(), f"{what}: non-finite GPU output"
    err = np.abs(g - o)
    bound = abs_tol + rel_tol * np.abs(o)
    worst = float((err - bound).max())
    p = psnr(g, o, peak)
    assert worst <= 0, f"{what}: max abs err {float(err.max()):.3e} exceeds tolerance (worst excess {worst:.3e})"
    assert p >= min_psnr, f"{what}: PSNR {p:.1f} dB < {min_psnr}"
    return float(err.max()), p
