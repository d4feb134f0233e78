"""GPU, two or more devices: the fused gathers of K3 over torch symmetric memory -- P2P stores (pbrcu_shade_ggx_ibl_replicated) and
NVSwitch multicast stores (pbrcu_shade_ggx_ibl_multicast) -- leave the whole frame on every rank, bit-identical to the unsharded
frame.  Spawns one rank per GPU (tests/multicast_ranks.py under torch.distributed.run); skipped on a one-GPU box, where
tests/test_gpu_mg.py and test_shade_replicated_stores_equal_the_plain_frame cover the store path."""
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def test_replicated_and_multicast_stores_across_ranks():
    import ctypes as C

    import pbr_b200 as pbr
    count = C.c_int()
    pbr.load().device_count(C.byref(count))  # (torch stays out of this process: the ranks import it)
    n = count.value
    if n < 2:
        pytest.skip("needs two GPUs")
    n = min(n, 4)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1", "--master-port", "29533",
                        str(ROOT / "tests" / "multicast_ranks.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("rank ")]
    assert len(lines) == n and all(" ok" in ln for ln in lines), "\n".join(lines)
    print("\n".join(lines))
