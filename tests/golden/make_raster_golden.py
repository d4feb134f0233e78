"""Generates tests/golden/raster_golden.npz: one 160x90 frame of the seeded raster_scene through the CPU oracle
of the geometry pass (oracle/pbr_oracle_raster.c), whose vertex / fragment stages are checked bit-for-bit
against the reference's GLSL (tests/test_oracle_raster.py).  Run from the repo root: python tests/golden/make_raster_golden.py"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT / "tests"))
from oracle import Oracle, col_major, look_at_rh, perspective_rh_no, raster_scene  # noqa: E402

view, proj = col_major(look_at_rh((0, -1, -10), (0, 0, 0))), col_major(perspective_rh_no(np.pi / 2, 16 / 9))
out = Oracle().geometry_pass(raster_scene(1), view, proj, 160, 90, extended=True)
np.savez_compressed(ROOT / "tests" / "golden" / "raster_golden.npz", **out)
print({k: v.shape for k, v in out.items()}, "covered", float((out["prim"] != 0xFFFFFFFF).mean()))
