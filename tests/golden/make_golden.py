"""Regenerates tests/golden/tier_a_reference.npz from the REFERENCE'S OWN shaders compiled for the CPU
(oracle/_ref/libpbr_ref_shaders.so, built by oracle/Makefile from /root/reference/assets/shaders).

    python tests/golden/make_golden.py        # needs /root/reference (the CPU container)

The vectors pin the oracle (and through it the CUDA kernels) to the reference on machines where
/root/reference is not mounted (the GPU box)."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT / "tests"))
from oracle import RefShaders, synth_gbuffer  # noqa: E402


def main():
    r = RefShaders()
    out = {}
    # lighting.glsl::main over seeded G-buffers, several cameras (incl. the viewer's start state, the origin)
    for i, (w, h, cam) in enumerate([(48, 20, (0.0, 0.0, 0.0)), (33, 17, (1.5, -2.0, 12.0)), (16, 16, (-9.0, 4.0, -3.0))]):
        g = synth_gbuffer(w, h, seed=1000 + i)
        f32, f16 = r.lighting(g["pos"], g["nrm"], g["alb"], cam)
        out[f"light{i}_pos"] = g["pos"].view(np.uint16)
        out[f"light{i}_nrm"] = g["nrm"].view(np.uint16)
        out[f"light{i}_alb"] = g["alb"]
        out[f"light{i}_cam"] = np.asarray(cam, np.float32)
        out[f"light{i}_out_f32"] = f32
        out[f"light{i}_out_f16"] = f16
    # aces+gamma.glsl::main over an HDR ramp + random HDR values + exact blacks
    rng = np.random.default_rng(7)
    hdr = np.zeros((24, 40, 4), np.float16)
    hdr[:8] = np.linspace(0, 4, 8 * 40).reshape(8, 40, 1)
    hdr[8:] = (rng.random((16, 40, 4)) ** 3 * 3).astype(np.float16)
    hdr[12, :8] = 0
    hdr[..., 3] = rng.random((24, 40)) * 1.5
    out["tm_hdr"] = hdr.view(np.uint16)
    out["tm_ldr"] = r.tonemap(hdr.view(np.uint16))
    # library functions: blinnPhongLighting, gammaCorrect, acesTonemap
    n = 64
    frag = rng.random((n, 4)).astype(np.float32)
    N = rng.standard_normal((n, 3)).astype(np.float32); N /= np.linalg.norm(N, axis=1, keepdims=True)
    V = rng.standard_normal((n, 3)).astype(np.float32); V /= np.linalg.norm(V, axis=1, keepdims=True)
    L = rng.standard_normal((n, 3)).astype(np.float32); L /= np.linalg.norm(L, axis=1, keepdims=True)
    out["bp_frag"], out["bp_N"], out["bp_V"], out["bp_L"] = frag, N, V, L
    out["bp_out"] = np.stack([r.blinn_phong(frag[i], N[i], V[i], L[i]) for i in range(n)])
    rgb = (rng.random((n, 3)) * 2).astype(np.float32)
    out["tm_rgb"] = rgb
    out["tm_aces"] = np.stack([r.aces(c) for c in rgb])
    out["tm_gamma"] = np.stack([r.gamma_correct(c, 2.2) for c in rgb])
    out["fsv"] = np.stack([r.fullscreen_vertex(i) for i in range(3)])
    np.savez_compressed(ROOT / "tests" / "golden" / "tier_a_reference.npz", **out)
    print("wrote tier_a_reference.npz", sum(v.nbytes for v in out.values()), "bytes raw")


if __name__ == "__main__":
    main()
