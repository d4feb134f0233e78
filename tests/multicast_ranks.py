"""Worker of tests/test_gpu_multicast.py (one rank per GPU under torchrun; not a test module): every rank shades its row band of one
frame with pbrcu_shade_ggx_ibl_replicated (P2P stores into the peers' frames) and with pbrcu_shade_ggx_ibl_multicast (one multimem.st
per texel to the NVSwitch multicast address), frames in torch symmetric memory; both must leave the WHOLE frame on every rank,
bit-identical to the frame shaded unsharded.  Prints one line per rank: 'rank r ok' or the mismatch."""
import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import pbr_b200 as pbr  # noqa: E402
from pbr_b200.partition import PeerFrames, band_views, row_band  # noqa: E402
from pbr_b200.synth import synth_gbuffer, synth_lights  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = pbr.Device(local)
    lib = dev.lib
    w, h, nl = 160, 96, 24
    g = synth_gbuffer(w, h, seed=77, tier_b=True)
    gb = dev.gbuffer(g["pos"], g["nrm"], g["alb"], orm=g["orm"], emissive=g["emi"])
    results = []
    for tiled in (False, True):
        lights = synth_lights(nl if tiled else 4, seed=5, ranged=tiled)
        n = len(lights)
        ld = dev.upload(lights)
        tiles = None
        if tiled:
            tiles = pbr.LightTiles()
            lib.light_tiles_alloc(dev.ctx, w, h, n, C.byref(tiles))
            lib.cluster_lights(dev.ctx, C.byref(gb), ld.ptr, n, C.byref(tiles), None)
        cam = pbr.f3(0.5, -1.0, 12.0)
        whole = dev.alloc(w * h * 8)
        lib.shade_ggx_ibl(dev.ctx, C.byref(gb), cam, ld.ptr, n, C.byref(tiles) if tiled else None, None, pbr.WHOLE, whole.ptr, w * 8, None)
        dev.sync()
        want = whole.download((h, w, 4), np.uint16)
        y0, y1 = row_band(h, world, rank, align=16)
        band = band_views(pbr, gb, tiles, y0, y1)
        pf = PeerFrames(w * h * 8, torch.device("cuda", local), dist)
        stream = torch.cuda.current_stream().cuda_stream
        for mode in ("replicated", "multicast"):
            if mode == "multicast" and not pf.multicast:
                results.append(f"multicast unavailable (tiled={tiled})")
                continue
            pf.local.fill_(0x5A)
            torch.cuda.synchronize()
            dist.barrier()
            if y1 > y0:
                if mode == "replicated":
                    reps, nrep = pf.replica_array(y0 * w * 8)
                    lib.shade_ggx_ibl_replicated(dev.ctx, C.byref(band.gbuffer), cam, ld.ptr, n, C.byref(band.tiles) if tiled else None, None, pbr.WHOLE,
                                                 pf.local.data_ptr() + y0 * w * 8, w * 8, reps, nrep, stream)
                else:
                    lib.shade_ggx_ibl_multicast(dev.ctx, C.byref(band.gbuffer), cam, ld.ptr, n, C.byref(band.tiles) if tiled else None, None, pbr.WHOLE,
                                                pf.multicast + y0 * w * 8, w * 8, stream)
            pf.barrier()
            torch.cuda.synchronize()
            dist.barrier()
            got = pf.local.cpu().numpy().view(np.uint16).reshape(h, w, 4)
            if not np.array_equal(got, want):
                results.append(f"{mode} (tiled={tiled}): {np.count_nonzero(got != want)} values differ from the unsharded frame")
        del pf
    print(f"rank {rank} " + ("ok" if not [r for r in results if "differ" in r] else "FAIL") + (" | " + "; ".join(results) if results else ""), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
