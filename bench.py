#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native shading hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Metric (BASELINE.json): shaded Mpix/s at 4K GGX+IBL.  One "step" = one pass of the hot path over one
batch = K3 (`pbrcu_shade_ggx_ibl`: GGX/Smith/Schlick + 4 point lights + occlusion/emissive + split-sum
IBL) over one 3840x2160 synthetic G-buffer (BASELINE config[1]).  `value` is measured with the inputs
resident in HBM (CUDA events on the launching stream, max over ranks); `e2e` is the same metric through
the C-ABI call with HOST buffers (`pbrcu_shade_frame_host`: H2D of the five G-buffer planes, shade, D2H
of the RGBA16F result, all inside the timed region).

Multi-GPU (torchrun, one rank per GPU): frames (views) are independent, so every rank shades its own
frames with no data-path collective ("weak" scaling); torch.distributed is used for the barrier and the
max-over-ranks time only.

`--impl reference` times the CPU side on the box's host cores.  The reference has NO GGX/IBL code
(SURVEY.md section 0), so for this metric the CPU arm is the oracle port (oracle/pbr_oracle_ibl.c); the
Tier-A figure beside it (`tier_a`) is the reference's own GLSL compiled for the CPU (oracle/_ref).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

W4K, H4K = 3840, 2160
N_LIGHTS = 4
B_PER_PX_K3 = 44   # 36 read (pos 8 + nrm 8 + albedo 8 + emissive 8 + orm 4) + 8 written
B_PER_PX_K1 = 32   # 24 read + 8 written
# dram__bytes_read.sum + dram__bytes_write.sum of one shade_ggx_cl_kernel launch on this workload, from the committed
# `ncu --set full` capture (profiles/r02_k3_cl_ncu.txt): 520.4 MB + 60.5 MB; the 216 MB above the
# algorithmic 365 MB are the IBL footprint records / textures and write-allocate traffic
K3_DRAM_TRAFFIC_BYTES = 580.9e6
K3_KERNEL = "shade_ggx_cl_kernel<4 lights, records + texture-unit ambient>"
ENV_W, ENV_H = 4096, 2048
CUBE_RES, CUBE_MIPS, IBL_SAMPLES, IRR_RES, LUT_RES = 2048, 12, 1024, 64, 512
CAM = (0.0, -1.0, -12.0)
WORKLOAD = ("3840x2160 synthetic G-buffer (gbuf_spheres), GGX metallic-roughness + 4 point lights + precomputed IBL "
            "(2048^2 12-mip prefiltered cube, 64^2 irradiance, 512^2 LUT), 1 frame per step per GPU")
SEED = 0x5EED0002


def config_block(n_gpus):
    """`config` of the JSON line, identical in both arms (--impl ours / reference)."""
    return {"workload": WORKLOAD, "l2": "inputs larger than L2 (365 MB per frame) and 3 rotating buffer sets", "parallelism": f"frames x{n_gpus}"}


def roof(bound, achieved, peak, unit, note=None):
    """One roofline entry: achieved / peak of whatever bounds the kernel."""
    r = {"bound": bound, "achieved": round(achieved, 2), "peak": round(peak, 2), "unit": unit, "frac": round(achieved / peak, 4) if peak else None}
    if note:
        r["note"] = note
    return r


def fp32_peak_tflops(sm_count, mhz=1965.0):
    return sm_count * 128 * 2 * mhz * 1e6 / 1e12  # FMA = 2 flop, 128 lanes per SM


def tex_peak_gfetch(sm_count, mhz=1965.0):
    return sm_count * 2.0 * mhz * 1e6 / 1e9  # RGBA16F (64-bit texel) trilinear fetches: 2 per clock per SM (scratch/ubench.cu)


def measured_peak_gbs():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons through NVML while the timed regions run."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.h = None

    def run(self):
        if self.h is None:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=1.0)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
class Scene:
    """Device-resident benchmark inputs: rotating 4K G-buffer sets, lights, and the precomputed IBL."""

    def __init__(self, pbr, dev, torch, n_sets=3):
        self.pbr, self.dev, self.torch = pbr, dev, torch
        lib, ctx = dev.lib, dev.ctx
        cuda = torch.device("cuda", dev.ordinal)
        self.keep = []

        def buf(nbytes):
            t = torch.empty(int(nbytes), dtype=torch.uint8, device=cuda)
            self.keep.append(t)
            return t

        self.buf = buf
        self.gbufs, self.outs = [], []
        for i in range(n_sets):
            g = pbr.GBuffer()
            g.width, g.height, g.pitch8, g.pitch4 = W4K, H4K, W4K * 8, W4K * 4
            planes = {k: buf(W4K * H4K * (4 if k in ("depth", "orm") else 8))
                      for k in ("positions", "normals", "albedo", "depth", "orm", "emissive")}
            for k, t in planes.items():
                setattr(g, k, t.data_ptr())
            lib.synth_gbuffer(ctx, C.byref(g), 1, SEED + i, pbr.f3(*CAM), pbr.f3(0, 0, 0), None)
            self.gbufs.append((g, planes))
            self.outs.append(buf(W4K * H4K * 8))
        self.ldr = buf(W4K * H4K * 4)
        from pbr_b200.synth import synth_lights
        self.lights_host = synth_lights(N_LIGHTS, seed=2)
        self.lights = buf(self.lights_host.nbytes)
        self.lights.copy_(torch.from_numpy(self.lights_host.view(np.uint8).reshape(-1)))
        # ---- IBL precompute (config[2] shapes), each kernel timed once with CUDA events ----
        self.ibl_ms = {}
        env = buf(ENV_W * ENV_H * 16)
        lib.synth_environment(ctx, env.data_ptr(), ENV_W, ENV_H, SEED, None)
        chain_bytes = lib.cube_chain_texels(CUBE_RES, CUBE_MIPS) * 8
        self.src_chain, self.pref = buf(chain_bytes), buf(chain_bytes)
        self.irr, self.lut = buf(6 * IRR_RES * IRR_RES * 8), buf(LUT_RES * LUT_RES * 8)
        self.ham = buf(IBL_SAMPLES * 8)
        torch.cuda.synchronize()
        steps = [
            ("hammersley", lambda: lib.ibl_hammersley(ctx, IBL_SAMPLES, self.ham.data_ptr(), self.stream())),
            ("equirect_to_cube", lambda: lib.ibl_equirect_to_cube(ctx, env.data_ptr(), ENV_W, ENV_H, CUBE_RES, self.src_chain.data_ptr(), self.stream())),
            ("cube_mipgen", lambda: lib.ibl_cube_mipgen(ctx, self.src_chain.data_ptr(), CUBE_RES, CUBE_MIPS, self.stream())),
            ("prefilter_ggx", lambda: lib.ibl_prefilter_ggx(ctx, self.src_chain.data_ptr(), CUBE_RES, CUBE_MIPS, IBL_SAMPLES, self.pref.data_ptr(), 0, CUBE_MIPS - 1, 0, 1, self.stream())),
            ("irradiance", lambda: lib.ibl_irradiance(ctx, self.src_chain.data_ptr(), CUBE_RES, CUBE_MIPS, IRR_RES, IBL_SAMPLES, self.irr.data_ptr(), self.stream())),
            ("brdf_lut", lambda: lib.ibl_brdf_lut(ctx, LUT_RES, IBL_SAMPLES, self.lut.data_ptr(), self.stream())),
        ]
        for name, fn in steps:
            fn()  # warm (also the run whose outputs are used)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            self.ibl_ms[name] = round(a.elapsed_time(b), 4)
        # the same two convolutions from a PREPARED source (pbrcu_ibl_source: the flagged texture copy built once for both, as a
        # precompute that runs K6 and K7 from one environment does); results must equal the plain entry points' bit for bit
        self.source = C.c_void_p()
        check_pref, check_irr = buf(chain_bytes), buf(6 * IRR_RES * IRR_RES * 8)
        prepared = [
            ("source_prepare", lambda: (lib.ibl_source_free(ctx, self.source), lib.ibl_source_create(ctx, self.src_chain.data_ptr(), CUBE_RES, CUBE_MIPS, self.stream(), C.byref(self.source)))),
            ("prefilter_ggx_prepared", lambda: lib.ibl_prefilter_ggx_prepared(ctx, self.source, IBL_SAMPLES, check_pref.data_ptr(), 0, CUBE_MIPS - 1, 0, 1, None, 0, self.stream())),
            ("irradiance_prepared", lambda: lib.ibl_irradiance_prepared(ctx, self.source, IRR_RES, IBL_SAMPLES, check_irr.data_ptr(), self.stream())),
        ]
        for name, fn in prepared:
            fn()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            self.ibl_ms[name] = round(a.elapsed_time(b), 4)
        self.ibl_ms["prepared_equals_plain"] = bool(torch.equal(check_pref, self.pref) and torch.equal(check_irr, self.irr))
        self.ibl_ms["precompute_total_prepared"] = round(sum(self.ibl_ms[k] for k in ("hammersley", "equirect_to_cube", "cube_mipgen", "source_prepare",
                                                                                    "prefilter_ggx_prepared", "irradiance_prepared", "brdf_lut")), 4)
        self.keep = [t for t in self.keep if t is not check_pref and t is not check_irr]
        del check_pref, check_irr
        # the same prefilter with software fp32-weight filtering (pbrcu_ibl_set_exact_filtering): time, and
        # how far the texture-unit result is from it
        exact = buf(chain_bytes)
        lib.ibl_set_exact_filtering(ctx, 1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        lib.ibl_prefilter_ggx(ctx, self.src_chain.data_ptr(), CUBE_RES, CUBE_MIPS, IBL_SAMPLES, exact.data_ptr(), 0, CUBE_MIPS - 1, 0, 1, self.stream())
        b.record()
        torch.cuda.synchronize()
        lib.ibl_set_exact_filtering(ctx, 0)
        self.ibl_ms["prefilter_ggx_fp32_weights"] = round(a.elapsed_time(b), 4)
        lvl1 = lib.cube_mip_offset(CUBE_RES, 1) * 8
        f = self.pref[lvl1:].view(torch.float16).float().view(-1, 4)[:, :3]
        e = exact[lvl1:].view(torch.float16).float().view(-1, 4)[:, :3]
        err = (f - e).abs()
        rel = err / (e.abs() + 1e-3)
        self.prefilter_error = {"what": "texture-unit prefilter vs fp32-weight prefilter, levels 1..11, fp16 outputs",
                                "max_abs": float(err.max()), "max_rel": float(rel.max()),
                                "rel_p99.99": float(torch.quantile(rel.flatten()[:: max(1, rel.numel() // 4_000_000)], 0.9999)),
                                "violations_of_1e-3+2^-9": int((err > 1e-3 + 2.0 ** -9 * e.abs()).sum()), "values": int(err.numel()),
                                "psnr_db": float(10 * torch.log10(e.max() ** 2 / (err.double() ** 2).mean()))}
        del f, e, err, rel
        self.keep = [t for t in self.keep if t is not exact]
        del exact
        s = pbr.Ibl()
        s.irradiance, s.irr_res = self.irr.data_ptr(), IRR_RES
        s.prefiltered, s.pref_res, s.pref_mips = self.pref.data_ptr(), CUBE_RES, CUBE_MIPS
        s.lut, s.lut_res = self.lut.data_ptr(), LUT_RES
        # footprint-record copies of the three textures (built once, like the precompute; 1.07 GB)
        self.accel = C.c_void_p()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        if os.environ.get("PBRCU_BENCH_GATHER"):  # profiling runs of the opt-in gather mode (scratch/make_profiles.sh)
            s.flags = pbr.IBL_GATHER
        lib.ibl_accel_build(ctx, C.byref(s), C.byref(self.accel), self.stream())
        b.record()
        torch.cuda.synchronize()
        self.ibl_ms["accel_build"] = round(a.elapsed_time(b), 4)
        peak, _ = measured_peak_gbs()
        sms = dev.sm_count
        lvl_texels = lambda m: 6 * (CUBE_RES >> m) ** 2
        k7_fetches = sum(lvl_texels(m) for m in range(1, CUBE_MIPS)) * IBL_SAMPLES
        ms_of = lambda k: self.ibl_ms[k] * 1e-3
        self.ibl_roofline = {
            "equirect_to_cube": roof("hbm", (ENV_W * ENV_H * 16 + lvl_texels(0) * 8) / ms_of("equirect_to_cube") / 1e9, peak, "GB/s"),
            "cube_mipgen": roof("hbm", (chain_bytes + chain_bytes - lvl_texels(0) * 8) / ms_of("cube_mipgen") / 1e9, peak, "GB/s", "reads every level but the last, writes every level but the first"),
            "prefilter_ggx": roof("texture fetch rate", k7_fetches / ms_of("prefilter_ggx_prepared") / 1e9, tex_peak_gfetch(sms), "G trilinear RGBA16F fetches/s",
                                  "levels 1..11 x 1024 samples, one hardware trilinear cube fetch each; level 0 is a copy; prepared source (prefilter_ggx_prepared)"),
            "irradiance": roof("texture fetch rate", 6 * IRR_RES * IRR_RES * IBL_SAMPLES / ms_of("irradiance_prepared") / 1e9, tex_peak_gfetch(sms), "G trilinear RGBA16F fetches/s",
                               "24,576 texels: a single partial wave, latency-bound; prepared source (irradiance_prepared)"),
            "brdf_lut": roof("fp32", LUT_RES * LUT_RES * IBL_SAMPLES * 40 / ms_of("brdf_lut") / 1e12, fp32_peak_tflops(sms), "TFLOP/s",
                             "40 flop per sample (SURVEY 8(d)); deterministic math: no FMA contraction, correctly rounded sqrt / div"),
        }
        s.accel = self.accel
        self.ibl = s

    def stream(self):
        return self.torch.cuda.current_stream().cuda_stream

    def shade(self, i):
        g, _ = self.gbufs[i % len(self.gbufs)]
        # the four lights are passed as HOST data: they travel in the kernel's parameter block (shade_ggx_cl_kernel)
        self.dev.lib.shade_ggx_ibl(self.dev.ctx, C.byref(g), self.pbr.f3(*CAM), self.lights_host.ctypes.data, N_LIGHTS, None,
                                   C.byref(self.ibl), self.pbr.WHOLE, self.outs[i % len(self.outs)].data_ptr(), W4K * 8, self.stream())

    def lighting(self, i):
        g, _ = self.gbufs[i % len(self.gbufs)]
        self.dev.lib.lighting_blinnphong(self.dev.ctx, C.byref(g), self.pbr.f3(*CAM), self.pbr.WHOLE,
                                         self.outs[i % len(self.outs)].data_ptr(), W4K * 8, self.stream())

    def shade_then_tonemap(self, i):
        self.shade(i)
        self.dev.lib.tonemap_aces_gamma(self.dev.ctx, self.outs[i % len(self.outs)].data_ptr(), W4K * 8, W4K, H4K, self.pbr.WHOLE, self.ldr.data_ptr(), W4K * 4, self.stream())

    def shade_ldr(self, i):
        g, _ = self.gbufs[i % len(self.gbufs)]
        self.dev.lib.shade_ggx_ibl_ldr(self.dev.ctx, C.byref(g), self.pbr.f3(*CAM), self.lights_host.ctypes.data, N_LIGHTS, None, C.byref(self.ibl), self.pbr.WHOLE,
                                       self.ldr.data_ptr(), W4K * 4, None, 0, self.stream())

    def fused(self, i):
        g, _ = self.gbufs[i % len(self.gbufs)]
        self.dev.lib.lighting_tonemap_fused(self.dev.ctx, C.byref(g), self.pbr.f3(*CAM), self.pbr.WHOLE, None, 0,
                                            self.ldr.data_ptr(), W4K * 4, self.stream())


def extra_configs(scene, torch, dist, world, rank):
    """BASELINE configs[3] and [4], measured beside the headline (same timing protocol, fewer steps):
    [3] one 7680x4320 frame, 256 ranged point lights, K3c light clustering + K3, the frame's rows cut into
        16-aligned bands over the ranks (strong scaling: value = whole-frame pixels / max-over-ranks time);
    [4] 256 camera views at 1920x1080 (orbit rings around the sphere scene), contiguous view blocks per rank,
        K3 with 4 lights + IBL per view (no exchange: views are independent)."""
    pbr, dev, lib, ctx = scene.pbr, scene.dev, scene.dev.lib, scene.dev.ctx
    from pbr_b200.synth import synth_lights
    from pbr_b200.partition import band_views, cost_balanced_bands, row_band, view_block
    out = {}
    # ---- config[3]: 8K, 256 clustered lights, row bands ----
    W8, H8, NL = 7680, 4320, 256
    g = pbr.GBuffer()
    g.width, g.height, g.pitch8, g.pitch4 = W8, H8, W8 * 8, W8 * 4
    planes = {k: scene.buf(W8 * H8 * (4 if k in ("depth", "orm") else 8)) for k in ("positions", "normals", "albedo", "depth", "orm", "emissive")}
    for k, t in planes.items():
        setattr(g, k, t.data_ptr())
    lib.synth_gbuffer(ctx, C.byref(g), 1, SEED + 40, pbr.f3(*CAM), pbr.f3(0, 0, 0), None)
    hdr = scene.buf(W8 * H8 * 8)
    lights_host = synth_lights(NL, seed=4, ranged=True)
    lights = scene.buf(lights_host.nbytes)
    lights.copy_(torch.from_numpy(lights_host.view(np.uint8).reshape(-1)))
    tiles = pbr.LightTiles()
    lib.light_tiles_alloc(ctx, W8, H8, NL, C.byref(tiles))
    y0, y1 = row_band(H8, world, rank, align=16)
    bands_all = [row_band(H8, world, r, align=16) for r in range(world)]
    if world > 1:
        # bands of equal COST, not equal height: one untimed whole-frame clustering stands in for "the previous frame's light
        # lists"; every rank derives the same cuts from the same counts (cost of a tile ~ 9 light-equivalents of fixed work +
        # its lights), so nothing is exchanged
        lib.cluster_lights(ctx, C.byref(g), lights.data_ptr(), NL, C.byref(tiles), scene.stream())
        prev = np.empty(tiles.tiles_x * tiles.tiles_y, np.uint32)
        lib.memcpy_d2h(prev.ctypes.data, tiles.counts, prev.nbytes, scene.stream())
        torch.cuda.synchronize()
        row_costs = (9.0 + prev.reshape(tiles.tiles_y, tiles.tiles_x)).sum(axis=1)
        bands_all = cost_balanced_bands(row_costs, H8, world)
        y0, y1 = bands_all[rank]
    band = band_views(pbr, g, tiles, y0, y1)  # this rank's rows of the G-buffer and of the tile lists (pointer arithmetic only)

    def step8k(_):
        # each rank clusters and shades only its own band (per-band clustering == whole-frame clustering: tiles are 16-aligned)
        if y1 > y0:
            lib.cluster_lights(ctx, C.byref(band.gbuffer), lights.data_ptr(), NL, C.byref(band.tiles), scene.stream())
            lib.shade_ggx_ibl(ctx, C.byref(band.gbuffer), pbr.f3(*CAM), lights.data_ptr(), NL, C.byref(band.tiles), C.byref(scene.ibl), pbr.WHOLE,
                              hdr.data_ptr() + y0 * W8 * 8, W8 * 8, scene.stream())

    ms = timed(torch, dist, step8k, 5, 3)
    gathered = None
    if world > 1:
        # north_star: "NCCL over NVLink is used only to gather the finished tiles": the same step followed by the exchange that
        # leaves the whole HDR frame on every rank
        # ONE collective: every rank shades its band straight into its slot of a padded gather buffer (slots of the tallest band's
        # size -- the cost-balanced bands differ in height), one in-place ncclAllGather, then the bands are unpacked into the frame
        rb = W8 * 8
        slot = max(e - b for b, e in bands_all) * rb
        stage_all = scene.buf(world * slot)

        def step8k_gather(_):
            if y1 > y0:
                lib.cluster_lights(ctx, C.byref(band.gbuffer), lights.data_ptr(), NL, C.byref(band.tiles), scene.stream())
                lib.shade_ggx_ibl(ctx, C.byref(band.gbuffer), pbr.f3(*CAM), lights.data_ptr(), NL, C.byref(band.tiles), C.byref(scene.ibl), pbr.WHOLE,
                                  stage_all.data_ptr() + rank * slot, rb, scene.stream())
            dist.all_gather_into_tensor(stage_all, stage_all[rank * slot:(rank + 1) * slot])
            for r, (b, e) in enumerate(bands_all):
                if e > b:
                    hdr[b * rb:e * rb].copy_(stage_all[r * slot:r * slot + (e - b) * rb], non_blocking=True)

        ms_g = timed(torch, dist, step8k_gather, 5, 3)
        # the same frame with the gather fused into K3's stores: every rank's frame lives in symmetric memory, each rank
        # stores its finished band into its own and (over NVLink, from inside the shading kernel) into every peer's frame;
        # one device-side barrier orders the frame.  No exchange pass, no NCCL on the data path.
        fused = {}
        try:
            from pbr_b200.partition import PeerFrames
            pf = PeerFrames(W8 * H8 * 8, hdr.device, dist)
            reps, nrep = pf.replica_array(y0 * W8 * 8)

            def step8k_fused(_):
                if y1 > y0:
                    lib.cluster_lights(ctx, C.byref(band.gbuffer), lights.data_ptr(), NL, C.byref(band.tiles), scene.stream())
                    lib.shade_ggx_ibl_replicated(ctx, C.byref(band.gbuffer), pbr.f3(*CAM), lights.data_ptr(), NL, C.byref(band.tiles), C.byref(scene.ibl),
                                                 pbr.WHOLE, pf.local.data_ptr() + y0 * W8 * 8, W8 * 8, reps, nrep, scene.stream())
                pf.barrier()

            pf.local.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            ms_f = timed(torch, dist, step8k_fused, 5, 3)
            fused = {"ms_per_frame_fused_gather": round(ms_f / 5, 4), "fused_gather": "pbrcu_shade_ggx_ibl_replicated: K3 stores each finished texel "
                     "into the frame of every rank (torch symmetric memory, NVLink P2P stores) + one device-side barrier"}
            if pf.multicast:
                def step8k_mc(_):
                    if y1 > y0:
                        lib.cluster_lights(ctx, C.byref(band.gbuffer), lights.data_ptr(), NL, C.byref(band.tiles), scene.stream())
                        lib.shade_ggx_ibl_multicast(ctx, C.byref(band.gbuffer), pbr.f3(*CAM), lights.data_ptr(), NL, C.byref(band.tiles), C.byref(scene.ibl),
                                                    pbr.WHOLE, pf.multicast + y0 * W8 * 8, W8 * 8, scene.stream())
                    pf.barrier()

                check = scene.buf(W8 * H8 * 8)
                check.copy_(pf.local)  # the P2P-store result, compared below
                pf.local.zero_()
                torch.cuda.synchronize()
                dist.barrier()
                ms_m = timed(torch, dist, step8k_mc, 5, 3)
                fused.update({"ms_per_frame_multicast_gather": round(ms_m / 5, 4), "multicast_gather": "pbrcu_shade_ggx_ibl_multicast: one multimem.st per "
                              "texel to the NVSwitch multicast address of the frames + one device-side barrier",
                              "multicast_frame_equals_p2p_frame": bool(torch.equal(check, pf.local))})
                scene.keep = [t for t in scene.keep if t is not check]
                del check
            else:
                fused["multicast_gather_unavailable"] = "symmetric memory reports no multicast address on this box"
        except Exception as e:  # no symmetric memory on this box: the NCCL figures above stand alone
            pf = None
            fused = {"fused_gather_unavailable": f"{type(e).__name__}: {e}"[:200]}
        # the gathered frame against the whole frame shaded on this rank alone (untimed)
        whole = scene.buf(W8 * H8 * 8)
        lib.cluster_lights(ctx, C.byref(g), lights.data_ptr(), NL, C.byref(tiles), scene.stream())
        lib.shade_ggx_ibl(ctx, C.byref(g), pbr.f3(*CAM), lights.data_ptr(), NL, C.byref(tiles), C.byref(scene.ibl), pbr.WHOLE, whole.data_ptr(), W8 * 8, scene.stream())
        torch.cuda.synchronize()
        same = torch.tensor([int(torch.equal(whole, hdr)), int(pf is None or torch.equal(whole, pf.local))], device=hdr.device)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        gathered = {"ms_per_frame_with_gather": round(ms_g / 5, 4), "gather_bytes_per_frame": W8 * H8 * 8,
                    "gathered_frame_equals_unsharded": bool(same[0].item()), **fused}
        if pf is not None:
            gathered["fused_gathered_frame_equals_unsharded"] = bool(same[1].item())
        scene.keep = [t for t in scene.keep if t is not whole]
        del whole
    # tile counts live in library-owned memory: copy this rank's out through the C ABI
    host_counts = np.empty(max(band.tiles.tiles_x * band.tiles.tiles_y, 1), np.uint32)
    lib.memcpy_d2h(host_counts.ctypes.data, band.tiles.counts, band.tiles.tiles_x * band.tiles.tiles_y * 4, scene.stream())
    torch.cuda.synchronize()
    # (light, pixel) pairs the frame evaluates: every pixel of a tile loops its tile's list (90 flop per pair + 120 per pixel)
    pairs = torch.tensor([float(host_counts[:band.tiles.tiles_x * band.tiles.tiles_y].astype(np.float64).sum()) * 256.0], dtype=torch.float64, device=hdr.device)
    if dist is not None:
        dist.all_reduce(pairs)
    t8 = ms / 5 * 1e-3
    peak_hbm, _ = measured_peak_gbs()
    flop = float(pairs.item()) * 90.0 + W8 * H8 * 120.0
    # the same frame unsharded on this GPU alone (no barrier): the N = 1 time the strong-scaling efficiency refers to
    ms_whole = ms / 5
    if world > 1:
        tiles1 = pbr.LightTiles()
        lib.light_tiles_alloc(ctx, W8, H8, NL, C.byref(tiles1))
        whole = scene.buf(W8 * H8 * 8)

        def step_whole(_):
            lib.cluster_lights(ctx, C.byref(g), lights.data_ptr(), NL, C.byref(tiles1), scene.stream())
            lib.shade_ggx_ibl(ctx, C.byref(g), pbr.f3(*CAM), lights.data_ptr(), NL, C.byref(tiles1), C.byref(scene.ibl), pbr.WHOLE, whole.data_ptr(), W8 * 8, scene.stream())

        ms_whole = timed(torch, None, step_whole, 5, 3) / 5
        lib.light_tiles_free(ctx, C.byref(tiles1))
        scene.keep = [t for t in scene.keep if t is not whole]
        del whole
    out["config3_8k_256_lights"] = {"mpix_s": round(W8 * H8 * 5 / (ms * 1e-3) / 1e6, 1), "ms_per_frame": round(ms / 5, 4),
                                    "ms_per_frame_unsharded_on_one_gpu": round(ms_whole, 4),
                                    "roofline": roof("fp32", flop / t8 / 1e12, fp32_peak_tflops(dev.sm_count) * world, "TFLOP/s",
                                                     f"{pairs.item():.4g} (light, pixel) pairs x 90 flop + 120 flop per pixel over {world} GPU(s); peak = SMs x 128 lanes x 2 x 1.965 GHz"),
                                    "roofline_hbm": roof("hbm", (B_PER_PX_K3 + 16) * W8 * H8 / t8 / 1e9, peak_hbm * world, "GB/s", "44 B/px K3 + 16 B/px K3c (positions + normals re-read)"),
                                    "rows_per_rank": y1 - y0, "mean_lights_per_tile": round(float(host_counts.mean()), 2),
                                    "max_lights_per_tile": int(host_counts.max()), "scaling": "strong (cost-balanced row bands from the previous frame's light lists, no exchange)",
                                    "what": "pbrcu_cluster_lights + pbrcu_shade_ggx_ibl on the rank's row band, 256 point lights range U[2,6], IBL"}
    if gathered:
        out["config3_8k_256_lights"].update(gathered)
    lib.light_tiles_free(ctx, C.byref(tiles))
    # ---- config[4]: 256 views x 1080p, view blocks ----
    WV, HV, NV = 1920, 1080, 256
    v0, v1 = view_block(NV, world, rank)
    nv = v1 - v0
    # one allocation per plane with the rank's views back to back (view stride = one view): the batched entry shades
    # them all in ONE launch
    stride8, stride4 = WV * HV * 8, WV * HV * 4
    planes = {k2: scene.buf(max(nv, 1) * (stride4 if k2 == "orm" else stride8)) for k2 in ("positions", "normals", "albedo", "orm", "emissive")}
    outs = scene.buf(max(nv, 1) * stride8)
    cams = np.zeros((max(nv, 1), 3), np.float32)
    g0 = pbr.GBuffer()
    g0.width, g0.height, g0.pitch8, g0.pitch4 = WV, HV, WV * 8, WV * 4
    for k2, t in planes.items():
        setattr(g0, k2, t.data_ptr())
    for i, v in enumerate(range(v0, v1)):
        ring, k = divmod(v, NV // 2)
        ang = 2 * np.pi * k / (NV // 2)
        cams[i] = (12.0 * np.cos(ang), -1.0 - 4.0 * ring, 12.0 * np.sin(ang))
        gv = pbr.GBuffer()
        gv.width, gv.height, gv.pitch8, gv.pitch4 = WV, HV, WV * 8, WV * 4
        for k2, t in planes.items():
            setattr(gv, k2, t.data_ptr() + i * (stride4 if k2 == "orm" else stride8))
        lib.synth_gbuffer(ctx, C.byref(gv), 1, SEED + 50, pbr.f3(*cams[i]), pbr.f3(0, 0, 0), None)  # one scene, many cameras
    torch.cuda.synchronize()
    cam_ptr = cams.ctypes.data_as(C.POINTER(C.c_float))

    def step_views(_):
        if nv:
            lib.shade_ggx_ibl_batched(ctx, C.byref(g0), stride8, stride4, nv, cam_ptr, scene.lights_host.ctypes.data, N_LIGHTS, C.byref(scene.ibl), outs.data_ptr(), WV * 8,
                                      stride8, scene.stream())

    ms = timed(torch, dist, step_views, 3, 3)
    out["config4_256_views_1080p"] = {"mpix_s": round(NV * WV * HV * 3 / (ms * 1e-3) / 1e6, 1), "ms_per_pass": round(ms / 3, 3),
                                      "roofline": roof("hbm", B_PER_PX_K3 * NV * WV * HV / (ms / 3 * 1e-3) / 1e9, measured_peak_gbs()[0] * world, "GB/s", "44 B/px, all views, all ranks"),
                                      "views_per_rank": nv, "scaling": "strong (view blocks, no exchange)",
                                      "what": "pbrcu_shade_ggx_ibl_batched: the rank's views in one launch, 4 point lights + IBL"}
    # the reference's own per-view frame over the same views (Tier A, pinned to its shaders): lighting (K1) batched, and lighting + tonemap fused
    # (K1+K2, LDR only) batched -- one launch each for the rank's block
    ldr_views = scene.buf(max(nv, 1) * stride4)

    def step_views_k1(_):
        if nv:
            lib.lighting_blinnphong_batched(ctx, C.byref(g0), stride8, stride4, nv, cam_ptr, outs.data_ptr(), WV * 8, stride8, scene.stream())

    def step_views_fused(_):
        if nv:
            lib.lighting_tonemap_fused_batched(ctx, C.byref(g0), stride8, nv, cam_ptr, None, 0, 0, ldr_views.data_ptr(), WV * 4, stride4, scene.stream())

    ms_k1 = timed(torch, dist, step_views_k1, 3, 3) / 3
    ms_fu = timed(torch, dist, step_views_fused, 3, 3) / 3
    peak_v = measured_peak_gbs()[0] * world
    out["config4_256_views_1080p_tier_a"] = {
        "what": "the reference's per-view frame (lighting.glsl, then aces+gamma.glsl) over the same 256 views: pbrcu_lighting_blinnphong_batched and "
                "pbrcu_lighting_tonemap_fused_batched (LDR only), the rank's views in one launch each",
        "ms_lighting": round(ms_k1, 3), "ms_lighting_tonemap_fused_ldr": round(ms_fu, 3), "views_per_rank": nv,
        "roofline_lighting": roof("hbm", B_PER_PX_K1 * NV * WV * HV / (ms_k1 * 1e-3) / 1e9, peak_v, "GB/s", "32 B/px, all views, all ranks"),
        "roofline_fused_ldr": roof("hbm", 28 * NV * WV * HV / (ms_fu * 1e-3) / 1e9, peak_v, "GB/s", "28 B/px (24 read + 4 written), all views, all ranks")}
    scene.keep = [t for t in scene.keep if t is not ldr_views]
    del ldr_views
    out["geometry_pass_1080p"] = geometry_bench(scene, torch, dist, world, rank)
    out["geometry_pass_1080p"]["config4_256_views_gltf"] = gltf_views_bench(torch, dist, world, rank)
    out["config2_prefilter_sharded"] = prefilter_sharded(scene, torch, dist, world, rank)
    return out


def prefilter_sharded(scene, torch, dist, world, rank):
    """BASELINE configs[2]'s prefilter (2048^2, 12 mips, 1024 spp) over the ranks: every rank convolves its row
    block of every level (pbrcu_ibl_prefilter_ggx shard/nshards) from its own replica of the source chain, then
    the finished chain is all-gathered over NCCL so every GPU can shade with it.  The source is PREPARED once
    (pbrcu_ibl_source, outside the timed region like the source chain itself: it is the same on every rank).  Strong
    scaling; the gather is inside the timed region, and the gathered chain is checked against the unsharded one."""
    from pbr_b200.partition import gather_chain
    lib, ctx = scene.dev.lib, scene.dev.ctx
    dst = scene.buf(scene.pref.numel())

    def conv(_):
        # level 0 (a copy of the source level, which every rank holds) whole on every rank; levels 1.. by row blocks
        lib.ibl_prefilter_ggx_prepared(ctx, scene.source, IBL_SAMPLES, dst.data_ptr(), 0, 0, 0, 1, None, 0, scene.stream())
        lib.ibl_prefilter_ggx_prepared(ctx, scene.source, IBL_SAMPLES, dst.data_ptr(), 1, CUBE_MIPS - 1, rank, world, None, 0, scene.stream())

    def step(i):
        conv(i)
        gather_chain(dst, CUBE_RES, CUBE_MIPS, world, rank, dist if world > 1 else None, first_level=1)

    ms_conv = timed(torch, dist, conv, 3, 2) / 3
    dst.zero_()
    ms = timed(torch, dist, step, 3, 2) / 3
    same = bool(torch.equal(dst, scene.pref))
    fused = {}
    if world > 1:
        # the same with the gather fused into K7's stores: chains in symmetric memory, every finished texel also stored
        # into the peers' chains (NVLink P2P stores), one device-side barrier -- no NCCL on the data path
        try:
            from pbr_b200.partition import PeerFrames
            pf = PeerFrames(dst.numel(), dst.device, dist)
            reps, nrep = pf.replica_array(0)

            def step_fused(_):
                lib.ibl_prefilter_ggx_prepared(ctx, scene.source, IBL_SAMPLES, pf.local.data_ptr(), 0, 0, 0, 1, None, 0, scene.stream())
                lib.ibl_prefilter_ggx_prepared(ctx, scene.source, IBL_SAMPLES, pf.local.data_ptr(), 1, CUBE_MIPS - 1, rank, world, reps, nrep, scene.stream())
                pf.barrier()

            pf.local.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            ms_f = timed(torch, dist, step_fused, 3, 2) / 3
            ok = torch.tensor([int(torch.equal(pf.local, scene.pref))], device=dst.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            fused = {"ms_prefilter_fused_gather": round(ms_f, 3), "fused_chain_equals_unsharded": bool(ok.item()),
                     "fused_gather": "pbrcu_ibl_prefilter_ggx_replicated: K7 stores each finished texel into every rank's chain (symmetric memory, NVLink P2P stores)"}
            del pf
        except Exception as e:
            fused = {"fused_gather_unavailable": f"{type(e).__name__}: {e}"[:200]}
    return {**fused, "ms_prefilter_and_gather": round(ms, 3), "ms_prefilter_shard_only": round(ms_conv, 3), "shards": world,
            "gathered_chain_equals_unsharded": same, "gather_bytes_per_rank": int((dst.numel() - 6 * CUBE_RES * CUBE_RES * 8) * (world - 1) // world),
            "scaling": "strong (row blocks of levels 1..; level 0 copied whole on every rank; NCCL all-gather of levels 1..)"}


def gltf_views_bench(torch, dist, world, rank):
    """BASELINE configs[4] through the glTF path: the 65-draw scene written as a .glb (tests/gltfwriter.py), imported by the C++ facade
    (pbr::gltf::Loader -> Asset::loadScene -> TransferStager) inside tools/gltf_viewer, and 256 orbit cameras rendered by
    PbrRenderSystem::renderViews (one batched geometry pass + one batched GGX shading pass per rank's view block).  The viewer is a process
    of its own on this rank's GPU; it times enqueue + synchronise with the host clock (min of 3 repeats after a warm-up)."""
    import subprocess
    import tempfile
    viewer = ROOT / "tools" / "gltf_viewer" / "build" / "gltf_viewer"
    if not viewer.exists():
        return {"unavailable": "tools/gltf_viewer/build/gltf_viewer is not built (python -c 'import __graft_entry__ as g; g.build()')"}
    from pbr_b200.partition import view_block
    NV, W, H = 256, 1920, 1080
    v0, v1 = view_block(NV, world, rank)
    out = None
    with tempfile.TemporaryDirectory() as tmp:
        path = Path(tmp) / "scene.glb"
        write_views_glb(path)
        glb_bytes = path.stat().st_size
        if v1 > v0:
            env = dict(os.environ)
            local = int(os.environ.get("LOCAL_RANK", "0"))
            visible = [d for d in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if d]
            env["CUDA_VISIBLE_DEVICES"] = visible[local] if local < len(visible) else str(local)
            r = subprocess.run([str(viewer), str(path), "--size", f"{W}x{H}", "--ggx", "--views", str(v1 - v0), "--view-first", str(v0), "--view-total", str(NV), "--frames", "3",
                                "--orbit", "12,-2", "--times", str(Path(tmp) / "t.json")], capture_output=True, text=True, env=env, timeout=600)
            if r.returncode != 0:
                return {"unavailable": f"gltf_viewer failed: {r.stderr.strip()[-300:]}"}
            out = json.loads((Path(tmp) / "t.json").read_text())
    return _gltf_views_entry(torch, dist, world, out, NV, W, H, v1 - v0, glb_bytes)


def write_views_glb(path):
    """The 65-draw scene of geometry_bench as a .glb: 64 textured spheres (one set of accessors, 64 materials and node transforms) and a ground
    plane that crosses the near plane; 589,828 triangles."""
    sys.path.insert(0, str(ROOT / "tests"))  # the glTF writer is test tooling (not the checker)
    from gltfwriter import Builder
    from pbr_b200.synth import checker_texture, mesh_vertex, uv_sphere
    rng = np.random.default_rng(7)
    b = Builder()
    sv, si = uv_sphere(1.0, 48, 96)
    sv = np.asarray(sv, np.float32).reshape(-1, 12)
    g = 40.0
    quad = np.array([mesh_vertex((-g, 3.0, -g), (0, -1, 0), (1, 0, 0, 1), (0, 0)), mesh_vertex((g, 3.0, -g), (0, -1, 0), (1, 0, 0, 1), (16, 0)),
                     mesh_vertex((g, 3.0, g), (0, -1, 0), (1, 0, 0, 1), (16, 16)), mesh_vertex((-g, 3.0, g), (0, -1, 0), (1, 0, 0, 1), (0, 16))], np.float32)
    qi = np.array([0, 1, 2, 0, 2, 3, 0, 2, 1, 0, 3, 2], np.uint16)
    ct, nt, flat = b.image(checker_texture(64, 64, 1), "checker"), b.image(checker_texture(32, 32, 2, normal_map=True), "bumps"), b.image(np.array([[[128, 128, 255, 255]]], np.uint8), "flat")
    ball = b.primitive(sv, si, 0)  # one set of accessors, 64 materials
    nodes = []
    for k in range(64):
        t, sc, axis, ang = rng.uniform(-8, 8, 3), rng.uniform(0.5, 1.5, 3), rng.standard_normal(3), rng.uniform(0, 6.28)
        axis = axis / np.linalg.norm(axis)
        m = b.material(f"m{k}", rng.uniform(0.3, 1.0, 4), ct, nt, metallic=float(rng.integers(0, 2)), roughness=float(rng.uniform(0.05, 1.0)))
        nodes.append(b.node(f"ball{k}", mesh=b.mesh(f"ball{k}", [dict(ball, material=m)]), translation=t, rotation_xyzw=(*(axis * np.sin(ang / 2)), np.cos(ang / 2)), scale=sc))
    mg = b.material("ground", (0.8, 0.8, 0.8, 1.0), ct, flat, metallic=0.0, roughness=0.6)
    nodes.append(b.node("ground", mesh=b.mesh("ground", [b.primitive(quad, qi, mg)])))
    b.scene(nodes)
    b.write_glb(Path(path))


def _gltf_views_entry(torch, dist, world, out, NV, W, H, nv, glb_bytes):
    ms = torch.tensor([out["ms_geometry"] if out else 0.0, out["ms_geometry_plus_shading"] if out else 0.0], dtype=torch.float64, device="cuda")
    if dist is not None and world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    tri = torch.tensor([out["triangles_per_view"] if out else 0, out["draws"] if out else 0], dtype=torch.int64, device="cuda")
    if dist is not None and world > 1:
        dist.all_reduce(tri, op=dist.ReduceOp.MAX)
    ms_geo, ms_both = float(ms[0]), float(ms[1])
    return {"views": NV, "views_per_rank": nv, "ms_geometry_all_views": round(ms_geo, 3), "ms_geometry_plus_shading_all_views": round(ms_both, 3),
            "us_geometry_per_view": round(ms_geo * 1e3 / max(nv, 1), 2), "mpix_s": round(NV * W * H / (ms_both * 1e-3) / 1e6, 1),
            "triangles_per_view": int(tri[0]), "draws": int(tri[1]), "glb_bytes": glb_bytes, "scaling": "strong (view blocks, no exchange)",
            "timing": "host clock around enqueue + stream synchronise inside the viewer process, min of 3 repeats after a warm-up, max over ranks",
            "what": "tests/gltfwriter.py scene (.glb: 64 textured spheres + a ground plane that crosses the near plane) -> pbr::gltf::Loader -> Asset::loadScene -> "
                    "PbrRenderSystem::renderViews = pbrcu_geometry_pass_batched + pbrcu_shade_ggx_ibl_batched, 4 point lights + IBL, 1920x1080; parity of the path: "
                    "tests/test_gpu_viewer.py::test_viewer_views_mode_matches_oracle (first / middle / last view bit-exact vs the oracle's geometry pass)"}


def geometry_bench(scene, torch, dist, world, rank):
    """K10 (pbrcu_geometry_pass, the CUDA visibility-buffer rasteriser that FILLS the G-buffer): 64 textured,
    normal-mapped spheres of 9216 triangles each + two large quads = 589,828 triangles at 1920x1080, one camera per
    step on an orbit; then the same frame end to end: geometry pass -> GGX shading (4 lights + IBL) -> the HDR image."""
    pbr, lib, ctx = scene.pbr, scene.dev.lib, scene.dev.ctx
    from pbr_b200.synth import checker_texture, col_major, look_at_rh, mesh_vertex, model_push_constant, perspective_rh_no, uv_sphere
    W, H = 1920, 1080
    rng = np.random.default_rng(7)

    def dev_of(a):
        t = scene.buf(a.nbytes)
        t.copy_(torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)))
        return t

    sv, si = uv_sphere(1.0, 48, 96)
    g = 40.0
    quad = np.array([mesh_vertex((-g, 3.0, -g), (0, -1, 0), (1, 0, 0, 1), (0, 0)), mesh_vertex((g, 3.0, -g), (0, -1, 0), (1, 0, 0, 1), (16, 0)),
                     mesh_vertex((g, 3.0, g), (0, -1, 0), (1, 0, 0, 1), (16, 16)), mesh_vertex((-g, 3.0, g), (0, -1, 0), (1, 0, 0, 1), (0, 16))], np.float32)
    qi = np.array([0, 1, 2, 0, 2, 3, 0, 2, 1, 0, 3, 2], np.uint16)
    vb, ib, qvb, qib = dev_of(sv), dev_of(si), dev_of(quad), dev_of(qi)
    ctex, ntex, flat = dev_of(checker_texture(64, 64, 1)), dev_of(checker_texture(32, 32, 2, normal_map=True)), dev_of(np.array([[[128, 128, 255, 255]]], np.uint8))
    n_draws = 65
    draws = (pbr.Draw * n_draws)()

    def fill(d, vbuf, ibuf, nidx, model, color, ct, cshape, nt, nshape, rough, metal):
        d.vertices, d.indices, d.first_index, d.index_count, d.vertex_offset = vbuf.data_ptr(), ibuf.data_ptr(), 0, nidx, 0
        d.vertex_count = vbuf.numel() // 48
        d.model[:] = model[0].tolist(); d.normal_model[:] = model[1].tolist()
        m = d.material
        m.color[:] = [float(c) for c in color]
        m.color_tex.texels, m.color_tex.width, m.color_tex.height, m.color_tex.pitch = ct.data_ptr(), cshape[1], cshape[0], cshape[1] * 4
        m.normal_tex.texels, m.normal_tex.width, m.normal_tex.height, m.normal_tex.pitch = nt.data_ptr(), nshape[1], nshape[0], nshape[1] * 4
        m.occlusion, m.roughness, m.metallic = 1.0, rough, metal

    for k in range(64):
        fill(draws[k], vb, ib, len(si), model_push_constant(rng.uniform(-8, 8, 3), rng.uniform(0.5, 1.5, 3), rng.standard_normal(3), rng.uniform(0, 6.28)),
             rng.uniform(0.3, 1.0, 4), ctex, (64, 64), ntex, (32, 32), float(rng.uniform(0.05, 1.0)), float(rng.integers(0, 2)))
    fill(draws[64], qvb, qib, len(qi), model_push_constant(), (0.8, 0.8, 0.8, 1.0), ctex, (64, 64), flat, (1, 1), 0.6, 0.0)
    n_tris = 64 * (len(si) // 3) + len(qi) // 3
    gb = pbr.GBuffer()
    gb.width, gb.height, gb.pitch8, gb.pitch4 = W, H, W * 8, W * 4
    planes = {k: scene.buf(W * H * (4 if k in ("depth", "orm") else 8)) for k in ("positions", "normals", "albedo", "depth", "orm", "emissive")}
    for k, t in planes.items():
        setattr(gb, k, t.data_ptr())
    hdr = scene.buf(W * H * 8)
    proj = col_major(perspective_rh_no(np.pi / 2, W / H))
    cams = [(12.0 * np.cos(a), -2.0, 12.0 * np.sin(a)) for a in np.linspace(0, 2 * np.pi, 16, endpoint=False)]
    views = [col_major(look_at_rh(c, (0, 0, 0))) for c in cams]

    def geo(i):
        lib.geometry_pass(ctx, C.byref(gb), views[i % 16].ctypes.data, proj.ctypes.data, draws, n_draws, None, scene.stream())

    def frame(i):
        geo(i)
        lib.shade_ggx_ibl(ctx, C.byref(gb), pbr.f3(*cams[i % 16]), scene.lights_host.ctypes.data, N_LIGHTS, None, C.byref(scene.ibl), pbr.WHOLE, hdr.data_ptr(), W * 8, scene.stream())

    ms_geo = timed(torch, dist, geo, 16, 4) / 16
    ms_frame = timed(torch, dist, frame, 16, 4) / 16
    # ---- BASELINE configs[4] as written: 256 camera views of the SCENE -- every view rasterised (pbrcu_geometry_pass_batched) and shaded
    # (pbrcu_shade_ggx_ibl_batched), contiguous view blocks over the ranks, no exchange ----
    from pbr_b200.partition import view_block
    NV = 256
    v0, v1 = view_block(NV, world, rank)
    nv = v1 - v0
    s8, s4 = W * H * 8, W * H * 4
    vplanes = {k: scene.buf(max(nv, 1) * (s4 if k in ("depth", "orm") else s8)) for k in ("positions", "normals", "albedo", "depth", "orm", "emissive")}
    vout = scene.buf(max(nv, 1) * s8)
    gv = pbr.GBuffer()
    gv.width, gv.height, gv.pitch8, gv.pitch4 = W, H, W * 8, W * 4
    for k, t in vplanes.items():
        setattr(gv, k, t.data_ptr())
    orbit = []
    for v in range(v0, v1):
        ring, k = divmod(v, NV // 2)
        ang = 2 * np.pi * k / (NV // 2)
        orbit.append((12.0 * np.cos(ang), -1.0 - 4.0 * ring, 12.0 * np.sin(ang)))
    vcams = np.ascontiguousarray(np.array(orbit or [(0, 0, 1)], np.float32))
    vviews = np.ascontiguousarray(np.stack([col_major(look_at_rh(c, (0, 0, 0))) for c in (orbit or [(0, 0, 1)])]).astype(np.float32))
    vprojs = np.ascontiguousarray(np.tile(np.asarray(proj, np.float32).reshape(1, 16), (max(nv, 1), 1)))

    def geo_views(_):
        if nv:
            lib.geometry_pass_batched(ctx, C.byref(gv), s8, s4, nv, vviews.ctypes.data, vprojs.ctypes.data, draws, n_draws, scene.stream())

    def frames_views(i):
        geo_views(i)
        if nv:
            lib.shade_ggx_ibl_batched(ctx, C.byref(gv), s8, s4, nv, vcams.ctypes.data_as(C.POINTER(C.c_float)), scene.lights_host.ctypes.data, N_LIGHTS, C.byref(scene.ibl),
                                      vout.data_ptr(), W * 8, s8, scene.stream())

    ms_geo_views = timed(torch, dist, geo_views, 2, 1) / 2
    ms_frames_views = timed(torch, dist, frames_views, 2, 1) / 2
    # three sampled views of the batch against the single-view entry point, bit for bit
    same = True
    for k in sorted({0, nv // 2, nv - 1} if nv else set()):
        lib.geometry_pass(ctx, C.byref(gb), vviews[k].ctypes.data, proj.ctypes.data, draws, n_draws, None, scene.stream())
        torch.cuda.synchronize()
        for name in ("positions", "normals", "albedo", "orm", "emissive", "depth"):
            n = s4 if name in ("depth", "orm") else s8
            same = same and bool(torch.equal(planes[name], vplanes[name][k * n:(k + 1) * n]))
    views_entry = {"views": NV, "views_per_rank": nv, "ms_geometry_all_views": round(ms_geo_views, 3), "ms_geometry_plus_shading_all_views": round(ms_frames_views, 3),
                   "us_geometry_per_view": round(ms_geo_views * 1e3 / max(nv, 1), 2), "mpix_s": round(NV * W * H / (ms_frames_views * 1e-3) / 1e6, 1),
                   "sampled_views_equal_single_pass": same, "scaling": "strong (view blocks, no exchange)",
                   "what": "pbrcu_geometry_pass_batched (draw tables uploaded once, camera-independent vertex attributes computed by view 0 and shared, views overlapped on 4 internal streams) + pbrcu_shade_ggx_ibl_batched, "
                           f"{n_tris} triangles per view at 1920x1080, 4 point lights + IBL; the scene is the mesh generator's (the reference's bundled .glb is "
                           "missing from its snapshot); glTF -> device import is exercised by tools/gltf_viewer and its tests"}
    for t in list(vplanes.values()) + [vout]:
        scene.keep = [x for x in scene.keep if x is not t]
    del vplanes, vout
    covered = float((planes["normals"].view(torch.int16).view(-1, 4)[:, :3] != 0).any(dim=1).float().mean())
    return {"roofline": roof("hbm (traffic floor; the pass is latency-bound: dependent gathers key -> record -> vertices -> texel)",
                             (W * H * (28 + 12 + 16) + n_tris * (16 + 48)) / (ms_geo * 1e-3) / 1e9, measured_peak_gbs()[0], "GB/s",
                             "per pixel 28 B of planes + 12 B of ORM / emissive written, 8 B visibility word written and read; per triangle a 48 B set-up + 16 B record"),
            "config4_256_views_scene": views_entry,
            "triangles": n_tris, "draws": n_draws, "ms_geometry_pass": round(ms_geo, 4), "mtri_s": round(n_tris / (ms_geo * 1e-3) / 1e6, 1),
            "ms_frame_geometry_plus_ggx_shading": round(ms_frame, 4), "frame_mpix_s": round(W * H / (ms_frame * 1e-3) / 1e6, 1),
            "covered_fraction_last_view": round(covered, 3),
            "what": "pbrcu_geometry_pass (setup, cover, chunk queue, resolve) per view; then + pbrcu_shade_ggx_ibl on the rasterised G-buffer"}


def strong_scaling_block(extra, scene, world):
    """Fixed total work over N GPUs: time at N, time on one GPU and efficiency = t1 / (N tN).  The headline `value` is N
    independent frames (weak scaling, no data shared), so its scaling says nothing about these designs.  t1: measured in
    this run where the unsharded work fits beside the sharded one (8K frame, prefilter), else the committed N = 1 line."""
    if not extra:
        return None
    cached = {}
    try:
        cached = json.loads((ROOT / "profiles" / "r02_bench_n1.json").read_text().strip().splitlines()[-1])["other_configs"]
    except Exception:
        pass
    c3, c4, pf = extra["config3_8k_256_lights"], extra["config4_256_views_1080p"], extra["config2_prefilter_sharded"]

    def entry(t1, tn, src):
        return {"n_gpus": world, "ms": tn, "ms_one_gpu": t1, "efficiency": (round(t1 / (world * tn), 3) if t1 and tn else None), "ms_one_gpu_source": src}

    t1_views = c4["ms_per_pass"] if world == 1 else cached.get("config4_256_views_1080p", {}).get("ms_per_pass")
    out = {"256_views_1080p": entry(t1_views, c4["ms_per_pass"], "this run" if world == 1 else "profiles/r02_bench_n1.json"),
           "8k_256_lights_sharded": entry(c3["ms_per_frame_unsharded_on_one_gpu"], c3["ms_per_frame"], "this run (same frame unsharded on this GPU)"),
           "prefilter_sharded": entry(scene.ibl_ms["prefilter_ggx_prepared"], pf["ms_prefilter_shard_only"], "this run (ibl_precompute_ms.prefilter_ggx_prepared)")}
    for key, name in (("ms_per_frame_with_gather", "8k_256_lights_sharded_nccl_gather"), ("ms_per_frame_fused_gather", "8k_256_lights_sharded_fused_gather_p2p"),
                      ("ms_per_frame_multicast_gather", "8k_256_lights_sharded_fused_gather_multicast")):
        if key in c3:
            out[name] = entry(c3["ms_per_frame_unsharded_on_one_gpu"], c3[key], "this run (same frame unsharded on this GPU)")
    # BASELINE configs[4] as written: the 256 views rasterised and shaded (mesh generator's buffers / the .glb through the viewer)
    geo = extra.get("geometry_pass_1080p", {})
    for key, name in (("config4_256_views_scene", "256_views_1080p_rasterised_and_shaded"), ("config4_256_views_gltf", "256_views_1080p_gltf_rasterised_and_shaded")):
        tn = geo.get(key, {}).get("ms_geometry_plus_shading_all_views")
        if tn:
            t1 = tn if world == 1 else cached.get("geometry_pass_1080p", {}).get(key, {}).get("ms_geometry_plus_shading_all_views")
            out[name] = entry(t1, tn, "this run" if world == 1 else "profiles/r02_bench_n1.json")
    for key, name in (("ms_prefilter_and_gather", "prefilter_sharded_nccl_gather"), ("ms_prefilter_fused_gather", "prefilter_sharded_fused_gather")):
        if key in pf and world > 1:
            out[name] = entry(scene.ibl_ms["prefilter_ggx_prepared"], pf[key], "this run (ibl_precompute_ms.prefilter_ggx_prepared)")
    return out


def mg_host_leg(scene, torch, n_dev):
    """`--mg-host c`: BASELINE configs[3] (7680x4320, 256 clustered lights, row bands) driven by the C multi-GPU host -- ONE process,
    one host thread, n_dev devices (pbrcu_mg_*: csrc/mg.cu) -- instead of one torch.distributed rank per GPU: no torch symmetric
    memory, no NCCL through torch.  Per frame every member clusters and shades its cost-balanced band; the gather is fused into K3's
    stores (P2P), done by grouped ncclBroadcasts, or left out.  Timed with CUDA events on member 0's stream (the call's closing event
    barrier makes it wait for every member), so the host's launch overhead for n_dev devices is inside the number."""
    pbr, lib = scene.pbr, scene.dev.lib
    W8, H8, NL = 7680, 4320, 256
    from pbr_b200.synth import synth_lights
    lights = synth_lights(NL, seed=4, ranged=True)
    mg = C.c_void_p()
    lib.mg_create((C.c_int * n_dev)(*range(n_dev)), n_dev, C.byref(mg))
    keep, gbs, ibls, tiles, frames = [], (pbr.GBuffer * n_dev)(), (pbr.Ibl * n_dev)(), (pbr.LightTiles * n_dev)(), []
    for i in range(n_dev):
        lib.mg_set_device(mg, i)
        ctx = lib.mg_ctx(mg, i)
        lib.gbuffer_alloc(ctx, W8, H8, 1, C.byref(gbs[i]))
        lib.synth_gbuffer(ctx, C.byref(gbs[i]), 1, SEED + 40, pbr.f3(*CAM), pbr.f3(0, 0, 0), None)
        cuda_i = torch.device("cuda", lib.mg_device(mg, i))
        copies = [t.to(cuda_i, copy=True) for t in (scene.irr, scene.pref, scene.lut)]  # every member holds the IBL set
        keep += copies
        ibls[i].irradiance, ibls[i].irr_res = copies[0].data_ptr(), IRR_RES
        ibls[i].prefiltered, ibls[i].pref_res, ibls[i].pref_mips = copies[1].data_ptr(), CUBE_RES, CUBE_MIPS
        ibls[i].lut, ibls[i].lut_res = copies[2].data_ptr(), LUT_RES
        a = C.c_void_p()
        lib.ibl_accel_build(ctx, C.byref(ibls[i]), C.byref(a), None)
        ibls[i].accel = a
        lib.light_tiles_alloc(ctx, W8, H8, NL, C.byref(tiles[i]))
        f = torch.zeros(W8 * H8 * 8, dtype=torch.uint8, device=cuda_i)
        keep.append(f)
        frames.append(f)
        lib.stream_synchronize(None)
    frame_ptrs = (C.c_void_p * n_dev)(*[f.data_ptr() for f in frames])
    # "the previous frame's light lists": one whole-frame clustering on member 0 gives the tile-row costs every band cut derives from
    lib.mg_set_device(mg, 0)
    ctx0 = lib.mg_ctx(mg, 0)
    ld = scene.buf(lights.nbytes) if scene.dev.ordinal == lib.mg_device(mg, 0) else torch.empty(lights.nbytes, dtype=torch.uint8, device=frames[0].device)
    ld.copy_(torch.from_numpy(lights.view(np.uint8).reshape(-1)))
    lib.cluster_lights(ctx0, C.byref(gbs[0]), ld.data_ptr(), NL, C.byref(tiles[0]), None)
    counts = np.empty(tiles[0].tiles_x * tiles[0].tiles_y, np.uint32)
    lib.memcpy_d2h(counts.ctypes.data, tiles[0].counts, counts.nbytes, None)
    lib.stream_synchronize(None)
    costs = (9.0 + counts.reshape(tiles[0].tiles_y, tiles[0].tiles_x)).sum(axis=1).astype(np.float32)
    bands = (C.c_uint32 * (2 * n_dev))()
    lib.mg_row_bands(H8, n_dev, costs.ctypes.data_as(C.POINTER(C.c_float)), bands)
    # the unsharded frame on member 0 (untimed): what every gathered frame must equal bit for bit
    whole = torch.empty(W8 * H8 * 8, dtype=torch.uint8, device=frames[0].device)
    lib.shade_ggx_ibl(ctx0, C.byref(gbs[0]), pbr.f3(*CAM), ld.data_ptr(), NL, C.byref(tiles[0]), C.byref(ibls[0]), pbr.WHOLE, whole.data_ptr(), W8 * 8, None)
    lib.stream_synchronize(None)
    timer = C.c_void_p()
    lib.timer_create(C.byref(timer))
    out = {"n_devices": n_dev, "host": "one process, one host thread, pbrcu_mg_shade_frame (csrc/mg.cu); CUDA events on member 0's stream around 10 frames"}
    modes = [("ms_per_frame_no_gather", 0), ("ms_per_frame_fused_gather_p2p", 1)] + ([("ms_per_frame_nccl_gather", 2)] if n_dev > 1 else [])
    for key, gather in modes:
        try:
            for _ in range(3):
                lib.mg_shade_frame(mg, gbs, pbr.f3(*CAM), lights.ctypes.data, NL, tiles, ibls, bands, frame_ptrs, W8 * 8, gather)
            lib.mg_synchronize(mg)
            lib.mg_set_device(mg, 0)
            lib.timer_start(timer, lib.mg_stream(mg, 0))
            for _ in range(10):
                lib.mg_shade_frame(mg, gbs, pbr.f3(*CAM), lights.ctypes.data, NL, tiles, ibls, bands, frame_ptrs, W8 * 8, gather)
            lib.mg_set_device(mg, 0)
            lib.timer_stop(timer, lib.mg_stream(mg, 0))
            ms = C.c_float()
            lib.timer_elapsed_ms(timer, C.byref(ms))
            lib.mg_synchronize(mg)
            out[key] = round(ms.value / 10, 4)
            if gather:
                out[key.replace("ms_per_frame_", "") + "_equals_unsharded"] = all(bool(torch.equal(f.to(whole.device), whole)) for f in frames)
                for f in frames:
                    f.zero_()
        except pbr.PbrcuError as e:
            out[key] = None
            out[key + "_error"] = str(e)[:160]
    # display-ready frames: the tonemap fused into K3's stores, 4-byte texels on the links (pbrcu_mg_shade_frame_ldr)
    try:
        ldr_frames = [torch.zeros(W8 * H8 * 4, dtype=torch.uint8, device=f.device) for f in frames]
        ldr_ptrs = (C.c_void_p * n_dev)(*[f.data_ptr() for f in ldr_frames])
        lib.mg_set_device(mg, 0)
        whole_ldr = torch.empty(W8 * H8 * 4, dtype=torch.uint8, device=frames[0].device)
        lib.tonemap_aces_gamma(ctx0, whole.data_ptr(), W8 * 8, W8, H8, pbr.WHOLE, whole_ldr.data_ptr(), W8 * 4, None)
        lib.stream_synchronize(None)
        for _ in range(3):
            lib.mg_shade_frame_ldr(mg, gbs, pbr.f3(*CAM), lights.ctypes.data, NL, tiles, ibls, bands, ldr_ptrs, W8 * 4, 1)
        lib.mg_synchronize(mg)
        lib.mg_set_device(mg, 0)
        lib.timer_start(timer, lib.mg_stream(mg, 0))
        for _ in range(10):
            lib.mg_shade_frame_ldr(mg, gbs, pbr.f3(*CAM), lights.ctypes.data, NL, tiles, ibls, bands, ldr_ptrs, W8 * 4, 1)
        lib.mg_set_device(mg, 0)
        lib.timer_stop(timer, lib.mg_stream(mg, 0))
        ms = C.c_float()
        lib.timer_elapsed_ms(timer, C.byref(ms))
        lib.mg_synchronize(mg)
        out["ms_per_frame_fused_gather_p2p_ldr"] = round(ms.value / 10, 4)
        out["fused_gather_p2p_ldr_equals_tonemap_of_unsharded"] = all(bool(torch.equal(f.to(whole_ldr.device), whole_ldr)) for f in ldr_frames)
        out["fused_gather_p2p_ldr_differing_bytes"] = int(sum(int((f.to(whole_ldr.device) != whole_ldr).sum()) for f in ldr_frames))
        del ldr_frames, whole_ldr
    except pbr.PbrcuError as e:
        out["ms_per_frame_fused_gather_p2p_ldr_error"] = str(e)[:160]
    lib.timer_destroy(timer)
    for i in range(n_dev):
        lib.mg_set_device(mg, i)
        ctx = lib.mg_ctx(mg, i)
        lib.light_tiles_free(ctx, C.byref(tiles[i]))
        lib.ibl_accel_free(ctx, ibls[i].accel)
        lib.gbuffer_free(ctx, C.byref(gbs[i]))
    lib.mg_destroy(mg)
    torch.cuda.set_device(scene.dev.ordinal)
    return out



def timed(torch, dist, fn, steps, warmup):
    """W untimed + exactly K timed calls of fn(i), barrier + synchronize on both sides, CUDA events on the
    launching stream; returns max-over-ranks milliseconds for the K steps."""
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        fn(warmup + i)
    b.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = a.elapsed_time(b)
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def host_frame(scene, torch):
    """Pinned host copies of G-buffer set 0 (the e2e inputs) and a pinned result buffer."""
    g, planes = scene.gbufs[0]
    host = {k: torch.empty(t.numel(), dtype=torch.uint8, pin_memory=True) for k, t in planes.items() if k != "depth"}
    for k, t in host.items():
        t.copy_(planes[k])
    out = torch.empty(W4K * H4K * 8, dtype=torch.uint8, pin_memory=True)
    torch.cuda.synchronize()
    return host, out


def run_e2e(scene, torch, dist, host, out, steps, warmup):
    lib, ctx, pbr = scene.dev.lib, scene.dev.ctx, scene.pbr
    lights = scene.lights_host
    cam = pbr.f3(*CAM)

    def step(_):
        lib.shade_frame_host(ctx, host["positions"].data_ptr(), host["normals"].data_ptr(), host["albedo"].data_ptr(),
                             host["orm"].data_ptr(), host["emissive"].data_ptr(), W4K, H4K, cam, lights.ctypes.data,
                             N_LIGHTS, C.byref(scene.ibl), out.data_ptr(), 4, scene.stream())

    return timed(torch, dist, step, steps, warmup)


def cpu_sample(scene, torch, budget_s=12.0):
    """Times the CPU oracle (port of the published formulas; the reference has no GGX/IBL code) on a
    bounded band of rows of the SAME 4K G-buffer and IBL textures, all host threads; also returns the
    GPU-vs-oracle parity of that band."""
    from oracle import Oracle
    o = Oracle()
    g, planes = scene.gbufs[0]
    np_of = lambda t, dt, shape: t.cpu().numpy().view(dt).reshape(shape)
    pos = np_of(planes["positions"], np.float16, (H4K, W4K, 4))
    nrm = np_of(planes["normals"], np.float16, (H4K, W4K, 4))
    alb = np_of(planes["albedo"], np.uint16, (H4K, W4K, 4))
    orm = np_of(planes["orm"], np.uint8, (H4K, W4K, 4))
    emi = np_of(planes["emissive"], np.float16, (H4K, W4K, 4))
    ibl = dict(irr=np_of(scene.irr, np.uint16, (6, IRR_RES, IRR_RES, 4)), pref=np_of(scene.pref, np.uint16, (-1, 4)),
               pref_res=CUBE_RES, pref_mips=CUBE_MIPS, lut=np_of(scene.lut, np.float32, (LUT_RES, LUT_RES, 2)))
    rows, y0 = 16, H4K // 2
    t0 = time.perf_counter()
    o.shade(pos, nrm, alb, orm, emi, CAM, scene.lights_host, ibl, rect=(0, y0, W4K, rows))
    dt = time.perf_counter() - t0
    # bounded sample: a band of the frame (at most its lower half), repeated until about budget_s of CPU work
    rows = int(min(H4K // 2, max(16, rows * budget_s / max(dt, 1e-6))))
    t0 = time.perf_counter()
    _, want = o.shade(pos, nrm, alb, orm, emi, CAM, scene.lights_host, ibl, rect=(0, y0, W4K, rows))
    dt1 = time.perf_counter() - t0
    reps = int(max(1, min(400, budget_s / max(dt1, 1e-6))))
    t0 = time.perf_counter()
    for _ in range(reps):
        o.shade(pos, nrm, alb, orm, emi, CAM, scene.lights_host, ibl, rect=(0, y0, W4K, rows))
    dt = (time.perf_counter() - t0) / reps
    wf = want[y0:y0 + rows].view(np.float16).astype(np.float32)
    peak = float(max(wf.max(), 1.0))

    def compare(exact, rel_bits):
        # default path: irradiance + LUT through the texture unit's 8-bit weights (tolerance 1e-3 + 2^-9 |ref|, two fp16 ulps);
        # exact path: fp32 weights everywhere (1e-3 + 2^-10 |ref|: one fp16 ulp, what the output format forces above 2.0)
        scene.dev.lib.ibl_set_exact_filtering(scene.dev.ctx, exact)
        scene.shade(0)
        torch.cuda.synchronize()
        scene.dev.lib.ibl_set_exact_filtering(scene.dev.ctx, 0)
        gf = np_of(scene.outs[0], np.uint16, (H4K, W4K, 4))[y0:y0 + rows].view(np.float16).astype(np.float32)
        err = np.abs(gf - wf)
        bound = 1e-3 + 2.0 ** -rel_bits * np.abs(wf)
        mse = float(np.mean((gf.astype(np.float64) - wf) ** 2))
        return {"tolerance": f"1e-3 + 2^-{rel_bits}*|ref|", "violations": int((err > bound).sum()), "values": int(err.size), "non_finite": int((~np.isfinite(gf)).sum()),
                "max_abs_err": float(err.max()), "max_err_over_tolerance": round(float((err / bound).max()), 3),
                "psnr_db": (None if mse == 0 else round(10 * np.log10(peak * peak / mse), 1))}

    parity = {"rows": rows, "vs": "oracle/pbr_oracle_ibl.c (no reference counterpart: parity unpinned vs the reference)",
              "default_path": compare(0, 9), "exact_path": compare(1, 10)}
    parity["within"] = parity["default_path"]["violations"] == 0 and parity["exact_path"]["violations"] == 0
    mpix = rows * W4K / dt / 1e6
    return {"value": round(mpix, 3), "unit": "Mpix/s", "cores": o.max_threads(), "kind": "port",
            "sample": f"{reps} passes over {rows} rows x {W4K} px of the 4K frame, {N_LIGHTS} lights + IBL, oracle/pbr_oracle_ibl.c, {dt * reps:.1f} s of CPU work"}, parity


def bind_near_gpu(torch, local):
    """Pins this rank to the NUMA node of its GPU (`partition.bind_host_near_device`); a no-op on one node."""
    from pbr_b200 import partition as part
    pr = torch.cuda.get_device_properties(local)
    try:
        bus = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
    except AttributeError:
        return {"bound": False, "why_not": "torch does not report the PCI address"}
    return part.bind_host_near_device(bus)


def main_ours(args):
    import torch
    import pbr_b200 as pbr

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    torch.cuda.set_device(local)
    host_placement = bind_near_gpu(torch, local)  # before any pinned allocation: first touch decides the node
    if world > 1:
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    dev = pbr.Device(local)  # raises when libpbrcu.so or the GPU is missing: there is no fallback
    scene = Scene(pbr, dev, torch)
    sampler = ClockSampler(local)
    sampler.start()

    ms = timed(torch, dist, scene.shade, args.steps, args.warmup)
    px_step = W4K * H4K
    value = world * px_step * args.steps / (ms * 1e-3) / 1e6
    launch_ms = ms / args.steps  # one K3 launch per step
    peak, peak_src = measured_peak_gbs()
    achieved = B_PER_PX_K3 * px_step / (launch_ms * 1e-3) / 1e9

    # Tier A (reference-pinned) kernels, same protocol
    ms_k1 = timed(torch, dist, scene.lighting, args.steps, args.warmup)
    ms_fused = timed(torch, dist, scene.fused, args.steps, args.warmup)
    # Tier B display frame (GGX + IBL -> RGBA8): K3 then K2, and K3 with the tonemap fused into its stores
    ms_b_two = timed(torch, dist, scene.shade_then_tonemap, args.steps, args.warmup)
    ms_b_ldr = timed(torch, dist, scene.shade_ldr, args.steps, args.warmup)

    host, out = host_frame(scene, torch)
    e2e_steps = max(3, min(args.steps, 20))
    ms_e2e = run_e2e(scene, torch, dist, host, out, e2e_steps, 3)
    e2e_value = world * px_step * e2e_steps / (ms_e2e * 1e-3) / 1e6
    # the reference's own frame (Tier A) through host buffers, same protocol
    ldr_host = torch.empty(W4K * H4K * 4, dtype=torch.uint8, pin_memory=True)

    def step_a_host(_):
        dev.lib.frame_host(dev.ctx, host["positions"].data_ptr(), host["normals"].data_ptr(), host["albedo"].data_ptr(), W4K, H4K, pbr.f3(*CAM),
                           out.data_ptr(), ldr_host.data_ptr(), scene.stream())

    ms_a_e2e = timed(torch, dist, step_a_host, e2e_steps, 3)
    # the floor of the e2e step on this box: every rank moves one frame's bytes both ways AT THE SAME TIME (pinned memory, two
    # streams per rank, no kernels) -- what the PCIe links and the host's memory system deliver to N GPUs at once
    dev_in = scene.buf(px_step * 36)
    host_in = torch.empty(px_step * 36, dtype=torch.uint8, pin_memory=True)
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

    def step_copy(_):
        cur = torch.cuda.current_stream()
        s_in.wait_stream(cur); s_out.wait_stream(cur)
        with torch.cuda.stream(s_in):
            dev_in.copy_(host_in, non_blocking=True)
        with torch.cuda.stream(s_out):
            out.copy_(scene.outs[0], non_blocking=True)
        cur.wait_stream(s_in); cur.wait_stream(s_out)

    ms_floor = timed(torch, dist, step_copy, e2e_steps, 3) / e2e_steps
    clocks = sampler.stop()

    extra = None if args.no_extra else extra_configs(scene, torch, dist, world, rank)

    mg_host = None
    if args.mg_host == "c" and not args.no_extra:
        flag = Path(os.environ.get("TMPDIR", "/tmp")) / f"pbr_bench_mg_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
        if rank == 0:
            try:
                mg_host = mg_host_leg(scene, torch, min(world, torch.cuda.device_count()))
            except Exception as e:  # the leg must not take the headline down with it
                mg_host = {"error": f"{type(e).__name__}: {e}"[:300]}
            if world > 1:
                flag.write_text("done")
        else:  # the other ranks keep their GPUs idle until rank 0 is through (a host-side wait, no spinning collective)
            t_wait = time.time()
            while not flag.exists() and time.time() - t_wait < 900:
                time.sleep(0.05)
        if dist is not None:
            dist.barrier()
            if rank == 0:
                flag.unlink(missing_ok=True)
    cpu, parity = (None, None)
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu, parity = cpu_sample(scene, torch)

    if rank == 0:
        k1_gbs = B_PER_PX_K1 * px_step * args.steps / (ms_k1 * 1e-3) / 1e9
        line = {
            "metric": "shaded Mpix/s at 4K GGX+IBL", "value": round(value, 1), "unit": "Mpix/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 5),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_block(world),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "traffic": K3_DRAM_TRAFFIC_BYTES, "traffic_unit": "bytes per launch (ncu, profiles/)",
                         "algorithmic_bytes_per_launch": B_PER_PX_K3 * px_step, "peak_source": peak_src,
                         "kernel": K3_KERNEL, "algorithmic_bytes_per_px": B_PER_PX_K3},
            "e2e": {"value": round(e2e_value, 1), "unit": "Mpix/s", "h2d_bytes_per_step": px_step * 36,
                    "d2h_bytes_per_step": px_step * 8, "steps": e2e_steps, "api": "pbrcu_shade_frame_host (pinned host planes, 4 pipelined row bands, tapered 8:4:2:1)",
                    "host_placement": host_placement, "ms_per_step": round(ms_e2e / e2e_steps, 4), "floor_ms": round(ms_floor, 4),
                    "floor": f"{world} rank(s) copying 299 MB in and 66 MB out concurrently from / to pinned memory, no kernels (max over ranks)"},
            "gpu_launches": args.steps,
            "clocks": clocks,
            "cpu_baseline": cpu,
            "parity": parity,
            "tier_a": {"note": "reference-pinned passes (lighting.glsl / aces+gamma.glsl) at 4K, same protocol",
                       "lighting_mpix_s": round(world * px_step * args.steps / (ms_k1 * 1e-3) / 1e6, 1),
                       "lighting_gbs": round(k1_gbs, 1), "lighting_roofline_frac": round(k1_gbs / peak, 4),
                       "fused_ldr_mpix_s": round(world * px_step * args.steps / (ms_fused * 1e-3) / 1e6, 1),
                       "fused_ldr_roofline_frac": round(28 * px_step * args.steps / (ms_fused * 1e-3) / 1e9 / peak, 4),
                       "frame_host_mpix_s": round(world * px_step * e2e_steps / (ms_a_e2e * 1e-3) / 1e6, 1),
                       "frame_host": "pbrcu_frame_host: three pinned host planes in (24 B/px), lighting + ACES tonemap, HDR and LDR frames out (12 B/px), 4 tapered row bands"},
            "display_frame_ggx_ibl_4k": {"what": "GGX + 4 lights + IBL to a display-ready RGBA8 frame: pbrcu_shade_ggx_ibl + pbrcu_tonemap_aces_gamma (44 + 12 B/px) vs "
                                                 "pbrcu_shade_ggx_ibl_ldr (tonemap fused into K3's stores, 40 B/px), bit-identical",
                                         "ms_two_passes": round(ms_b_two / args.steps, 5), "ms_fused": round(ms_b_ldr / args.steps, 5),
                                         "mpix_s_fused": round(world * px_step * args.steps / (ms_b_ldr * 1e-3) / 1e6, 1)},
            "other_configs": extra,
            "strong_scaling": strong_scaling_block(extra, scene, world),
            "mg_host_c": mg_host,
            "ibl_precompute_ms": scene.ibl_ms,
            "ibl_precompute_roofline": scene.ibl_roofline,
            "prefilter_error": scene.prefilter_error,
        }
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------
# CPU arm (--impl reference)
# ---------------------------------------------------------------------------------------------------
def main_reference(args):
    """The reference's CPU side of the path on the box's host cores.  For GGX+IBL the reference has no
    implementation at all, so this arm times the oracle port; the reference's own shaders
    (oracle/_ref, lighting.glsl compiled for the CPU) are timed beside it as `tier_a`."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from oracle import Oracle, RefShaders, synth_environment, synth_gbuffer, synth_lights
    o = Oracle()
    # every core this process may run on -- NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1 to its workers
    threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    rows_total = H4K  # the whole 4K frame is generated; each step shades a bounded band of it
    g = synth_gbuffer(W4K, rows_total, seed=SEED & 0xFFFF, background=0.25, tier_b=True)
    lights = synth_lights(N_LIGHTS, seed=2)
    # IBL textures of the GPU arm's SHAPES (2048^2 x 12 mips, 64^2, 512^2): what the shading pass costs depends on the shapes, not on
    # how the texels were made, so they are filled cheaply -- the box-filtered chain of the environment stands in for the
    # prefiltered chain, irradiance and LUT take 32 samples (the full-sample precompute would take the CPU many minutes)
    env = synth_environment(1024, 512, 1)
    chain = o.make_chain(o.equirect_to_cube(env, CUBE_RES, nthreads=threads), CUBE_MIPS, nthreads=threads)
    ibl = dict(irr=o.irradiance(chain, CUBE_RES, CUBE_MIPS, IRR_RES, 32, nthreads=threads), pref=chain, pref_res=CUBE_RES, pref_mips=CUBE_MIPS,
               lut=o.brdf_lut(LUT_RES, 32, nthreads=threads))

    def shade(rows):
        o.shade(g["pos"], g["nrm"], g["alb"], g["orm"], g["emi"], CAM, lights, ibl, rect=(0, 0, W4K, rows), nthreads=threads)

    t0 = time.perf_counter(); shade(8); dt8 = time.perf_counter() - t0
    per_step = min(2.0, 150.0 / (args.steps + args.warmup))
    rows = int(max(1, min(rows_total, 8 * per_step / max(dt8, 1e-6))))
    for _ in range(args.warmup):
        shade(rows)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        shade(rows)
    dt = time.perf_counter() - t0
    value = rows * W4K * args.steps / dt / 1e6
    sample = (f"{rows} rows x {W4K} px per step of the 4K synthetic G-buffer, {N_LIGHTS} lights + IBL textures of the GPU arm's shapes "
              f"({CUBE_RES}^2 x {CUBE_MIPS} mips, {IRR_RES}^2, {LUT_RES}^2; cheap CPU fill), {threads} host threads")
    tier_a = None
    try:
        r = RefShaders()
        t0 = time.perf_counter(); r.lighting(g["pos"], g["nrm"], g["alb"], CAM, nthreads=threads); d = time.perf_counter() - t0
        tier_a = {"lighting_mpix_s": round(rows_total * W4K / d / 1e6, 2), "kind": "reference",
                  "what": f"assets/shaders/pbr/lighting.glsl compiled for the CPU (oracle/_ref), {threads} host threads"}
    except Exception as e:  # oracle/_ref not built
        tier_a = {"unavailable": str(e)}
    line = {
        "impl": "reference", "metric": "shaded Mpix/s at 4K GGX+IBL", "value": round(value, 3), "unit": "Mpix/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_block(args.gpus),  # the GPU arm's configuration; the CPU steps over a bounded sample of it (cpu_baseline.sample)
        "cpu_baseline": {"value": round(value, 3), "unit": "Mpix/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": round(value, 3), "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "tier_a": tier_a,
        "note": "the reference has no GGX/IBL implementation and its Vulkan path is unavailable on this box; CPU arm = oracle port",
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--mg-host", choices=["none", "c"], default="c",
                    help="c: also run configs[3] through the C multi-GPU host (pbrcu_mg_*), one process over the run's GPUs (rank 0)")
    ap.add_argument("--no-extra", action="store_true", help="skip the config[3]/[4] measurements (profiling runs)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        main_reference(args)
    else:
        main_ours(args)


if __name__ == "__main__":
    main()
