"""config 3 (8K, 256 clustered lights) exact (fp32 weights everywhere) vs default (irradiance + LUT through the texture unit)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench, pbr_b200 as pbr
dev = pbr.Device(0); scene = bench.Scene(pbr, dev, torch)
for v in (1, 0, 1, 0):
    dev.lib.ibl_set_exact_filtering(dev.ctx, v)
    ex = bench.extra_configs(scene, torch, None, 1, 0)
    print("exact", v, ex["config3_8k_256_lights"]["ms_per_frame"], ex["config4_256_views_1080p"]["ms_per_pass"], flush=True)
