"""config 3 (8K, 256 clustered lights) with the tiled-list kernel variants: PBRCU_K3_TILED = 0 (old kernel), 4, 5 CTAs/SM."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench, pbr_b200 as pbr
dev = pbr.Device(0); scene = bench.Scene(pbr, dev, torch)
for v in sys.argv[1:] or ["0", "4", "5"]:
    os.environ["PBRCU_K3_TILED"] = v
    ex = bench.extra_configs(scene, torch, None, 1, 0)
    print("tiled variant", v, json.dumps(ex["config3_8k_256_lights"])[:160], flush=True)
