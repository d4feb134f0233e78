"""bench.gltf_views_bench on its own (one GPU)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
print(json.dumps(bench.gltf_views_bench(torch, None, 1, 0), indent=1))
