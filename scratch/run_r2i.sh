#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_raster.py tests/test_gpu_tier_a.py -m gpu -x -q > gpurun_out/pytest_raster.txt 2>&1; echo "raster+tierA exit $?"; tail -5 gpurun_out/pytest_raster.txt
timeout 1500 python bench.py --steps 50 --warmup 10 --no-cpu > gpurun_out/bench_n1b.json 2> gpurun_out/bench_n1b.err; echo "bench exit $?"; tail -5 gpurun_out/bench_n1b.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_n1b.json").read().strip().splitlines()[-1])
print(json.dumps(d["other_configs"]["geometry_pass_1080p"])[:2500])
print(json.dumps(d["mg_host_c"])[:600])
print(d["value"], d["other_configs"]["config4_256_views_1080p"]["ms_per_pass"])
PY
