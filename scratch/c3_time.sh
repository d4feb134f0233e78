#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_tier_b.py tests/test_gpu_closed_form.py -x -q 2>&1 | tail -2
python bench.py --steps 5 --warmup 3 --no-cpu --mg-host none > gpurun_out/c3.json 2> gpurun_out/c3.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/c3.json').read().strip().splitlines()[-1])
c3=d["other_configs"]["config3_8k_256_lights"]
print("config3 ms", c3["ms_per_frame"], "fp32 frac", c3["roofline"]["frac"], "| headline us", d["ms_per_step"]*1e3, "| 256 views ms", d["other_configs"]["config4_256_views_1080p"]["ms_per_pass"])
PY
