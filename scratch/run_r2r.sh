#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_tier_b.py tests/test_gpu_closed_form.py -m gpu -x -q > gpurun_out/pytest_flags.txt 2>&1; echo "tests exit $?"; tail -4 gpurun_out/pytest_flags.txt
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu --no-extra > gpurun_out/bench_flags.json 2> gpurun_out/bench_flags.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_flags.json").read().strip().splitlines()[-1])
print(d["ibl_precompute_ms"]); print(d["prefilter_error"])
PY
