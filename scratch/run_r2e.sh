#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"; tail -2 gpurun_out/bench_ref.err
timeout 1500 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; tail -5 gpurun_out/bench_n1.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_n1.json").read().strip().splitlines()[-1])
for k in ("value", "ms_per_step", "roofline", "e2e", "parity", "cpu_baseline", "strong_scaling", "ibl_precompute_ms", "ibl_precompute_roofline", "prefilter_error"):
    print(k, json.dumps(d[k])[:900])
for k, v in d["other_configs"].items():
    print(k, json.dumps(v)[:1200])
r = json.loads(open("gpurun_out/bench_ref.json").read().strip().splitlines()[-1])
print("REF", r["value"], r["cpu_baseline"])
PY
