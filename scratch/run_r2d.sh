#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "suite exit $?"
tail -6 gpurun_out/pytest_gpu.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.txt 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.txt
timeout 900 python bench.py --steps 50 --warmup 10 --no-extra > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "bench exit $?"; tail -3 gpurun_out/bench_short.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_short.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step", "roofline", "e2e", "parity", "cpu_baseline")})
PY
python bench.py --steps 3 --warmup 3 --no-cpu --no-extra > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:shade_ggx_cl_kernel -s 3 -c 2 -o gpurun_out/r02_k3_cl python bench.py --steps 3 --warmup 3 --no-cpu --no-extra > gpurun_out/ncu.log 2>&1; echo "ncu exit $?"; tail -3 gpurun_out/ncu.log
