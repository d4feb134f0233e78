#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 50 --warmup 10 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 exit $?"; tail -5 gpurun_out/bench_n2.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_n2.json").read().strip().splitlines()[-1])
for k in ("value", "ms_per_step", "e2e", "strong_scaling", "mg_host_c"):
    print(k, json.dumps(d[k])[:1500])
print(json.dumps(d["other_configs"]["config3_8k_256_lights"])[:1500])
print(json.dumps(d["other_configs"]["config2_prefilter_sharded"])[:900])
PY
