#!/bin/bash
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
if [ $N -gt 1 ]; then
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29527 bench.py --gpus $N --steps 50 --warmup 10 --no-cpu > gpurun_out/bench_ldr_n$N.json 2> gpurun_out/bench_ldr_n$N.err; echo "bench exit $?"
else
  timeout 1500 python bench.py --steps 100 --warmup 10 --no-cpu > gpurun_out/bench_ldr_n$N.json 2> gpurun_out/bench_ldr_n$N.err; echo "bench exit $?"
fi
tail -3 gpurun_out/bench_ldr_n$N.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_ldr_n$N.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["display_frame_ggx_ibl_4k"])
print(d["mg_host_c"])
PY
