#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tier_b.py tests/test_gpu_mg.py tests/test_gpu_multicast.py -m gpu -x -q > gpurun_out/pytest_pairs.txt 2>&1; echo "tests exit $?"; tail -5 gpurun_out/pytest_pairs.txt
N=$(nvidia-smi -L | wc -l)
for PS in 0 1; do
PBRCU_PAIR_STORES=$PS timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu > gpurun_out/bench_pairs_$PS.json 2> gpurun_out/bench_pairs_$PS.err; echo "bench pairs=$PS exit $?"
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_pairs_$PS.json").read().strip().splitlines()[-1])
c = d["other_configs"]["config3_8k_256_lights"]
print("pairs=$PS", {k: c.get(k) for k in ("ms_per_frame", "ms_per_frame_with_gather", "ms_per_frame_fused_gather", "ms_per_frame_multicast_gather", "fused_gathered_frame_equals_unsharded", "multicast_frame_equals_p2p_frame")})
print("   mg_host_c", d["mg_host_c"])
PY
done
