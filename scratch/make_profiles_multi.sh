#!/bin/bash
# The scaling lines of profiles/: gpurun --gpus N -- 'bash scratch/make_profiles_multi.sh'  (N = 2, 4, 8; profiles/r02_bench_n1.json must exist:
# the strong_scaling block reads the one-GPU time of the 256 views from it).  At N = 8 the 8K frame is also run with 8-byte peer stores only.
N=$(nvidia-smi -L | wc -l)
O=gpurun_out/prof
mkdir -p $O
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench n$N exit $?"
[ "${PROFILES_LIGHT:-0}" = "1" ] && { tail -c 300 $O/bench_n$N.err; exit 0; }   # the reference arm and the 8-byte-store run only change with their code
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --impl reference --gpus $N --steps 5 --warmup 3 > $O/bench_reference_n$N.json 2> $O/bench_reference_n$N.err; echo "reference n$N exit $?"
if [ $N -ge 8 ]; then
  PBRCU_PAIR_STORES=0 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29525 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu > $O/bench_n${N}_8byte_stores.json 2> $O/bench_n${N}_8byte_stores.err; echo "8-byte stores exit $?"
fi
tail -c 600 $O/bench_n$N.err
