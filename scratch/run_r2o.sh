#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_tier_b.py tests/test_gpu_tier_a.py -m gpu -x -q > gpurun_out/pytest_ldr2.txt 2>&1; echo "exit $?"; tail -8 gpurun_out/pytest_ldr2.txt
bash scratch/run_r2n.sh
