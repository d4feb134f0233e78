#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L | head -3
timeout 1200 python -m pytest tests/test_gpu_mg.py tests/test_gpu_tier_b.py -m gpu -x -q > gpurun_out/pytest_mg.txt 2>&1; echo "mg+tierb exit $?"
tail -25 gpurun_out/pytest_mg.txt
