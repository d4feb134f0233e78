#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scratch/ldr_diag.py > gpurun_out/ldr_diag.txt 2>&1; echo "diag exit $?"; tail -4 gpurun_out/ldr_diag.txt
timeout 1800 python -m pytest tests/test_gpu_tier_b.py tests/test_gpu_tier_a.py tests/test_gpu_facade.py tests/test_gpu_viewer.py -m gpu -x -q > gpurun_out/pytest_ldr3.txt 2>&1; echo "tests exit $?"; tail -5 gpurun_out/pytest_ldr3.txt
