#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_tier_b.py tests/test_gpu_mg.py tests/test_gpu_closed_form.py -m gpu -x -q > gpurun_out/pytest_ldr.txt 2>&1; echo "exit $?"; tail -12 gpurun_out/pytest_ldr.txt
