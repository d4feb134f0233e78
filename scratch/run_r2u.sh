#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_mg.py tests/test_gpu_facade.py -m gpu -x -q > gpurun_out/pytest_mg2.txt 2>&1; echo "exit $?"; tail -12 gpurun_out/pytest_mg2.txt
