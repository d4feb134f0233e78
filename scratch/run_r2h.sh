#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q -rs > gpurun_out/pytest_gpu.txt 2>&1; echo "suite exit $?"
tail -15 gpurun_out/pytest_gpu.txt
