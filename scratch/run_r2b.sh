#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.txt 2>&1; echo "alone exit $?"
tail -5 gpurun_out/pytest_noalloc_alone.txt
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "suite exit $?"
tail -25 gpurun_out/pytest_gpu.txt
