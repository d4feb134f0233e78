#!/bin/bash
# first GPU call of round 2: tld4 probe, the K3 A/B, the GPU test-suite
mkdir -p gpurun_out
scratch/bin/gather_probe > gpurun_out/gather_probe.txt 2>&1
cat gpurun_out/gather_probe.txt
timeout 900 python scratch/k3_cl_exp.py > gpurun_out/k3_cl_exp.txt 2>&1; echo "k3_cl_exp exit $?"
tail -30 gpurun_out/k3_cl_exp.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest exit $?"
tail -15 gpurun_out/pytest_gpu.txt
