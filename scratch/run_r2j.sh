#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scratch/k3_l2_exp.py > gpurun_out/k3_l2_exp.txt 2>&1; echo "exit $?"; tail -12 gpurun_out/k3_l2_exp.txt
