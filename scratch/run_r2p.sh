#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scratch/ldr_diag.py > gpurun_out/ldr_diag.txt 2>&1; echo "exit $?"; tail -12 gpurun_out/ldr_diag.txt
