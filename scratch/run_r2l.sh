#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scratch/batched_ab.py > gpurun_out/batched_ab.txt 2>&1; echo "exit $?"; tail -8 gpurun_out/batched_ab.txt
