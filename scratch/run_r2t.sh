#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scratch/k3_pipe_exp.py > gpurun_out/k3_pipe_exp.txt 2>&1; echo "exit $?"; tail -10 gpurun_out/k3_pipe_exp.txt
