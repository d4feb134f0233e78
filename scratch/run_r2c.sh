#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python scratch/k3_cl_exp.py > gpurun_out/k3_cl_exp2.txt 2>&1; echo "k3_cl_exp exit $?"
grep -v "^nl=" gpurun_out/k3_cl_exp2.txt | tail -40
