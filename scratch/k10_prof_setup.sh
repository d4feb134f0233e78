#!/bin/bash
set -x
timeout 600 python scratch/prof_k10.py > gpurun_out/k10_plain.log 2>&1 || { tail -5 gpurun_out/k10_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:raster_setup -s 1 -c 1 -o gpurun_out/k10_s -f python scratch/prof_k10.py > gpurun_out/k10_ncu.log 2>&1
ncu -i gpurun_out/k10_s.ncu-rep --page raw --csv > gpurun_out/k10_s_raw.csv 2>/dev/null
ncu -i gpurun_out/k10_s.ncu-rep --page source --csv > gpurun_out/k10_s_source.csv 2>/dev/null
rm -f gpurun_out/k10_s.ncu-rep
