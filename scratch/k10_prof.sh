#!/bin/bash
# K10 after the narrow edge functions and the new launch bounds: tests, plain run, then one ncu capture of the second pass's kernels
set -x
timeout 600 python -m pytest tests/test_gpu_raster.py -x -q 2>&1 | tail -2
timeout 600 python scratch/prof_k10.py > gpurun_out/k10_plain.log 2>&1 || { tail -5 gpurun_out/k10_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:raster_ -s 7 -c 7 -o gpurun_out/k10_r2 -f python scratch/prof_k10.py > gpurun_out/k10_ncu.log 2>&1
ncu -i gpurun_out/k10_r2.ncu-rep --page raw --csv > gpurun_out/k10_r2_raw.csv 2>/dev/null
ncu -i gpurun_out/k10_r2.ncu-rep --page source --csv --kernel-name regex:raster_setup > gpurun_out/k10_r2_setup_source.csv 2>/dev/null
ncu -i gpurun_out/k10_r2.ncu-rep --page source --csv --kernel-name regex:raster_resolve > gpurun_out/k10_r2_resolve_source.csv 2>/dev/null
ncu -i gpurun_out/k10_r2.ncu-rep --page source --csv --kernel-name regex:raster_cover > gpurun_out/k10_r2_cover_source.csv 2>/dev/null
ls -la gpurun_out/k10_r2*; rm -f gpurun_out/k10_r2.ncu-rep
