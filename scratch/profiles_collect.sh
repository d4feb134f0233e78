#!/bin/bash
# Turns gpurun_out/prof/ (scratch/make_profiles.sh) into the committed summaries under profiles/ (run here: ncu -i needs no GPU).
set -u
O=gpurun_out/prof
P=profiles
tail -1 $O/bench_n1.json > $P/r02_bench_n1.json
tail -1 $O/bench_reference.json > $P/r02_bench_reference.json
cp $O/launches.csv $P/r02_launches_bench.csv
python scratch/launch_summary.py $O/launches.csv > $P/r02_launches_summary.txt
for rep in k3_cl k3_cl_gather; do
  [ -f $O/$rep.raw.csv ] && python scratch/summarize_ncu.py $O/$rep.raw.csv $O/$rep.src.csv > $P/r02_${rep}_ncu.txt
done
if [ -f $O/k10.raw.csv ]; then
  { echo "# ncu --set full --clock-control none --import-source on -k regex:raster_ -s 7 -c 7 python scratch/prof_k10.py   (scratch/make_profiles.sh)";
    echo "# (second single-view geometry pass of the bench scene: 589,828 triangles in 65 draws at 1920x1080; final kernels of round 2; summary by scratch/summarize_ncu.py)";
    python scratch/summarize_ncu.py $O/k10.raw.csv "" 30; } > $P/r02_k10_ncu.txt
fi
for rep in k7 misc; do
  [ -f $O/$rep.raw.csv ] && python scratch/summarize_ncu.py $O/$rep.raw.csv "" 30 > $P/r02_${rep}_ncu.txt
done
ls -la $P | grep r02
