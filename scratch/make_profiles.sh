#!/bin/bash
# Regenerates every round-2 file under profiles/ that comes from ONE GPU, from the build in the tree:
#     gpurun --timeout 3000 -- 'bash scratch/make_profiles.sh'      (then: bash scratch/profiles_collect.sh here, where ncu -i works too)
# Writes raw material to gpurun_out/prof/ (bench lines, the ncu launch list, .ncu-rep captures).  Each ncu command runs only after the
# same command exited 0 without ncu.  Numbers printed under ncu are never bench values.
set -u
O=gpurun_out/prof
mkdir -p $O
python bench.py --impl reference > $O/bench_reference.json 2> $O/bench_reference.err; echo "reference arm exit $?"
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench n1 exit $?"
CMD="python bench.py --steps 5 --warmup 3 --no-cpu --mg-host none"
$CMD > $O/plain_launches.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1; echo "launch list exit $?"
# K10 (one pass of the bench scene through pbrcu_geometry_pass; scratch/prof_k10.py)
python scratch/prof_k10.py > $O/plain_k10.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:raster_ -s 7 -c 7 -o $O/k10 python scratch/prof_k10.py > $O/ncu_k10.log 2>&1; echo "ncu k10 exit $?"
[ -f $O/k10.ncu-rep ] && ncu -i $O/k10.ncu-rep --page raw --csv > $O/k10.raw.csv 2>/dev/null
rm -f $O/k10.ncu-rep
if [ "${PROFILES_LIGHT:-0}" = "1" ]; then ls -la $O; exit 0; fi   # the K3 / K7 / misc captures only change when those kernels do
CMD2="python bench.py --steps 3 --warmup 3 --no-cpu --no-extra"
$CMD2 > $O/plain_k3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:shade_ggx_cl_kernel -s 3 -c 2 -o $O/k3_cl $CMD2 > $O/ncu_k3.log 2>&1; echo "ncu k3 exit $?"
PBRCU_BENCH_GATHER=1 $CMD2 > $O/plain_k3_gather.log 2>&1 &&
PBRCU_BENCH_GATHER=1 ncu --set full --clock-control none --import-source on -k regex:shade_ggx_cl_kernel -s 3 -c 1 -o $O/k3_cl_gather $CMD2 > $O/ncu_k3g.log 2>&1; echo "ncu k3 gather exit $?"
$CMD2 > $O/plain_k7.log 2>&1 &&
ncu --set full --clock-control none -k regex:conv_kernel -c 14 -o $O/k7 $CMD2 > $O/ncu_k7.log 2>&1; echo "ncu k7 exit $?"
CMD3="python bench.py --steps 3 --warmup 3 --no-cpu --mg-host none"
$CMD3 > $O/plain_misc.log 2>&1 &&
ncu --set full --clock-control none -k regex:'equirect_to_cube_kernel|cube_mip_kernel' -c 24 -o $O/misc $CMD3 > $O/ncu_misc.log 2>&1; echo "ncu misc exit $?"
bash scratch/prof_misc2.sh   # K3c, the tiled K3 kernel on the 8K frame, K8 -> misc2.raw.csv (profiles/r02_k3_tiled_8k_ncu.txt)
# the reports are too large to travel (64 MiB limit on gpurun_out/): export what profiles_collect.sh reads, keep only the K3 reports
for rep in k3_cl k3_cl_gather k7 misc; do
  [ -f $O/$rep.ncu-rep ] || continue
  ncu -i $O/$rep.ncu-rep --page raw --csv > $O/$rep.raw.csv 2>/dev/null
done
for rep in k3_cl k3_cl_gather; do
  [ -f $O/$rep.ncu-rep ] && ncu -i $O/$rep.ncu-rep --page source --csv --print-source sass > $O/$rep.src.csv 2>/dev/null
done
rm -f $O/k7.ncu-rep $O/misc.ncu-rep
rm -f gpurun_out/*.ncu-rep
ls -la $O
