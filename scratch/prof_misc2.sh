#!/bin/bash
# K3c + the tiled K3 kernel on the 8K / 256-light frame + K8 (the first 24 launches matched by make_profiles.sh's misc pass are all K4 / K5)
O=gpurun_out/prof
mkdir -p $O
CMD3="python bench.py --steps 3 --warmup 3 --no-cpu --mg-host none"
$CMD3 > $O/plain_misc2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'cluster_lights_kernel|shade_ggx_kernel|brdf_lut_kernel' -c 7 -o $O/misc2 $CMD3 > $O/ncu_misc2.log 2>&1; echo "ncu misc2 exit $?"
ncu -i $O/misc2.ncu-rep --page raw --csv > $O/misc2.raw.csv 2>/dev/null
ncu -i $O/misc2.ncu-rep --page source --csv --print-source sass --kernel-name regex:shade_ggx_kernel > $O/misc2.src.csv 2>/dev/null
rm -f $O/misc2.ncu-rep; ls -la $O | grep misc2
