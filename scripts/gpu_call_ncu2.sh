#!/bin/bash
set -x
O=gpurun_out/r01j
mkdir -p $O
cap() { # name, kernel regex, skip, count
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c $4 -f -o $O/$1 python scripts/profile_target.py venice 4 > $O/ncu_$1.log 2>&1
  python scripts/ncu_summary.py $O/$1.ncu-rep $O/ncu_$1.csv
}
cap tile_lin k_tile_lin 1 1
cap camera_pass k_camera_pass 0 2
cap pcg_tail k_pcg_tail 1 1
cap backsub 'k_tile_pass<1' 0 1
