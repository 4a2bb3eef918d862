#!/bin/bash
# Round-end records on ONE GPU (run on the GPU box through gpurun): bench line (venice), reference arm (short budget), ncu launch
# list of the bench command, one `ncu --set full` capture of the hot kernels, per-kernel DRAM traffic for the bench's `traffic`.
#   bash scripts/r02_profile.sh <tag>        -> gpurun_out/<tag>/...
set -x
R=${1:-r02}
O=gpurun_out/$R
mkdir -p $O
timeout 900 python bench.py > $O/bench_venice.json 2> $O/bench_venice.err
python scripts/bench_summary.py $O/bench_venice.json
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 --ref-budget-s 150 > $O/bench_venice_ref.json 2> $O/bench_venice_ref.err
cut -c1-400 $O/bench_venice_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_venice.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --extras none > $O/ncu_launches.log 2>&1
python scripts/launch_summary.py $O/launches_venice.csv > $O/launches_venice_summary.txt; cat $O/launches_venice_summary.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_tile_pass|k_tile_lin|k_pcg_tail|k_camera_pass|k_point_prepare' \
    -c 14 -f -o $O/hot_kernels python scripts/profile_target.py venice 3 > $O/ncu_full.log 2>&1
python scripts/ncu_summary.py $O/hot_kernels.ncu-rep $O/ncu_hot_kernels.csv | tee $O/ncu_hot_kernels.txt
ls -la $O
