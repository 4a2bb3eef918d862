#!/bin/bash
# round-end records: bench line (venice, 1 GPU), reference arm, ncu launch list of the bench command, one full ncu capture of
# the three kernels of the PCG / linearisation hot path
set -x
R=${1:-r01j}
O=gpurun_out/$R
mkdir -p $O
timeout 600 python bench.py > $O/bench_venice.json 2> $O/bench_venice.err
python scripts/bench_summary.py $O/bench_venice.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_venice_ref.json 2> $O/bench_venice_ref.err
cat $O/bench_venice_ref.json | cut -c1-300
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/launches_venice.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_tile_pass|k_tile_lin|k_pcg_tail' -s 6 -c 6 -f -o $O/tile_kernels \
    python scripts/profile_target.py venice 4 > $O/ncu_full.log 2>&1
python scripts/ncu_summary.py $O/tile_kernels.ncu-rep $O/ncu_tile_kernels.csv
ls -la $O
