#!/bin/bash
set -x
O=gpurun_out/r01j
mkdir -p $O
timeout 600 python bench.py > $O/bench_venice.json 2> $O/bench_venice.err
python scripts/bench_summary.py $O/bench_venice.json | head -3
timeout 200 python bench.py --workload bal50 > $O/bench_bal50.json 2> $O/bench_bal50.err
timeout 300 python bench.py --workload ladybug > $O/bench_ladybug.json 2> $O/bench_ladybug.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/launches_venice.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_launches.log 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; tail -2 $O/smoke.txt
