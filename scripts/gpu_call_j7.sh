#!/bin/bash
set -x
O=gpurun_out/r01j
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_devloop.txt 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_devloop.txt
tail -5 $O/pytest_devloop.txt
timeout 300 python scripts/ab_solve.py venice - BA_B200_HOST_LOOP=1 - BA_B200_HOST_LOOP=1 > $O/ab_devloop_venice.txt 2>&1
grep "solve ms" $O/ab_devloop_venice.txt
timeout 200 python scripts/ab_solve.py bal50 - BA_B200_HOST_LOOP=1 - > $O/ab_devloop_bal50.txt 2>&1
grep "solve ms" $O/ab_devloop_bal50.txt
timeout 200 python scripts/ab_solve.py ladybug - BA_B200_HOST_LOOP=1 - > $O/ab_devloop_ladybug.txt 2>&1
grep "solve ms" $O/ab_devloop_ladybug.txt
timeout 120 python scripts/run_latency.py > $O/run_latency_devloop.txt 2>&1
BA_B200_HOST_LOOP=1 timeout 120 python scripts/run_latency.py > $O/run_latency_hostloop.txt 2>&1
head -4 $O/run_latency_devloop.txt; head -4 $O/run_latency_hostloop.txt
