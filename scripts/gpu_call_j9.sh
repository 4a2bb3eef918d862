#!/bin/bash
set -x
O=gpurun_out/r01j
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_j9.txt 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_j9.txt
tail -3 $O/pytest_j9.txt
python scripts/profile_window.py | tail -16
timeout 120 python scripts/run_latency.py > $O/run_latency_j9.txt 2>&1; head -4 $O/run_latency_j9.txt
timeout 200 python scripts/ab_solve.py bal50 - > $O/ab_j9_bal50.txt 2>&1; cat $O/ab_j9_bal50.txt
timeout 200 python scripts/ab_solve.py ladybug - > $O/ab_j9_ladybug.txt 2>&1; cat $O/ab_j9_ladybug.txt
