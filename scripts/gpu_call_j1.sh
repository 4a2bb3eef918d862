#!/bin/bash
# round 1j, call 1: A/B of the PDL / prefetch variants, GPU test-suite with defaults and with the variants on,
# bench records of the two small single-GPU configurations
set -x
O=gpurun_out/r01j
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/smi.txt 2>&1
timeout 300 python scripts/ab_solve.py venice - BA_B200_CM_PREFETCH=1 BA_B200_CM_PREFETCH=2 BA_B200_PDL=1 BA_B200_PDL=1,BA_B200_CM_PREFETCH=2 - > $O/ab_venice.txt 2>&1
tail -20 $O/ab_venice.txt
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_default.txt 2>&1; echo "pytest default rc=$?" | tee -a $O/pytest_default.txt
tail -3 $O/pytest_default.txt
BA_B200_PDL=1 BA_B200_CM_PREFETCH=2 timeout 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "matvec or iterates or bitwise or converged or full_size or normal_equation" > $O/pytest_variants.txt 2>&1; echo "pytest variants rc=$?" | tee -a $O/pytest_variants.txt
tail -3 $O/pytest_variants.txt
timeout 200 python bench.py --workload bal50 > $O/bench_bal50.json 2> $O/bench_bal50.err
timeout 300 python bench.py --workload ladybug > $O/bench_ladybug.json 2> $O/bench_ladybug.err
timeout 120 python scripts/run_latency.py > $O/run_latency.txt 2>&1
tail -5 $O/run_latency.txt
