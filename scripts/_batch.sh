timeout 500 python scripts/time_setup.py 2>&1 | tail -5
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 500 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r01e.json 2> gpurun_out/bench_r01e.err; python scripts/bench_summary.py gpurun_out/bench_r01e.json 2>/dev/null | head -30
