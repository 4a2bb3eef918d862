set -x
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r01f.json 2> gpurun_out/bench_r01f.err; echo rc=$?
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r01f_ref.json 2> gpurun_out/bench_r01f_ref.err; echo rc=$?
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r01f_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1; echo rc=$?
timeout 300 python scripts/profile_target.py venice 4 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tile_pass -s 2 -c 1 -f -o gpurun_out/prof_tile9 python scripts/profile_target.py venice 4 > gpurun_out/ncu.log 2>&1; echo rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_pcg_tail -s 2 -c 1 -f -o gpurun_out/prof_tail9 python scripts/profile_target.py venice 4 > gpurun_out/ncu2.log 2>&1; echo rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_lin_points -s 0 -c 1 -f -o gpurun_out/prof_linp9 python scripts/profile_target.py venice 4 > gpurun_out/ncu3.log 2>&1; echo rc=$?
python scripts/bench_summary.py gpurun_out/bench_r01f.json 2>/dev/null | head -24
cat gpurun_out/bench_r01f_ref.json
