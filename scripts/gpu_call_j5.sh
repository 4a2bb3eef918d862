#!/bin/bash
# 8-GPU records with the LL exchange: weak scaling on the venice shape (what the driver's scaling run does) and the
# 50 M-observation problem sharded 8 ways
set -x
O=gpurun_out/r01j
mkdir -p $O
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 \
    bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_8gpu_weak.json 2> $O/bench_8gpu_weak.err
python scripts/bench_summary.py $O/bench_8gpu_weak.json | grep -E "ms/step|schur_matvec_tiles|pcg_tail|lin_|rhs"
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 \
    bench.py --gpus 8 --steps 2 --warmup 3 --workload shard50m --scaling strong --no-cpu-baseline > $O/bench_8gpu_50m.json 2> $O/bench_8gpu_50m.err
python scripts/bench_summary.py $O/bench_8gpu_50m.json | grep -E "ms/step|schur_matvec_tiles|pcg_tail|lin_|rhs"
