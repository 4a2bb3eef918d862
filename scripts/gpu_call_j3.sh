#!/bin/bash
# 2-GPU A/B of the peer exchange: pull (remote loads) against push (remote stores), strong scaling on the venice shape
set -x
O=gpurun_out/r01j
mkdir -p $O
run() { # name, env...
  name=$1; shift
  env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus 2 --steps 3 --warmup 3 --scaling strong --no-cpu-baseline > $O/bench_2gpu_strong_$name.json 2> $O/bench_2gpu_strong_$name.err
  python scripts/bench_summary.py $O/bench_2gpu_strong_$name.json | grep -E "ms/step|schur_matvec_tiles|pcg_tail"
}
run ll BA_B200_XCHG=ll
run push BA_B200_XCHG=push
run ll2 BA_B200_XCHG=ll
run push2 BA_B200_XCHG=push
