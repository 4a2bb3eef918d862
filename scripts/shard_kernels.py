"""Per-kernel times of one rank's share of a sharded problem, on ONE GPU (what a rank of an N-GPU run computes locally).

    python scripts/shard_kernels.py [shape] [world]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, sharding, synthetic as S  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "shard50m"
world = int(sys.argv[2]) if len(sys.argv) > 2 else 8
fp = S.make_named_problem(shape, keep_truth=False)
shard, (lo, hi) = sharding.shard_problem(fp, 0, world)
s = FlatSolver(0)
s.set_flat(shard)
print(s.layout_info())
for rep in range(2):
    s.restore_initial()
    st = s.solve(max_iters=3, verbose=2, pcg_tol=1e-30, pcg_max_iters=8, pcg_check_every=100)
for k, (n, ms) in s.kernel_stats().items():
    if n:
        print(f"{k:24s} {n:6d} {ms:9.3f} ms {1e3 * ms / n:8.1f} us")
