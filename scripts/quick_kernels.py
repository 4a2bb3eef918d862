"""Per-kernel CUDA-event times of a short solve (3 LM iterations x 8 PCG iterations, none skipped) on a BASELINE shape.

    python scripts/quick_kernels.py [shape] [schur_path]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
path = sys.argv[2] if len(sys.argv) > 2 else "tiles"
fp = S.make_named_problem(shape, keep_truth=False)
s = FlatSolver(0)
s.set_flat(fp)
print(s.layout_info())
for rep in range(2):
    s.restore_initial()
    st = s.solve(max_iters=3, verbose=2, schur_path=path, pcg_tol=1e-30, pcg_max_iters=8, pcg_check_every=100)
ks = s.kernel_stats()
print(shape, st["iterations"], st["pcg_iterations"], round(st["solve_ms"], 2), "ms", st["final_cost"])
for k, (n, ms) in ks.items():
    if n:
        print(f"{k:24s} {n:6d} {ms:9.3f} ms {1e3 * ms / n:8.1f} us")
