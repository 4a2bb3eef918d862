"""Per-kernel device times (CUDA events around every launch, verbose = 2) of a few LM iterations on a BASELINE shape.

    python scripts/kernel_times.py venice [max_iters] [option=value ...]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 6
kw = {}
for kv in sys.argv[3:]:
    k, v = kv.split("=")
    kw[k] = float(v) if ("." in v or "e" in v) else int(v)
fp = S.make_named_problem(shape, keep_truth=False)
s = FlatSolver(0)
s.set_flat(fp)
for rep in range(2):
    s.restore_initial()
    st = s.solve(max_iters=iters, verbose=2, **kw)
ks = s.kernel_stats()
print(f"{shape}: {st['iterations']} LM / {st['pcg_iterations']} PCG iterations, {st['solve_ms']:.2f} ms (with per-launch events)")
for k, (n, ms) in ks.items():
    if n and ms > 0:
        print(f"  {k:24s} x{n:5d}  {1e3 * ms / n:8.1f} us")
