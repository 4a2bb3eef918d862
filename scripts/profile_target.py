"""Short, fixed workload for ncu: one LM iteration with a few PCG iterations on a BASELINE shape.

    python scripts/profile_target.py [shape] [pcg_iters]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
pcg = int(sys.argv[2]) if len(sys.argv) > 2 else 3
fp = S.make_named_problem(shape, keep_truth=False)
s = FlatSolver(0)
s.set_flat(fp)
st = s.solve(max_iters=1, pcg_max_iters=pcg, pcg_check_every=100, pcg_tol=1e-30)
print(shape, fp.n_cams, fp.n_pts, fp.n_obs, st["iterations"], st["pcg_iterations"], st["launches"], st["solve_ms"])
