"""Round-2 first GPU call: Schur-Jacobi against the two-level preconditioner on one problem (whole-solve device time,
iteration counts, final cost), with the LM trace on stderr (verbose bit 0).

    python scripts/r02_first.py [shape] [max_iters]
"""
import os
import sys
import json

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
max_iters = int(sys.argv[2]) if len(sys.argv) > 2 else 50
fp = S.make_named_problem(shape, keep_truth=False)
s = FlatSolver(0)
s.set_flat(fp)
out = {}
for name, kw in (("schur_jacobi", {}), ("two_level", {"precond": "two_level"})):
    for rep in range(2):
        s.restore_initial()
        st = s.solve(max_iters=max_iters, device_loop=0, verbose=1 if rep == 1 else 0, **kw)
    out[name] = st
    print(name, json.dumps({k: st[k] for k in ("status_text", "iterations", "pcg_iterations", "final_cost", "solve_ms", "launches")}), flush=True)
