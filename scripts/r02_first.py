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
keys = ("status_text", "precond_text", "iterations", "pcg_iterations", "coarse_builds", "coarse_matvecs", "final_cost", "solve_ms", "launches")
for name, kw in (("schur_jacobi ftol=0", {"precond": "schur_jacobi", "ftol": 0.0, "xtol": 0.0}), ("default ftol=0", {"ftol": 0.0, "xtol": 0.0}),
                 ("default", {})):
    for rep in range(3):
        s.restore_initial()
        st = s.solve(max_iters=max_iters, verbose=1 if rep == 2 and name == "default" else 0, **kw)
    out[name] = st
    tr = s.solve_trace()
    print("   ms per LM iteration:", " ".join(f"{b - a:.2f}" for a, b in zip(tr["ms"][:-1], tr["ms"][1:])), flush=True)
    print("   pcg per LM iteration:", " ".join(str(b - a) for a, b in zip(tr["pcg_total"][:-1], tr["pcg_total"][1:])), flush=True)
    print(name, json.dumps({k: st[k] for k in keys}), flush=True)
s.restore_initial()
st = s.solve(max_iters=max_iters, verbose=2, pcg_check_every=1)
ks = s.kernel_stats()
print("per-kernel us:", "  ".join(f"{k} {1e3 * ms_ / n:.1f}x{n}" for k, (n, ms_) in ks.items() if n and ms_ > 0.0), flush=True)
