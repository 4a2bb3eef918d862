"""Device-stamped LM trajectory of one BASELINE shape under a few option sets (ba_solve_trace).

    python scripts/r02_trace.py shard50m "precond=schur_jacobi" "-" "point_linesearch=0"
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

shape = sys.argv[1]
variants = sys.argv[2:] or ["-"]
fp = S.make_named_problem(shape, keep_truth=False)
s = FlatSolver(0)
s.set_flat(fp)
keys = ("status_text", "precond_text", "iterations", "pcg_iterations", "coarse_builds", "coarse_matvecs", "final_cost", "solve_ms")
for var in variants:
    kw = {}
    if var != "-":
        for kv in var.split(","):
            k, v = kv.split("=")
            kw[k] = v if not v.replace(".", "").replace("-", "").replace("e", "").isdigit() else (float(v) if ("." in v or "e" in v) else int(v))
    for rep in range(2):
        s.restore_initial()
        st = s.solve(**kw)
    tr = s.solve_trace()
    print(f"{var:40s}", json.dumps({k: st[k] for k in keys}), flush=True)
    print("   ms :", " ".join(f"{b - a:.1f}" for a, b in zip(tr["ms"][:-1], tr["ms"][1:])))
    print("   pcg:", " ".join(str(b - a) for a, b in zip(tr["pcg_total"][:-1], tr["pcg_total"][1:])))
    print("   lam:", " ".join(f"{v:.0e}" for v in tr["lambda"][1:]))
    print("   cost:", " ".join(f"{c:.5e}" for c in tr["cost"][::5]), flush=True)
