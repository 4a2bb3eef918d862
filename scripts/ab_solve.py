"""A/B of environment-selected kernel variants on one problem in one process (the variables are read by
``ba_set_problem``): whole-solve device times of the default solve, and per-kernel CUDA-event times of a short one.

    python scripts/ab_solve.py [shape] VAR=VALUE[,VAR=VALUE...] [VAR=VALUE ...]     ("-" = no variable set)
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
variants = sys.argv[2:] or ["-"]
fp = S.make_named_problem(shape, keep_truth=False)
s = FlatSolver(0)
touched = set()
for var in variants:
    for k in touched:
        os.environ.pop(k, None)
    if var != "-":
        for kv in var.split(","):
            k, v = kv.split("=")
            os.environ[k] = v
            touched.add(k)
    s.set_flat(fp)
    ms = []
    for rep in range(4):
        s.restore_initial()
        st = s.solve()
        ms.append(st["solve_ms"])
    print(f"{var:40s} solve ms {' '.join(f'{m:8.2f}' for m in ms)}  iters {st['iterations']} pcg {st['pcg_iterations']} "
          f"cost {st['final_cost']:.6f}", flush=True)
    s.restore_initial()
    st = s.solve(max_iters=3, verbose=2, pcg_tol=1e-30, pcg_max_iters=8, pcg_check_every=100)
    ks = s.kernel_stats()
    print("    " + "  ".join(f"{k} {1e3 * ms_ / n:.1f}us" for k, (n, ms_) in ks.items() if n and ms_ / n > 0.02), flush=True)
