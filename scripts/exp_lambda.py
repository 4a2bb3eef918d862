import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bundle_adjustment_b200 import FlatSolver, synthetic as S
from oracle import ba_oracle as O
s = FlatSolver(0)
def drift(fp, pts):
    c = -O.rodrigues(fp.fixed_rvec).T @ fp.fixed_tvec
    return float(np.median(np.linalg.norm(pts - c, axis=1) / np.linalg.norm(fp.pts - c, axis=1)))
for name, fp in [("small", S.make_flat_problem(5, 80, 240, seed=6)), ("bal50", S.make_named_problem("bal50")), ("ladybug", S.make_named_problem("ladybug"))]:
    s.set_flat(fp)
    for lam_min in [1e-12, 1e-8, 1e-6, 1e-5, 1e-4]:
        for pcg_tol in [1e-1, 1e-2]:
            for iters in [50, 300]:
                s.restore_initial()
                st = s.solve(max_iters=iters, lambda_min=lam_min, pcg_tol=pcg_tol, ftol=1e-6 if iters == 50 else 1e-12)
                _, _, pts = s.get_params()
                print(f"{name:8s} lam_min {lam_min:7.0e} pcg_tol {pcg_tol:5.0e} max_it {iters:4d}: it {st['iterations']:4d} rej {st['rejected']:3d} pcg {st['pcg_iterations']:6d} "
                      f"cost {st['final_cost']:.8e} status {st['status']} ms {st['solve_ms']:9.2f} drift {drift(fp, pts):.5f}", flush=True)
