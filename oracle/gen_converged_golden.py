"""Converged-cost references for the BASELINE.json shapes (tier "T1-tight" of SURVEY.md section 8c).  TEST INFRASTRUCTURE.

Runs ``oracle/lm_sparse.py`` (LM with exact sparse solves of the reduced camera system, IRLS Huber weights - the same
objective as the reference's ``least_squares(loss='huber')`` call, ``/root/reference/src/bundle_adjuster.py:170-174``) from the
generator's initial guess on the WELL-POSED variant of a shape (``synthetic.make_named_problem(..., well_posed=True)``: every
point keeps two inlier observations, so a minimiser exists) until the cost decrease per iteration is below ``stop`` relative,
and stores the final cost, the cost trace, the cameras and a sample of the points under ``tests/golden/``.

    python -m oracle.gen_converged_golden bal50 [max_iters] [stop]
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bundle_adjustment_b200 import synthetic as S  # noqa: E402
from oracle import ba_oracle as O, lm_sparse as LS, lm_schur as L  # noqa: E402


def main():
    shape = sys.argv[1]
    max_iters = int(sys.argv[2]) if len(sys.argv) > 2 else 150
    stop = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-9
    fp = S.make_named_problem(shape, keep_truth=False, well_posed=True)
    prob = O.Problem.from_flat(fp)
    t0 = time.perf_counter()
    res = LS.solve(prob, fp.rvec, fp.tvec, fp.pts, lam_min=1e-10, max_iters=max_iters, gtol=1e-8, ftol=stop, verbose=True)
    lin = L.linearize(prob, res.rvec, res.tvec, res.pts)
    sample = np.arange(fp.n_pts)
    # SciPy-side cross-check (small shapes): the reference's own solver, given the analytic Jacobian and tight tolerances and
    # started AT the converged point, must not find a lower cost than it by more than the 1e-6 the parity tests allow
    scipy_cost = scipy_nfev = scipy_status = -1
    if fp.n_obs <= 200_000:
        sres = LS.solve_scipy_tight(prob, res.x(), max_nfev=60)
        scipy_cost, scipy_nfev, scipy_status = float(sres.cost), int(sres.nfev), int(sres.status)
        print(f"scipy from the converged point: cost {scipy_cost:.9e} (ours {res.cost:.9e}, rel diff "
              f"{(res.cost - scipy_cost) / res.cost:.2e}), nfev {scipy_nfev}, status {scipy_status}", flush=True)
    out = os.path.join(ROOT, "tests", "golden", f"converged_{shape}.npz")
    np.savez_compressed(
        out, shape=shape, n_cams=fp.n_cams, n_pts=fp.n_pts, n_obs=fp.n_obs, cost=res.cost, raw=res.raw, status=res.status,
        iterations=res.iterations, gmax=LS.gradient_max(lin), rvec=res.rvec, tvec=res.tvec, pts_sample_idx=sample,
        pts_sample=res.pts[sample], trace_cost=np.array([h["cost_new"] if h["accepted"] else h["cost"] for h in res.history]),
        trace_gmax=np.array([h["gmax"] for h in res.history]), x0_cost=res.history[0]["cost"],
        last_decrease=res.history[-1]["cost"] - res.history[-1]["cost_new"],
        scipy_cost=scipy_cost, scipy_nfev=scipy_nfev, scipy_status=scipy_status,
        seconds=time.perf_counter() - t0, generator="oracle/gen_converged_golden.py")
    print(json.dumps({"shape": shape, "cost": res.cost, "status": res.status, "iterations": res.iterations,
                      "gmax": LS.gradient_max(lin), "seconds": time.perf_counter() - t0, "file": out}))


if __name__ == "__main__":
    main()
