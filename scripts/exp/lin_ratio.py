"""Ratios of test_full_size_properties step (4) on venice: device linearisation against the oracle's, per block set."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S
from oracle import ba_oracle as O, lm_schur as L
fp = S.make_named_problem("venice", keep_truth=False)
prob = O.Problem.from_flat(fp)
for env in sys.argv[1:] or ["-"]:
    os.environ.pop("BA_B200_NO_SLOT_BALANCE", None)
    if env != "-":
        k, v = env.split("="); os.environ[k] = v
    s = FlatSolver(0)
    s.set_flat(fp)
    st = s.solve(max_iters=12)
    rvec, tvec, pts = s.get_params()
    nb = s.normal_blocks()
    lin = L.linearize(prob, rvec, tvec, pts)
    out = []
    for key, ref in (("gc", lin.gc), ("gp", lin.gp), ("U", lin.U), ("V", lin.V)):
        out.append(f"{key} {np.linalg.norm(nb[key] - ref) / np.linalg.norm(ref):.3e}")
        if key in ("gp", "V"):
            d = np.abs(nb[key] - ref).reshape(fp.n_pts, -1).max(axis=1)
            w = int(np.argmax(d)); out.append(f"(worst point {w}: {d[w]:.3e}, ref {np.abs(ref[w]).max():.3e}, track {int((fp.obs_pt == w).sum())})")
    print(env, st["final_cost"], " ".join(out), flush=True)
