"""Host-side timing of ba_set_problem / ba_solve / ba_get_params on a BASELINE shape (pinned host buffers)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bundle_adjustment_b200 import FlatSolver, synthetic as S
shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
fp = S.make_named_problem(shape, keep_truth=False)
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
for k in ("rvec", "tvec", "pts", "obs_cam", "obs_pt", "obs_uv"):
    setattr(fp, k, pin(getattr(fp, k)))
s = FlatSolver(0)
for rep in range(3):
    t0 = time.perf_counter(); s.set_flat(fp); t1 = time.perf_counter()
    print(f"set_problem {1e3*(t1-t0):.2f} ms", flush=True)
t0 = time.perf_counter(); s.get_params(); print(f"get_params {1e3*(time.perf_counter()-t0):.2f} ms")
