"""Warm stage times of ba_set_problem (BA_B200_TRACE) and of the end-to-end step (upload + build + solve + read-back) on a
BASELINE shape, host buffers pinned as in bench.py's e2e leg.

    python scripts/set_problem_stages.py [shape]
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
fp = S.make_named_problem(shape, keep_truth=False)
keep, host = [], {}
for k in ("rvec", "tvec", "pts", "obs_cam", "obs_pt", "obs_uv"):
    t = torch.from_numpy(np.ascontiguousarray(getattr(fp, k))).pin_memory()
    keep.append(t)
    host[k] = t.numpy()
s = FlatSolver(0)


def upload():
    s.set_problem(fp.fx, fp.fy, fp.cx, fp.cy, host["rvec"], host["tvec"], fp.fixed_rvec, fp.fixed_tvec, host["pts"],
                  host["obs_cam"], host["obs_pt"], host["obs_uv"])


for rep in range(3):
    upload()
    s.solve()
os.environ["BA_B200_TRACE"] = "1"
sys.stderr.write("---- warm call, stage by stage (each mark synchronises) ----\n")
upload()
os.environ.pop("BA_B200_TRACE")
for rep in range(3):
    t0 = time.perf_counter()
    upload()
    t1 = time.perf_counter()
    st = s.solve()
    t2 = time.perf_counter()
    s.get_params()
    t3 = time.perf_counter()
    print(f"set_problem {1e3 * (t1 - t0):7.2f} ms  solve {1e3 * (t2 - t1):7.2f} ms (device {st['solve_ms']:.2f})  "
          f"read-back {1e3 * (t3 - t2):6.2f} ms  total {1e3 * (t3 - t0):7.2f} ms", flush=True)
