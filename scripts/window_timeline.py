"""Where a window-sized solve (SURVEY config 1) spends its time: host wall clock of every stage of FlatSolver use, and the
device-stamped time of each LM iteration (ba_solve_trace: entry 0 = initial point, stamped when the first cost is known).

    python scripts/window_timeline.py
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

fp = S.make_flat_problem(5, 2800, 6000, seed=S.SEEDS["window"])  # 1 fixed + 4 adjustable cameras, as the pipeline's windows
print(f"window: {fp.n_cams} adjustable cameras, {fp.n_pts} points, {fp.n_obs} observations")
s = FlatSolver(0)
for kw in ({}, {"max_iters": 10}, {"max_iters": 1}, {"device_loop": 0}):
    ts = []
    for rep in range(6):
        t0 = time.perf_counter()
        s.set_flat(fp)
        t1 = time.perf_counter()
        st = s.solve(**kw)
        t2 = time.perf_counter()
        ts.append((t1 - t0, t2 - t1, st["solve_ms"]))
    tr = s.solve_trace()
    a = np.median(np.array(ts[2:]), axis=0)
    print(f"{str(kw):24s} set_problem {1e3 * a[0]:.2f} ms | solve() wall {1e3 * a[1]:.2f} ms, events {a[2]:.2f} ms | {st['iterations']} LM iterations, "
          f"first stamp {tr['ms'][0]:.3f} ms, last {tr['ms'][-1]:.3f} ms, per iteration {np.median(np.diff(tr['ms'])) * 1e3:.1f} us")
