"""Short workload of pipeline-window size for ncu launch lists: 5 keyframes / 2800 points / 6000 observations."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

n_cams, n_pts, n_obs = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (5, 2800, 6000)
fp = S.make_flat_problem(n_cams, n_pts, n_obs, seed=1001)
s = FlatSolver(0)
s.set_flat(fp)
st = s.solve(max_iters=3, device_loop=0)
print(st["iterations"], st["pcg_iterations"], st["launches"], st["solve_ms"])
st = s.solve(verbose=2)
for k, (n, ms) in s.kernel_stats().items():
    if n:
        print(f"{k:24s} {n:6d} {ms:9.3f} ms {1e3 * ms / n:8.1f} us")
