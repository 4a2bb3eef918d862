"""Mat-vec kernel time against problem size and shape (single GPU).

    python scripts/size_sweep.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

CASES = [(1778, 993_000, 5_000_000), (10_000, 1_000_000, 5_000_000), (1778, 2_500_000, 12_500_000)]
s = FlatSolver(0)
for n_cams, n_pts, n_obs in CASES:
    fp = S.make_flat_problem(n_cams, n_pts, n_obs, 77, keep_truth=False)
    s.set_flat(fp)
    li = s.layout_info()
    for rep in range(2):
        s.restore_initial()
        st = s.solve(max_iters=3, verbose=2, pcg_tol=1e-30, pcg_max_iters=8, pcg_check_every=100)
    ks = s.kernel_stats()
    n, ms = ks["schur_matvec_tiles"]
    us = 1e3 * ms / n
    print(li)
    print(f"cams {n_cams:6d} pts {n_pts:8d} obs {n_obs:9d} tiles {li['point_tiles']:6d} segs {li['segments']:8d} "
          f"matvec {us:7.1f} us  {us / (n_obs / 1e6):6.2f} us/Mobs  rows_finish {1e3 * ks['rows_finish'][1] / ks['rows_finish'][0]:6.1f} us", flush=True)
