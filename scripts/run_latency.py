"""Where BundleAdjuster.run spends its time on a pipeline-sized window (SURVEY config 1).

    python scripts/run_latency.py [n_kfs] [n_pts]
"""
import io
import os
import sys
import time
from contextlib import redirect_stdout

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import BundleAdjuster, synthetic as S  # noqa: E402
from bundle_adjustment_b200.map_structures import Keyframe, KeyPoint, Map, MapPoint  # noqa: E402

n_kfs = int(sys.argv[1]) if len(sys.argv) > 1 else 10
n_pts = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
spec = S.make_window_spec(n_kfs=n_kfs, n_pts=n_pts)
adj = BundleAdjuster(spec["camera_matrix"], window_size=5)
for rep in range(4):
    gmap = S.build_map(spec, Map, Keyframe, MapPoint, KeyPoint.from_xy)
    t0 = time.perf_counter()
    with redirect_stdout(io.StringIO()) as out:
        adj.run(gmap)
    dt = time.perf_counter() - t0
    st = adj.last_stats
    print(f"run {rep}: {1e3 * dt:7.2f} ms total | device solve {st['solve_ms']:6.2f} ms, {st['iterations']} LM / {st['pcg_iterations']} PCG iterations, "
          f"launches {st['launches']} | {out.getvalue().strip().splitlines()[-1].strip()}")
# stage by stage
gmap = S.build_map(spec, Map, Keyframe, MapPoint, KeyPoint.from_xy)
ids = sorted(gmap.keyframes)[-6:-1]
t = [time.perf_counter()]
mp_ids, obs, kp = adj._gather_local_data(gmap, ids); t.append(time.perf_counter())
oc, op, ouv = adj._index_observations(ids[1:], mp_ids, obs, kp, ids[0]); t.append(time.perf_counter())
from bundle_adjustment_b200.geometry import matrix_to_rvec
rv = np.array([matrix_to_rvec(gmap.keyframes[i].R) for i in ids[1:]]); tv = np.array([gmap.keyframes[i].t.ravel() for i in ids[1:]])
pts = np.array([gmap.map_points[i].position.ravel() for i in mp_ids]); t.append(time.perf_counter())
s = adj._get_solver()
K = spec["camera_matrix"]
s.set_problem(K[0, 0], K[1, 1], K[0, 2], K[1, 2], rv, tv, matrix_to_rvec(gmap.keyframes[ids[0]].R), gmap.keyframes[ids[0]].t.ravel(), pts, oc, op, ouv)
t.append(time.perf_counter())
st = s.solve(**adj.solver_options); t.append(time.perf_counter())
x = s.get_x(); t.append(time.perf_counter())
adj._update_map(gmap, x, ids[1:], mp_ids); t.append(time.perf_counter())
names = ["gather", "index", "pack", "set_problem", "solve", "get_x", "update_map"]
print(f"window: {len(ids)} keyframes, {len(mp_ids)} points, {len(obs)} observations")
for n, a, b in zip(names, t[:-1], t[1:]):
    print(f"  {n:12s} {1e3 * (b - a):8.3f} ms")
# per-kernel device times of the same window (host-driven loop, CUDA events around every launch)
s.set_problem(K[0, 0], K[1, 1], K[0, 2], K[1, 2], rv, tv, matrix_to_rvec(gmap.keyframes[ids[0]].R), gmap.keyframes[ids[0]].t.ravel(), pts, oc, op, ouv)
st = s.solve(verbose=2, device_loop=0)
ks = s.kernel_stats()
print(f"per-kernel (host loop, {st['iterations']} LM iterations, {st['solve_ms']:.2f} ms): " +
      "  ".join(f"{k} {1e3 * ms_ / n:.1f}us x{n}" for k, (n, ms_) in ks.items() if n and ms_ > 0.0))
