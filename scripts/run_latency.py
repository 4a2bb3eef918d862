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
    print(f"run {rep} (a NEW map every call: everything is uploaded): {1e3 * dt:7.2f} ms total | device solve {st['solve_ms']:6.2f} ms, "
          f"{st['iterations']} LM / {st['pcg_iterations']} PCG iterations, launches {st['launches']} | {out.getvalue().strip().splitlines()[-1].strip()}")


# the pipeline's pattern: ONE map that gains a keyframe (and the points it brings) between calls
def growing(resident, n_total=14):
    import cProfile, pstats
    sp = S.make_window_spec(n_kfs=n_total, n_pts=n_pts * n_total // 10)
    full = S.build_map(sp, Map, Keyframe, MapPoint, KeyPoint.from_xy)
    gm, a, out_ms = Map(), BundleAdjuster(sp["camera_matrix"], window_size=5, resident=resident), []
    prof = cProfile.Profile()
    for step, kf_id in enumerate(sorted(full.keyframes)):
        kf = full.keyframes[kf_id]
        for mp_id, _ in kf.observations:
            if mp_id not in gm.map_points:
                gm.add_map_point(full.map_points[mp_id])
        gm.add_keyframe(kf)
        if step < 5:
            continue
        t0 = time.perf_counter()
        if step == n_total - 1:
            prof.enable()
        with redirect_stdout(io.StringIO()):
            a.run(gm)
        prof.disable()
        out_ms.append((1e3 * (time.perf_counter() - t0), a.last_stats["solve_ms"], a.last_stats["iterations"], a.last_upload))
    return out_ms, prof


for resident in (True, False):
    ms, prof = growing(resident)
    print(f"growing map, resident={resident}: per call " + "  ".join(f"{m:.2f}" for m, *_ in ms) + " ms")
    tail = ms[2:-1]
    print(f"  median of calls 3.. : {np.median([m for m, *_ in tail]):.2f} ms total, device solve {np.median([m[1] for m in tail]):.2f} ms "
          f"({tail[-1][2]} LM iterations), last upload {tail[-1][3]}")
    if resident:
        import pstats
        st_ = pstats.Stats(prof, stream=sys.stdout).sort_stats("cumulative")
        st_.print_stats(14)
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
