"""Tier T0 of SURVEY.md section 8d: ONE verbatim ``_cost_function`` call of the UNMODIFIED reference
(``/root/reference/src/bundle_adjuster.py:24-72``: a Python loop with one ``cv2.projectPoints`` call per observation) on a
slice of the venice-shaped workload, and ONE verbatim ``BundleAdjuster.run`` on the config-1 window.  TEST INFRASTRUCTURE.

/root/reference exists only in the build container, so the figures are committed (``tests/golden/t0_baseline.json``) with their
provenance; ``bench.py`` quotes them next to the timed tier-T1 baseline and extrapolates the per-call time linearly in the
number of observations (the loop is per observation).  A full T0 solve of the 5 M-observation problem is infeasible
(1 + nfev + 9 njev calls of ~270 s each).

    python -m oracle.gen_t0_baseline [slice_observations]
"""
import io
import json
import os
import platform
import sys
import tempfile
import time
from contextlib import redirect_stdout

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cv2  # noqa: E402
from bundle_adjustment_b200 import synthetic as S  # noqa: E402
from oracle import ba_oracle as O  # noqa: E402
from oracle.reference_loader import load_reference  # noqa: E402


def main():
    n_slice = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
    ba, ms = load_reference()
    # a venice-shaped problem scaled to the slice size (same generator, same seed)
    fp = S.make_named_problem("venice", scale=n_slice / S.SHAPES["venice"][2], keep_truth=False)
    adj_ids = list(range(1, fp.n_cams + 1))
    mp_ids = list(range(fp.n_pts))
    observations = [(int(c) + 1, int(p)) for c, p in zip(fp.obs_cam, fp.obs_pt)]   # keyframe 0 = the fixed one
    kp2d = {(int(c) + 1, int(p)): (float(u), float(v)) for c, p, (u, v) in zip(fp.obs_cam, fp.obs_pt, fp.obs_uv)}
    fixed_pose = (cv2.Rodrigues(fp.fixed_rvec)[0], fp.fixed_tvec.reshape(3, 1))
    adj = ba.BundleAdjuster(S.CAMERA_MATRIX.copy(), window_size=5)
    x0 = fp.x0()
    t0 = time.perf_counter()
    r = adj._cost_function(x0, fixed_pose, 0, adj_ids, mp_ids, observations, kp2d)
    dt = time.perf_counter() - t0
    r_oracle = O.Problem.from_flat(fp).residuals(x0)
    err = float(np.linalg.norm(np.asarray(r) - r_oracle) / np.linalg.norm(r_oracle))
    # config 1: the pipeline's window through the reference's run()
    spec = S.make_window_spec(n_kfs=10, n_pts=3000)
    gmap = S.build_map(spec, ms.Map, ms.Keyframe, ms.MapPoint, lambda x, y: cv2.KeyPoint(float(x), float(y), 1.0))
    runner = ba.BundleAdjuster(spec["camera_matrix"], window_size=5)
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as d:
        os.chdir(d)
        os.makedirs("debug_sparsity", exist_ok=True); os.makedirs("output_map/lba_steps", exist_ok=True)
        t1 = time.perf_counter()
        with redirect_stdout(io.StringIO()) as out:
            runner.run(gmap)
        dt_run = time.perf_counter() - t1
        os.chdir(cwd)
    rec = {
        "what": "one verbatim reference _cost_function call (src/bundle_adjuster.py:24-72) on a venice-shaped slice",
        "slice_observations": fp.n_obs, "slice_cameras": fp.n_cams + 1, "slice_points": fp.n_pts,
        "seconds_per_call": dt, "us_per_observation_per_call": 1e6 * dt / fp.n_obs,
        "relative_difference_to_oracle_residuals": err,
        "calls_per_outer_iteration": 10, "calls_note": "1 residual evaluation + 9 finite-difference column groups (SURVEY 6.2)",
        "window_run": {"what": "verbatim BundleAdjuster.run, 10 keyframes / 3000 points, window_size 5",
                       "seconds": dt_run, "last_line": out.getvalue().strip().splitlines()[-1]},
        "host": {"cpu_count": os.cpu_count(), "machine": platform.machine(), "processor": platform.processor(),
                 "python": platform.python_version(), "cv2": cv2.__version__, "numpy": np.__version__},
        "threads": 1, "where": "build container (the reference sources do not travel to the GPU box)",
        "generator": "oracle/gen_t0_baseline.py",
    }
    path = os.path.join(ROOT, "tests", "golden", "t0_baseline.json")
    with open(path, "w") as f:
        json.dump(rec, f, indent=1)
    print(json.dumps(rec, indent=1))


if __name__ == "__main__":
    main()
