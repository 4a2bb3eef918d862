"""Generate ``tests/golden/*.npz`` by running the UNMODIFIED reference in the build container.

    python -m oracle.gen_golden            # from the repo root; needs /root/reference, cv2, scipy

TEST INFRASTRUCTURE.  For every fixture a synthetic sliding-window map is built through the
reference's own ``Map``/``Keyframe``/``MapPoint`` classes with real ``cv2.KeyPoint`` objects, and the
reference's ``BundleAdjuster`` methods are executed verbatim:

* ``_gather_local_data``  -> sorted map-point ids, observation rows, uv per row        (bit-exact golden)
* parameter packing (``run`` lines 157-162, ``cv2.Rodrigues``)                          (golden x0)
* ``_cost_function(x0)``                                                               (golden residuals)
* ``_prepare_sparsity_matrix(...).tocsr()``                                            (bit-exact golden)
* ``run(gmap)`` end to end, capturing the ``OptimizeResult`` of its ``least_squares`` call, the
  printed ``LBA Complete`` line and the mutated map                                    (golden solve)

The map description (``spec_*`` arrays) is stored next to the outputs so the GPU tests can rebuild the
identical map from the mirror classes without the reference being present.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from bundle_adjustment_b200 import synthetic as S          # noqa: E402
from oracle import ba_oracle as O                          # noqa: E402
from oracle.reference_loader import load_reference         # noqa: E402

FIXTURES = {
    # name: (make_window_spec kwargs, window_size used for run)
    "window_quirks": (dict(n_kfs=7, n_pts=120, seed=5, duplicates=2, dangling=2, kf_id_stride=2), 5),
    "window_short": (dict(n_kfs=5, n_pts=80, seed=6, mean_track=3.0), 5),      # exactly w keyframes: 1 fixed + 3 adjustable
    "window_cfg1": (dict(n_kfs=10, n_pts=1500, seed=S.SEEDS["window"], mean_track=4.0), 5),
    "global_cfg1": (dict(n_kfs=10, n_pts=600, seed=S.SEEDS["window"] + 1, mean_track=4.0), 10),  # main.py:81-87 global BA
}


def run_fixture(name, spec_kw, window_size, ba, ms, cv2):
    spec = S.make_window_spec(**spec_kw)
    gmap = S.build_map(spec, ms.Map, ms.Keyframe, ms.MapPoint, lambda x, y: cv2.KeyPoint(x, y, 1))
    adj = ba.BundleAdjuster(spec["camera_matrix"], window_size=window_size)

    all_ids = sorted(gmap.keyframes.keys())
    local = all_ids[-(window_size + 1):-1]
    fixed_id, adj_ids = local[0], local[1:]
    mp_ids, observations, kp2d = adj._gather_local_data(gmap, local)
    rvecs = np.array([cv2.Rodrigues(gmap.keyframes[i].R)[0].ravel() for i in adj_ids])
    tvecs = np.array([gmap.keyframes[i].t.ravel() for i in adj_ids])
    pts = np.array([gmap.map_points[i].position.ravel() for i in mp_ids])
    x0 = np.concatenate([rvecs.flatten(), tvecs.flatten(), pts.flatten()])
    fk = gmap.keyframes[fixed_id]
    r0 = adj._cost_function(x0, (fk.R, fk.t), fixed_id, adj_ids, mp_ids, observations, kp2d)
    A = adj._prepare_sparsity_matrix(len(adj_ids), len(mp_ids), adj_ids, mp_ids, observations).tocsr()

    captured = {}
    real_ls = ba.least_squares

    def capturing_least_squares(*a, **k):
        res = real_ls(*a, **k)
        captured["res"] = res
        return res

    ba.least_squares = capturing_least_squares
    out = io.StringIO()
    t0 = time.time()
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)
        try:
            with contextlib.redirect_stdout(out):
                adj.run(gmap)
        finally:
            os.chdir(cwd)
            ba.least_squares = real_ls
    wall = time.time() - t0
    res = captured["res"]
    lines = [ln for ln in out.getvalue().splitlines() if "LBA" in ln]

    # T1-tight: the same objective run to convergence from the same x0 (vectorised residual)
    flat = O.flatten_window(S.build_map(spec, ms.Map, ms.Keyframe, ms.MapPoint, lambda x, y: cv2.KeyPoint(x, y, 1)),
                            window_size)
    K = spec["camera_matrix"]
    prob = O.Problem(K[0, 0], K[1, 1], K[0, 2], K[1, 2], len(adj_ids), len(mp_ids), flat["fixed_rvec"],
                     flat["fixed_tvec"], flat["obs_cam"], flat["obs_pt"], flat["obs_uv"])
    tight = O.solve_reference(prob, x0, tight=True)

    gold = {f"spec_{k}": v for k, v in spec.items()}
    gold.update(
        window_size=np.int64(window_size),
        local_kf_ids=np.array(local, dtype=np.int64), fixed_kf_id=np.int64(fixed_id),
        adj_kf_ids=np.array(adj_ids, dtype=np.int64), mp_ids=np.array(mp_ids, dtype=np.int64),
        obs_kf=np.array([o[0] for o in observations], dtype=np.int64),
        obs_mp=np.array([o[1] for o in observations], dtype=np.int64),
        obs_uv=np.array([kp2d[o] for o in observations], dtype=np.float64),
        x0=x0, r0=np.asarray(r0, dtype=np.float64),
        sp_indptr=A.indptr.astype(np.int64), sp_indices=A.indices.astype(np.int64), sp_data=A.data.astype(np.int64),
        sp_shape=np.array(A.shape, dtype=np.int64),
        res_x=res.x, res_fun=res.fun, res_cost=np.float64(res.cost), res_status=np.int64(res.status),
        res_nfev=np.int64(res.nfev), res_njev=np.int64(res.njev), res_optimality=np.float64(res.optimality),
        initial_cost=np.float64(np.sum(r0 ** 2)), final_cost=np.float64(np.sum(res.fun ** 2)),
        log_lines=np.array(lines),
        out_kf_R=np.array([gmap.keyframes[i].R for i in all_ids]),
        out_kf_t=np.array([gmap.keyframes[i].t for i in all_ids]),
        out_mp_pos=np.array([gmap.map_points[i].position for i in sorted(gmap.map_points)]),
        tight_x=tight.x, tight_cost=np.float64(tight.cost), tight_raw=np.float64(np.sum(tight.fun ** 2)),
        tight_status=np.int64(tight.status), tight_nfev=np.int64(tight.nfev),
        tight_optimality=np.float64(tight.optimality),
        ref_wall_s=np.float64(wall),
    )
    path = os.path.join(ROOT, "tests", "golden", f"{name}.npz")
    np.savez_compressed(path, **gold)
    print(f"{name}: Nc={len(adj_ids)} Np={len(mp_ids)} No={len(observations)} ref run {wall:.1f}s "
          f"status={res.status} nfev={res.nfev} cost {0.5 * 0 + res.cost:.4f} | tight status={tight.status} "
          f"nfev={tight.nfev} cost {tight.cost:.6f} opt {tight.optimality:.2e} | {lines[-1].strip() if lines else ''}"
          f" -> {os.path.getsize(path) / 1024:.0f} KiB")


def main():
    import cv2
    import scipy
    ba, ms = load_reference()
    print(f"numpy {np.__version__} scipy {scipy.__version__} cv2 {cv2.__version__}")
    only = sys.argv[1:]
    for name, (kw, w) in FIXTURES.items():
        if only and name not in only:
            continue
        run_fixture(name, kw, w, ba, ms, cv2)


if __name__ == "__main__":
    main()
