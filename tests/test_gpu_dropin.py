"""GPU (B200): the drop-in ``BundleAdjuster`` on the same maps the reference was run on (golden fixtures),
called the way the reference's pipeline calls it (``src/pipeline.py:99``, ``src/main.py:81-87``)."""
import io
import re
from contextlib import redirect_stdout

import numpy as np
import pytest

from conftest import golden_names, load_golden, mirror_map, oracle_problem
from bundle_adjustment_b200 import BundleAdjuster
from oracle import ba_oracle as O

pytestmark = pytest.mark.gpu

LOG_RE = re.compile(r"-> LBA Complete\. Initial Cost: ([\d.]+), Final Cost: ([\d.]+), Improvement: ([\d.]+)%")


def _snapshot(gmap):
    return ({k: (kf.R.copy(), kf.t.copy()) for k, kf in gmap.keyframes.items()},
            {k: mp.position.copy() for k, mp in gmap.map_points.items()})


@pytest.mark.parametrize("name", golden_names())
def test_run_on_reference_windows(name):
    g, spec = load_golden(name)
    gmap = mirror_map(spec)
    before_kf, before_mp = _snapshot(gmap)
    adj = BundleAdjuster(spec["camera_matrix"], window_size=int(g["window_size"]))
    out = io.StringIO()
    with redirect_stdout(out):
        assert adj.run(gmap) is None
    text = out.getvalue()
    assert "--- Running Local Bundle Adjustment ---" in text
    m = LOG_RE.search(text)
    assert m, text                                              # the line src/analyze_log.py:42-45 parses
    ini, fin = float(m.group(1)), float(m.group(2))
    assert abs(ini - float(g["initial_cost"])) <= 0.011         # printed with 2 decimals; same x0 -> same raw cost
    assert fin < ini
    st = adj.last_stats
    # same objective, better minimiser: the Huber cost reached is below what the reference's call reached
    assert st["final_cost"] <= float(g["res_cost"])
    # map mutated in place exactly where the reference mutates it
    adj_ids, fixed_id = set(g["adj_kf_ids"].tolist()), int(g["fixed_kf_id"])
    window_mp = set(g["mp_ids"].tolist())
    for k, kf in gmap.keyframes.items():
        assert kf.R.shape == (3, 3) and kf.t.shape == (3, 1)
        changed = not (np.array_equal(kf.R, before_kf[k][0]) and np.array_equal(kf.t, before_kf[k][1]))
        assert changed == (k in adj_ids), k                     # fixed, older and the newest keyframe untouched
        assert np.abs(kf.R @ kf.R.T - np.eye(3)).max() < 1e-12
    for k, mp in gmap.map_points.items():
        assert mp.position.shape == (3, 1)
        if k not in window_mp:
            assert np.array_equal(mp.position, before_mp[k])
    # the mutated map evaluates (through the oracle) to the cost the device reported
    flat = O.flatten_window(gmap, int(g["window_size"]))
    prob = oracle_problem(spec, flat)
    x = np.concatenate([flat["rvec"].ravel(), flat["tvec"].ravel(), flat["pts"].ravel()])
    r = prob.residuals(x)
    assert abs(O.raw_cost(r) - st["final_raw_cost"]) <= 1e-8 * st["final_raw_cost"]
    assert abs(O.huber_cost(r) - st["final_cost"]) <= 1e-8 * st["final_cost"]


def test_private_helpers_side_by_side_with_reference_outputs():
    """_cost_function / _prepare_sparsity_matrix keep the reference's signatures and outputs."""
    g, spec = load_golden("window_quirks")
    gmap = mirror_map(spec)
    adj = BundleAdjuster(spec["camera_matrix"], window_size=int(g["window_size"]))
    local = g["local_kf_ids"].tolist()
    fixed_id, adj_ids = local[0], local[1:]
    mp_ids, obs, kp2d = adj._gather_local_data(gmap, local)
    fk = gmap.keyframes[fixed_id]
    r = adj._cost_function(g["x0"], (fk.R, fk.t), fixed_id, adj_ids, mp_ids, obs, kp2d)
    assert r.shape == g["r0"].shape
    assert np.linalg.norm(r - g["r0"]) <= 1e-9 * np.linalg.norm(g["r0"])
    A = adj._prepare_sparsity_matrix(len(adj_ids), len(mp_ids), adj_ids, mp_ids, obs).tocsr()
    assert A.shape == tuple(g["sp_shape"])
    assert np.array_equal(A.indptr, g["sp_indptr"]) and np.array_equal(A.indices, g["sp_indices"])
    assert np.array_equal(A.data, g["sp_data"])


def test_global_ba_call_pattern_and_divergence_guard(tmp_path):
    """main.py:81-87 sets window_size = number of keyframes for the final global BA, then restores it."""
    g, spec = load_golden("global_cfg1")
    gmap = mirror_map(spec)
    adj = BundleAdjuster(spec["camera_matrix"], window_size=5, lba_steps_dir=str(tmp_path))
    original = adj.window_size
    adj.window_size = len(gmap.keyframes)
    out = io.StringIO()
    with redirect_stdout(out):
        adj.run(gmap)
    adj.window_size = original
    assert LOG_RE.search(out.getvalue())
    assert (tmp_path / f"map_after_lba_kf_{int(g['fixed_kf_id'])}.pcd").is_file()
    # divergence guard (src/bundle_adjuster.py:177-179): with zero iterations final == initial -> discard
    gmap2 = mirror_map(spec)
    before_kf, before_mp = _snapshot(gmap2)
    adj0 = BundleAdjuster(spec["camera_matrix"], window_size=len(gmap2.keyframes), max_iters=0)
    out = io.StringIO()
    with redirect_stdout(out):
        adj0.run(gmap2)
    assert "LBA Diverged!" in out.getvalue() and "Discarding results." in out.getvalue()
    for k, kf in gmap2.keyframes.items():
        assert np.array_equal(kf.R, before_kf[k][0]) and np.array_equal(kf.t, before_kf[k][1])


# ---------------------------------------------------------------------------------------------------
# the maps the UNMODIFIED reference pipeline handed to its bundle adjuster on real frames
# (tests/golden/pipeline_desk.npz, recorded by oracle/gen_pipeline_golden.py), replayed call by call
# ---------------------------------------------------------------------------------------------------
def _huber_of(gmap, window_size, K):
    flat = O.flatten_window(gmap, window_size)
    prob = O.Problem(K[0, 0], K[1, 1], K[0, 2], K[1, 2], len(flat["adj_kf_ids"]), len(flat["mp_ids"]),
                     flat["fixed_rvec"], flat["fixed_tvec"], flat["obs_cam"], flat["obs_pt"], flat["obs_uv"])
    x = np.concatenate([flat["rvec"].ravel(), flat["tvec"].ravel(), flat["pts"].ravel()])
    r = prob.residuals(x)
    return O.huber_cost(r), O.raw_cost(r)


def test_replay_of_the_reference_pipeline_calls():
    from conftest import GOLDEN_DIR, recorded_map
    import os
    g = np.load(os.path.join(GOLDEN_DIR, "pipeline_desk.npz"))
    K = g["camera_matrix"]
    adj = BundleAdjuster(K, window_size=int(g["ba_window_size"]))        # src/pipeline.py:39
    n_solved = 0
    for k in range(int(g["n_calls"])):
        ref_lines = [str(s) for s in g[f"c{k}_lines"]]
        gmap = recorded_map(g, k)
        adj.window_size = int(g[f"c{k}_window_size"])                    # src/main.py:83-86 mutates the attribute
        out = io.StringIO()
        with redirect_stdout(out):
            assert adj.run(gmap) is None
        text = out.getvalue()
        ref_skipped = [ln for ln in ref_lines if "Skipped" in ln]
        if ref_skipped:
            assert ref_skipped[0].strip() in text                        # same skip, same words
            assert adj.last_stats is None
            continue
        ref_m = LOG_RE.search("\n".join(ref_lines))
        m = LOG_RE.search(text)
        assert ref_m and m, (ref_lines, text)
        ref_ini, ini, fin = float(ref_m.group(1)), float(m.group(1)), float(m.group(2))
        assert abs(ini - ref_ini) <= 0.011 + 1e-9 * ref_ini              # same map, same raw cost at the start
        assert fin < ini
        # same objective (Huber, 1 px): the device solve ends at or below what the reference's call reached
        ws = int(g[f"c{k}_window_size"])
        ref_huber, _ = _huber_of(recorded_map(g, k, "after"), ws, K)
        mine_huber, mine_raw = _huber_of(gmap, ws, K)
        st = adj.last_stats
        assert abs(mine_huber - st["final_cost"]) <= 1e-8 * st["final_cost"]
        assert abs(mine_raw - fin) <= 0.011 + 1e-8 * mine_raw
        assert mine_huber <= ref_huber * (1.0 + 1e-6), (k, mine_huber, ref_huber)
        n_solved += 1
    assert n_solved >= 3


# ---- device-resident map (ba_map_*) -------------------------------------------------------------------
def _device_problem(solver):
    """What the device holds of the uploaded problem, through the C ABI: residual vector (row order + keypoints + poses +
    points), the Jacobian's CSR structure (camera / point of every row) and the parameter vector."""
    indptr, indices, shape = solver.sparsity_csr()
    return solver.residuals(), indptr, indices, shape, solver.get_x()


def test_resident_window_is_the_host_flattened_problem_and_uploads_only_the_delta():
    """``ba_map_window`` (device-built window from the resident rows) against the host flattening (``_gather_flat`` +
    ``ba_set_problem``) AND against the CPU statement (oracle/map_window.py), on a map that grows between calls as the
    pipeline grows it: residual vectors bit-identical, same Jacobian structure, same x0; a call uploads exactly what was
    appended since the previous one."""
    from bundle_adjustment_b200 import FlatSolver, synthetic as S
    from bundle_adjustment_b200.geometry import matrix_to_rvec
    from bundle_adjustment_b200.map_structures import KeyPoint, MapPoint
    from oracle.map_window import MapStore
    spec = S.make_window_spec(n_kfs=8, n_pts=600, seed=23, duplicates=4, dangling=3, kf_id_stride=3)
    gmap = mirror_map(spec)
    K = spec["camera_matrix"]
    adj = BundleAdjuster(K, window_size=5)
    local = sorted(gmap.keyframes)[-6:-1]

    def check():
        dev = adj._get_solver()
        before = dev.map_stats()
        ids = adj._window_resident(gmap, local)
        after = dev.map_stats()
        got = _device_problem(dev)
        ids_h, oc, op, ouv = adj._gather_flat(gmap, local, local[0])
        assert ids == ids_h
        rv = np.array([matrix_to_rvec(gmap.keyframes[i].R) for i in local])
        tv = np.array([gmap.keyframes[i].t.ravel() for i in local])
        pts = np.array([gmap.map_points[i].position.ravel() for i in ids_h])
        with FlatSolver(0) as host:
            host.set_problem(K[0, 0], K[1, 1], K[0, 2], K[1, 2], rv[1:], tv[1:], rv[0], tv[0], pts, oc, op, ouv)
            want = _device_problem(host)
        for a, b in zip(got, want):
            assert np.array_equal(a, b)
        # the CPU statement builds the same rows from the same keyframe lists
        cpu = MapStore()
        for k in sorted(gmap.keyframes):
            obs, kps = gmap.keyframes[k].observations, gmap.keyframes[k].keypoints
            cpu.map_append_observations(k, [o[0] for o in obs], [kps[o[1]].pt for o in obs])
        cpu.xyz = {i: mp.position.ravel().copy() for i, mp in gmap.map_points.items()}
        cpu.map_window(K[0, 0], K[1, 1], K[0, 2], K[1, 2], local, rv, tv)
        f = cpu.flat
        assert f["ids"].tolist() == ids and np.array_equal(f["obs_cam"], oc) and np.array_equal(f["obs_pt"], op)
        assert np.array_equal(f["obs_uv"], ouv) and np.array_equal(f["pts"], pts)
        return after["rows_appended"] - before["rows_appended"], after["points_set"] - before["points_set"]

    n_rows = sum(len(gmap.keyframes[k].observations) for k in local)
    assert check() == (n_rows, len(gmap.map_points))
    assert check() == (0, 0)
    kf = gmap.keyframes[local[2]]
    new_id = int(spec["mp_ids"].max()) + 100
    gmap.add_map_point(MapPoint(id=new_id, position=np.array([[0.1], [0.2], [9.0]])))
    kf.keypoints.append(KeyPoint.from_xy(321.5, 222.25))
    kf.observations.append((new_id, len(kf.keypoints) - 1))
    kf.observations.append((kf.observations[0][0], len(kf.keypoints) - 1))     # duplicate pair: the last keypoint for both rows
    kf.observations.append((10 ** 6, 0))                                       # not in the map: dropped
    assert check() == (3, 1)
    local[:] = sorted(gmap.keyframes)[-5:]                                     # the window slides by one keyframe
    assert check() == (len(gmap.keyframes[local[-1]].observations), 0)
    gone = kf.observations[1][0]
    del gmap.map_points[gone]                                                  # a point leaves the map: its rows go
    assert check() == (0, 0)


def test_resident_runs_equal_host_runs_on_a_growing_map():
    """Successive ``run`` calls while keyframes and points are appended (what the pipeline does between calls): the resident
    path and the host-flattening path leave bit-identical maps, the resident one uploading only the new rows / points -
    positions of old points never travel again because ``ba_map_commit`` keeps the device's copy in step."""
    from bundle_adjustment_b200 import synthetic as S
    from bundle_adjustment_b200.map_structures import Keyframe, KeyPoint, Map, MapPoint
    spec = S.make_window_spec(n_kfs=10, n_pts=1200, seed=5, duplicates=2)
    full = [mirror_map(spec), mirror_map(spec)]
    maps = [Map(), Map()]
    adjs = [BundleAdjuster(spec["camera_matrix"], window_size=5, resident=True),
            BundleAdjuster(spec["camera_matrix"], window_size=5, resident=False)]
    kf_ids = sorted(full[0].keyframes)
    uploads = []
    for step, kf_id in enumerate(kf_ids):
        for gm, src in zip(maps, full):
            kf = src.keyframes[kf_id]
            for mp_id, _ in kf.observations:                    # the points this keyframe brings (first sighting)
                if mp_id in src.map_points and mp_id not in gm.map_points:
                    gm.add_map_point(src.map_points[mp_id])
            gm.add_keyframe(kf)
        if step < 5:
            continue
        texts = []
        for adj, gm in zip(adjs, maps):
            with redirect_stdout(io.StringIO()) as out:
                adj.run(gm)
            texts.append(out.getvalue())
        assert texts[0] == texts[1] and "LBA Complete" in texts[0]
        uploads.append(dict(adjs[0].last_upload))
        for k in maps[0].keyframes:
            assert np.array_equal(maps[0].keyframes[k].R, maps[1].keyframes[k].R)
            assert np.array_equal(maps[0].keyframes[k].t, maps[1].keyframes[k].t)
        for k in maps[0].map_points:
            assert np.array_equal(maps[0].map_points[k].position, maps[1].map_points[k].position)
    assert adjs[1].last_upload is None
    # first window: five keyframes; afterwards only the keyframe that entered the window (the newest one is not optimised,
    # so the one uploaded is the previous call's newest) and the points it brought
    assert uploads[0]["rows_appended"] == sum(len(maps[0].keyframes[k].observations) for k in kf_ids[:5])
    for step, up in enumerate(uploads[1:], start=1):
        entered = kf_ids[4 + step]
        assert up["rows_appended"] == len(maps[0].keyframes[entered].observations)
        assert 0 <= up["points_set"] <= len(maps[0].keyframes[entered].observations) + len(maps[0].keyframes[kf_ids[5 + step]].observations)
    assert sum(u["points_set"] for u in uploads) == len(maps[0].map_points)
