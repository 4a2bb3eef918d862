"""CPU: host-side logic of the drop-in (no compute calls), the C-ABI library's exports, sharding."""
import ctypes
import io
import os
import re
from contextlib import redirect_stdout

import numpy as np
import pytest

from conftest import ROOT, golden_names, load_golden, mirror_map
import bundle_adjustment_b200.bundle_adjuster
import bundle_adjustment_b200
from bundle_adjustment_b200 import BundleAdjuster, _capi, geometry, sharding, synthetic as S
from bundle_adjustment_b200.map_structures import Keyframe, KeyPoint, Map, MapPoint
from oracle import ba_oracle as O


def test_library_exports_every_declared_symbol():
    """libba_b200.so loads without a GPU and exports exactly what include/ba_b200.h declares."""
    header = open(os.path.join(ROOT, "include", "ba_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(ba_[a-z_0-9]+)\s*\(", header))
    assert declared, "no declarations parsed"
    assert declared == set(_capi.SYMBOLS), (declared ^ set(_capi.SYMBOLS))
    assert os.path.isfile(_capi.LIB_PATH), "build the library first (python -m bundle_adjustment_b200.build)"
    lib = ctypes.CDLL(_capi.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name


def test_default_options_without_gpu():
    o = _capi.Options()
    _capi.lib().ba_default_options(ctypes.byref(o))
    assert o.max_iters == 50 and o.huber_delta == 1.0 and o.robust_mode == _capi.ROBUST["irls"]
    assert o.precond == _capi.PRECOND["auto"] and o.pcg_max_iters > 0
    # the reference's own tolerances (src/bundle_adjuster.py:172)
    assert o.ftol == 1e-5 and o.xtol == 1e-5 and o.gtol == 1e-8


def test_struct_sizes_match_header():
    # int32 + 3 doubles ... layout must match the C structs (natural alignment)
    assert ctypes.sizeof(_capi.Options) % 8 == 0
    assert ctypes.sizeof(_capi.Stats) == 6 * 4 + 7 * 8 + 8 * 8


def test_struct_layout_against_the_compiled_header(tmp_path):
    """The ctypes mirrors of ba_options / ba_stats against the header itself, compiled with gcc: size and the offset of
    every field, and the defaults of the fields added last."""
    import shutil
    import subprocess
    if not shutil.which("gcc"):
        pytest.skip("no gcc")
    fields = {"ba_options": [n for n, _ in _capi.Options._fields_], "ba_stats": [n for n, _ in _capi.Stats._fields_]}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "ba_b200.h"', 'int main(void) {']
    for st, names in fields.items():
        lines.append(f'printf("{st} %zu\\n", sizeof({st}));')
        lines += [f'printf("{st}.{n} %zu\\n", offsetof({st}, {n}));' for n in names]
    lines += ['return 0; }']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.run(["gcc", "-I", os.path.join(root, "include"), "-o", str(exe), str(src)], check=True)
    out = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    for st, cls in (("ba_options", _capi.Options), ("ba_stats", _capi.Stats)):
        assert int(out[st]) == ctypes.sizeof(cls), st
        for n, _ in cls._fields_:
            assert int(out[f"{st}.{n}"]) == getattr(cls, n).offset, (st, n)
    o = _capi.Options()
    _capi.lib().ba_default_options(ctypes.byref(o))
    assert o.pcg_tail == 1 and o.device_loop == 1 and o.coarse_aggregate == 0 and o.coarse_after == 0 and o.point_linesearch == 1 and o.dense_schur == 1


def test_run_host_path_with_a_stub_solver():
    """Everything of ``run`` around the device calls (window selection, packing, write-back as (3, 1) arrays, the printed
    line), with the solver replaced by a stub that shifts every point by one: no GPU needed."""
    from bundle_adjustment_b200.map_structures import Keyframe, KeyPoint, Map, MapPoint
    spec = S.make_window_spec(n_kfs=10, n_pts=400)
    gmap = S.build_map(spec, Map, Keyframe, MapPoint, KeyPoint.from_xy)

    class Stub:
        def set_problem(self, fx, fy, cx, cy, rvecs, tvecs, frv, ftv, pts, oc, op, ouv):
            assert pts.dtype == np.float64 and pts.shape[1] == 3 and oc.min() == -1 and op.max() == len(pts) - 1
            self.n_cams, self.x = len(rvecs), np.concatenate([rvecs.ravel(), tvecs.ravel(), (pts + 1.0).ravel()])

        def solve(self, **kw):
            return {"initial_raw_cost": 2.0, "final_raw_cost": 1.0}

        def get_x(self):
            return self.x

    adj = BundleAdjuster(spec["camera_matrix"], window_size=5, resident=False)
    stub = Stub()
    adj._get_solver = lambda: stub
    before = {k: v.position.copy() for k, v in gmap.map_points.items()}
    # positions in the forms a caller may hold them: (3, 1) arrays (the pipeline), nested lists, flat arrays
    ids = sorted(gmap.keyframes)[-6:-1]
    seen = sorted({mp for kf in ids for mp, _ in gmap.keyframes[kf].observations})
    gmap.map_points[seen[0]].position = before[seen[0]].tolist()
    gmap.map_points[seen[1]].position = before[seen[1]].ravel()
    with redirect_stdout(io.StringIO()) as out:
        adj.run(gmap)
    assert "LBA Complete. Initial Cost: 2.00, Final Cost: 1.00, Improvement: 50.00%" in out.getvalue()
    assert stub.n_cams == 4
    for k, mp in gmap.map_points.items():
        assert mp.position.shape == (3, 1) and mp.position.dtype == np.float64
        want = before[k] + 1.0 if k in seen else before[k]
        assert np.array_equal(mp.position, want), k
    # write-back hands out independent rows: an in-place edit of one position leaves its neighbours alone
    gmap.map_points[seen[2]].position += 5.0
    assert np.array_equal(gmap.map_points[seen[3]].position, before[seen[3]] + 1.0)


def test_host_spd_inverse_against_numpy(tmp_path):
    """csrc/ba_host_math.h (inverse of the two-level preconditioner's coarse matrix), compiled with g++, against numpy:
    a well-conditioned and a badly scaled SPD matrix, and the refusal of an indefinite one."""
    import shutil
    import subprocess
    if not shutil.which("g++"):
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "inv.cpp"
    src.write_text('''
#include "ba_host_math.h"
#include <cstdio>
int main(int argc, char** argv) {
    int n = 0;
    FILE* f = fopen(argv[1], "rb");
    if (fread(&n, sizeof n, 1, f) != 1) return 2;
    std::vector<double> A((size_t)n * n), Ainv;
    if (fread(A.data(), sizeof(double), A.size(), f) != A.size()) return 2;
    fclose(f);
    const bool ok = ba::spd_inverse_host(A, n, Ainv);
    f = fopen(argv[2], "wb");
    int flag = ok ? 1 : 0;
    fwrite(&flag, sizeof flag, 1, f);
    if (ok) fwrite(Ainv.data(), sizeof(double), Ainv.size(), f);
    fclose(f);
    return 0;
}
''')
    exe = tmp_path / "inv"
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", os.path.join(root, "bundle_adjustment_b200", "csrc"), "-o", str(exe), str(src)], check=True)

    def run(A):
        n = A.shape[0]
        (tmp_path / "a.bin").write_bytes(np.int32(n).tobytes() + np.ascontiguousarray(A, dtype=np.float64).tobytes())
        subprocess.run([str(exe), str(tmp_path / "a.bin"), str(tmp_path / "o.bin")], check=True)
        raw = (tmp_path / "o.bin").read_bytes()
        ok = np.frombuffer(raw[:4], dtype=np.int32)[0] == 1
        return ok, (np.frombuffer(raw[4:], dtype=np.float64).reshape(n, n) if ok else None)

    rng = np.random.default_rng(3)
    B = rng.normal(size=(90, 90))
    A = B @ B.T + 90 * np.eye(90)
    A_noisy = A + 1e-12 * rng.normal(size=A.shape)            # probing leaves rounding-level asymmetry
    ok, Ai = run(A_noisy)
    assert ok and np.abs(Ai - np.linalg.inv(A)).max() <= 1e-9 * np.abs(np.linalg.inv(A)).max()
    assert np.array_equal(Ai, Ai.T)
    D = np.diag(10.0 ** rng.uniform(-4, 4, 90))                # rotation and translation blocks differ by orders of magnitude
    ok, Ai = run(D @ A @ D)
    assert ok and np.abs(Ai @ (D @ A @ D) - np.eye(90)).max() <= 1e-6
    A[5, 5] = -1.0
    ok, _ = run(A)
    assert not ok


def test_geometry_matches_oracle():
    rng = np.random.default_rng(1)
    for _ in range(100):
        r = rng.normal(0, 1, 3) * rng.choice([1e-10, 1e-2, 1.0, 2.5])
        R = geometry.rvec_to_matrix(r)
        assert np.abs(R - O.rodrigues(r)).max() < 1e-15
        assert np.abs(geometry.matrix_to_rvec(R) - O.rotation_to_rvec(R)).max() < 1e-13
        if np.linalg.norm(r) < 3.0:          # beyond pi the vector wraps
            assert np.abs(geometry.matrix_to_rvec(R) - r).max() < 1e-9


@pytest.mark.parametrize("name", golden_names())
def test_gather_and_indexing_match_reference(name):
    g, spec = load_golden(name)
    gmap = mirror_map(spec)
    adj = BundleAdjuster(spec["camera_matrix"], window_size=int(g["window_size"]))
    mp_ids, obs, kp2d = adj._gather_local_data(gmap, g["local_kf_ids"].tolist())
    assert mp_ids == g["mp_ids"].tolist()
    assert [o[0] for o in obs] == g["obs_kf"].tolist() and [o[1] for o in obs] == g["obs_mp"].tolist()
    assert np.array_equal(np.array([kp2d[o] for o in obs]), g["obs_uv"])
    cam, pt, uv = adj._index_observations(g["adj_kf_ids"].tolist(), mp_ids, obs, kp2d, int(g["fixed_kf_id"]))
    flat = O.flatten_window(gmap, int(g["window_size"]))
    assert np.array_equal(cam, flat["obs_cam"]) and np.array_equal(pt, flat["obs_pt"]) and np.array_equal(uv, flat["obs_uv"])


def test_update_map_shapes_and_values():
    g, spec = load_golden("window_short")
    gmap = mirror_map(spec)
    adj = BundleAdjuster(spec["camera_matrix"])
    adj._update_map(gmap, g["res_x"], g["adj_kf_ids"].tolist(), g["mp_ids"].tolist())
    all_ids = sorted(gmap.keyframes)
    R = np.array([gmap.keyframes[i].R for i in all_ids])
    t = np.array([gmap.keyframes[i].t for i in all_ids])
    pos = np.array([gmap.map_points[i].position for i in sorted(gmap.map_points)])
    assert t.shape[1:] == (3, 1) and pos.shape[1:] == (3, 1)
    assert np.abs(R - g["out_kf_R"]).max() < 1e-13        # cv2.Rodrigues vs numpy
    assert np.array_equal(t, g["out_kf_t"]) and np.array_equal(pos, g["out_mp_pos"])


def test_run_skips_like_the_reference():
    K = S.CAMERA_MATRIX
    adj = BundleAdjuster(K, window_size=5)
    gmap = Map()
    for i in range(3):
        gmap.add_keyframe(Keyframe(id=i, R=np.eye(3), t=np.zeros((3, 1))))
    out = io.StringIO()
    with redirect_stdout(out):
        assert adj.run(gmap) is None
    assert "LBA Skipped: Not enough keyframes." in out.getvalue()
    adj.window_size = 1                    # window = ids[-2:-1] -> one keyframe, nothing adjustable
    out = io.StringIO()
    with redirect_stdout(out):
        adj.run(gmap)
    assert "LBA Skipped: No adjustable keyframes." in out.getvalue()
    adj.window_size = 2                    # adjustable keyframes but no observations at all
    out = io.StringIO()
    with redirect_stdout(out):
        adj.run(gmap)
    assert "LBA Skipped: No points in the local window." in out.getvalue()


def test_map_rejects_duplicate_ids():
    gmap = Map()
    gmap.add_keyframe(Keyframe(id=0, R=np.eye(3), t=np.zeros((3, 1))))
    with pytest.raises(ValueError):
        gmap.add_keyframe(Keyframe(id=0, R=np.eye(3), t=np.zeros((3, 1))))
    gmap.add_map_point(MapPoint(id=4, position=np.zeros((3, 1))))
    with pytest.raises(ValueError):
        gmap.add_map_point(MapPoint(id=4, position=np.zeros((3, 1))))
    assert KeyPoint.from_xy(0.1, 0.2).pt == (float(np.float32(0.1)), float(np.float32(0.2)))


def test_synthetic_shapes_and_order():
    fp = S.make_named_problem("bal50")
    assert (fp.n_cams + 1, fp.n_pts, fp.n_obs) == S.SHAPES["bal50"]
    assert np.all(np.diff(fp.obs_cam) >= 0), "observations must be keyframe-major"
    assert fp.obs_cam.min() == -1 and fp.obs_cam.max() == fp.n_cams - 1
    assert np.bincount(fp.obs_pt, minlength=fp.n_pts).min() >= 2
    key = fp.obs_pt.astype(np.int64) * (fp.n_cams + 2) + fp.obs_cam + 1
    assert np.unique(key).size == fp.n_obs, "a point is seen at most once per camera"
    assert np.array_equal(fp.obs_uv, fp.obs_uv.astype(np.float32).astype(np.float64))
    fp2 = S.make_named_problem("bal50")
    assert np.array_equal(fp.obs_uv, fp2.obs_uv) and np.array_equal(fp.pts, fp2.pts)     # seeded


def test_partition_points_balanced_and_complete():
    fp = S.make_flat_problem(20, 5000, 24000, seed=9)
    for n in (1, 2, 3, 8):
        b = sharding.partition_points(fp.obs_pt, fp.n_pts, n)
        assert b[0] == 0 and b[-1] == fp.n_pts and np.all(np.diff(b) >= 0) and b.size == n + 1
        counts = np.bincount(fp.obs_pt, minlength=fp.n_pts)
        per = np.array([counts[b[r]:b[r + 1]].sum() for r in range(n)])
        assert per.sum() == fp.n_obs
        assert per.max() - per.min() <= 2 * counts.max()
        total = 0
        for r in range(n):
            sh, (lo, hi) = sharding.shard_problem(fp, r, n)
            assert sh.n_pts == hi - lo and sh.n_cams == fp.n_cams
            assert sh.obs_pt.min() >= 0 and sh.obs_pt.max() < sh.n_pts
            total += sh.n_obs
        assert total == fp.n_obs


def test_partition_points_never_leaves_a_rank_empty():
    """Skewed tracks (one point holds most observations) put several cuts on the same point; every rank still gets a point,
    and more ranks than points is the same ValueError on every rank (an empty shard would fail ba_set_problem on one rank only
    and leave the others waiting in a collective)."""
    obs_pt = np.concatenate([np.zeros(1000, dtype=np.int64), np.arange(1, 6)])
    b = sharding.partition_points(obs_pt, 6, 4)
    assert b[0] == 0 and b[-1] == 6 and np.all(np.diff(b) >= 1)
    with pytest.raises(ValueError, match="at least one point"):
        sharding.partition_points(obs_pt, 6, 7)


def test_injected_into_the_unmodified_reference_pipeline(tmp_path, monkeypatch, capsys):
    """SURVEY 8f-1 on the CPU side: the reference's VisualOdometryPipeline constructs and calls the drop-in
    class exactly as it does its own (src/pipeline.py:39, :99).  Real frames (desk_images) through the
    unmodified front end; the two calls made before five keyframes exist are skips, which need no device;
    the global-BA call of src/main.py:81-87 reaches the device library, which has no CPU path and must say so."""
    from oracle.reference_loader import load_reference, reference_available
    if not reference_available():
        pytest.skip("reference sources are not on this machine (GPU box)")
    cv2 = pytest.importorskip("cv2")
    import bundle_adjustment_b200
    from bundle_adjustment_b200 import _capi, bundle_adjuster as mine
    load_reference()
    import features, parameters, pipeline
    monkeypatch.setattr(pipeline, "BundleAdjuster", mine.BundleAdjuster)
    monkeypatch.chdir(tmp_path)
    for d in [parameters.OUTPUT_DIR, parameters.TRAJECTORY_DIR_2D, parameters.TRAJECTORY_DIR_3D, *parameters.DEBUG_DIRS.values()]:
        os.makedirs(d, exist_ok=True)
    K = np.array([[912.7820434570312, 0.0, 650.2929077148438], [0.0, 913.0294189453125, 362.7241516113281], [0.0, 0.0, 1.0]])
    crit = {"min_pixel_displacement": 20.0, "min_rotation": 0.15, "min_feature_ratio": 0.25, "min_parallax_deg": 1.0,
            "min_tracked_for_parallax": 20}
    vo = pipeline.VisualOdometryPipeline(camera_matrix=K, dist_coeffs=np.zeros(5), feature_extractor=features.ORBExtractor(n_features=4000),
                                         feature_matcher=features.BruteForceMatcher(), keyframe_criteria=crit)
    assert isinstance(vo.bundle_adjuster, mine.BundleAdjuster)
    assert vo.bundle_adjuster.window_size == parameters.BA_WINDOW_SIZE and vo.bundle_adjuster.camera_matrix is K
    for n in (11, 22, 33):
        vo.process_frame(cv2.imread(f"/root/reference/desk_images/image{n}.png"))
    text = capsys.readouterr().out
    assert vo.map.next_keyframe_id == 3 and len(vo.map.map_points) > 100
    assert text.count("--- Running Local Bundle Adjustment ---") == 2
    assert text.count("-> LBA Skipped: Not enough keyframes.") == 2
    import torch
    if not torch.cuda.is_available():
        vo.bundle_adjuster.window_size = vo.map.next_keyframe_id
        with pytest.raises(_capi.BaError, match="no CPU path"):
            vo.bundle_adjuster.run(vo.map)


@pytest.mark.parametrize("name", golden_names())
def test_incremental_flattening_equals_the_reference_gather(name):
    """the cached, vectorised window flattening that run() uses produces exactly what the
    reference-contract helpers produce (_gather_local_data + row indexing): ids, row order, duplicates with the
    last keypoint, dangling observations dropped."""
    g, spec = load_golden(name)
    gmap = mirror_map(spec)
    adj = BundleAdjuster(spec["camera_matrix"], window_size=int(g["window_size"]))
    local = g["local_kf_ids"].tolist()
    fixed, adjustable = local[0], local[1:]
    for rep in range(2):                                         # second pass is served from the cache
        mp_ids, obs, kp = adj._gather_local_data(gmap, local)
        oc, op, ouv = adj._index_observations(adjustable, mp_ids, obs, kp, fixed)
        ids2, oc2, op2, ouv2 = adj._gather_flat(gmap, local, fixed)
        assert ids2 == mp_ids == g["mp_ids"].tolist()
        assert np.array_equal(oc2, oc) and np.array_equal(op2, op) and np.array_equal(ouv2, ouv)
        assert oc2.dtype == np.int32 and op2.dtype == np.int32 and ouv2.dtype == np.float64


def test_incremental_flattening_follows_a_growing_map():
    """Observation lists grow between calls (the pipeline appends to the last keyframe when it triangulates,
    src/pipeline.py:305-306); cached prefixes are reused, tails converted, replaced lists rebuilt."""
    spec = S.make_window_spec(n_kfs=7, n_pts=200, seed=11, duplicates=3, dangling=2)
    gmap = mirror_map(spec)
    adj = BundleAdjuster(spec["camera_matrix"], window_size=5)
    local = sorted(gmap.keyframes)[-6:-1]

    def both():
        mp_ids, obs, kp = adj._gather_local_data(gmap, local)
        ref = adj._index_observations(local[1:], mp_ids, obs, kp, local[0])
        got = adj._gather_flat(gmap, local, local[0])
        assert got[0] == mp_ids
        for a, b in zip(ref, got[1:]):
            assert np.array_equal(a, b)

    both()
    kf = gmap.keyframes[local[2]]
    new_id = gmap.next_map_point_id
    gmap.add_map_point(MapPoint(id=new_id, position=np.array([[0.1], [0.2], [9.0]])))
    kf.keypoints.append(KeyPoint.from_xy(321.5, 222.25))
    kf.observations.append((new_id, len(kf.keypoints) - 1))      # appended: tail conversion
    kf.observations.append((kf.observations[0][0], len(kf.keypoints) - 1))   # duplicate pair: last keypoint wins for both rows
    both()
    kf2 = gmap.keyframes[local[3]]
    kf2.observations = list(reversed(kf2.observations))          # another list object: entry rebuilt
    both()
    kf2.observations.pop()                                       # shorter: rebuilt
    both()
    kf.observations.append((10 ** 6, 0))                         # observation of a map point that is not in the map: dropped
    both()


class _ResidentStub:
    """Stands in for FlatSolver in the CPU tests of the resident path: the map calls are the CPU statement of ``ba_map_*``
    (oracle/map_window.py); "solving" shifts every point by one."""

    def __new__(cls):
        from oracle.map_window import MapStore

        class Stub(MapStore):
            def solve(self, **kw):
                return {"initial_raw_cost": 2.0, "final_raw_cost": 1.0}

            def get_params(self):
                return self.flat["rvec"], self.flat["tvec"], self.flat["pts"]

            def get_x(self):
                f = self.flat
                return np.concatenate([f["rvec"].ravel(), f["tvec"].ravel(), (f["pts"] + 1.0).ravel()])

            def map_commit(self):
                MapStore.map_commit(self, self.flat["pts"] + 1.0)
        return Stub()


def test_resident_window_equals_the_host_flattening_and_uploads_only_the_delta():
    """SURVEY 8f-2, host logic of the device-resident map (no GPU: the device calls go to their CPU statement): on a map that
    grows between calls the window the resident path builds equals what the host flattening builds (ids, rows, order,
    duplicate pairs with the last keypoint, dangling observations dropped, positions, poses), and every call uploads
    exactly what was appended since the last one."""
    from bundle_adjustment_b200.geometry import matrix_to_rvec
    spec = S.make_window_spec(n_kfs=7, n_pts=200, seed=11, duplicates=3, dangling=2)
    gmap = mirror_map(spec)
    adj = BundleAdjuster(spec["camera_matrix"], window_size=5)
    stub = _ResidentStub()
    adj._get_solver = lambda: stub
    local = sorted(gmap.keyframes)[-6:-1]

    def both():
        before = stub.map_stats()
        ids = adj._window_resident(gmap, local)
        after = stub.map_stats()
        ids_h, oc, op, ouv = adj._gather_flat(gmap, local, local[0])
        f = stub.flat
        assert ids == ids_h and f["ids"].tolist() == ids_h
        assert np.array_equal(f["obs_cam"], oc) and np.array_equal(f["obs_pt"], op) and np.array_equal(f["obs_uv"], ouv)
        assert np.array_equal(f["pts"], np.array([gmap.map_points[i].position.ravel() for i in ids]))
        assert np.array_equal(f["fixed_rvec"], matrix_to_rvec(gmap.keyframes[local[0]].R))
        assert np.array_equal(f["tvec"], np.array([gmap.keyframes[i].t.ravel() for i in local[1:]]))
        return after["rows_appended"] - before["rows_appended"], after["points_set"] - before["points_set"]

    n_rows = sum(len(gmap.keyframes[k].observations) for k in local)
    # the first call uploads the window's keyframes and every point of the map
    assert both() == (n_rows, len(gmap.map_points))
    assert both() == (0, 0)                                      # nothing changed: nothing uploaded
    kf = gmap.keyframes[local[2]]
    new_id = gmap.next_map_point_id
    gmap.add_map_point(MapPoint(id=new_id, position=np.array([[0.1], [0.2], [9.0]])))
    kf.keypoints.append(KeyPoint.from_xy(321.5, 222.25))
    kf.observations.append((new_id, len(kf.keypoints) - 1))
    kf.observations.append((kf.observations[0][0], len(kf.keypoints) - 1))   # duplicate pair: last keypoint wins for both rows
    assert both() == (2, 1)                                      # two appended rows, one new point
    kf.observations.append((10 ** 6, 0))                         # observation of a map point that is not in the map: dropped
    assert both() == (1, 0)
    kf2 = gmap.keyframes[local[3]]
    kf2.observations = list(reversed(kf2.observations))          # another list object: the resident copy is rebuilt
    rows, pts = both()
    assert rows == sum(len(gmap.keyframes[k].observations) for k in local) and pts == len(gmap.map_points)
    kf2.observations.pop()                                       # shorter: rebuilt again
    assert both()[0] == sum(len(gmap.keyframes[k].observations) for k in local)
    # a point leaves the map (`mp_id in gmap.map_points`, src/bundle_adjuster.py:213): its rows go
    gone = kf.observations[1][0]
    del gmap.map_points[gone]
    assert both() == (0, 0) and gone not in stub.flat["ids"].tolist()
    # the window slides: only the keyframe that enters is uploaded
    local[:] = sorted(gmap.keyframes)[-5:]
    assert both() == (len(gmap.keyframes[local[-1]].observations), 0)
    # ids that cannot index a dense table: the call says so and run() flattens on the host
    with pytest.raises(bundle_adjustment_b200.bundle_adjuster._NotResident):
        adj._small_ids([3, 1 << 30])


def test_resident_run_keeps_the_device_positions_in_step_with_the_map():
    """run() on the resident path with the stub: the write-back reaches the map AND the resident positions (ba_map_commit),
    so the next window starts from what the last one produced without any position upload; "verify" re-uploads a position
    that somebody else replaced."""
    from bundle_adjustment_b200.map_structures import Keyframe, KeyPoint as KP, Map, MapPoint as MP
    spec = S.make_window_spec(n_kfs=10, n_pts=400)
    gmap = S.build_map(spec, Map, Keyframe, MP, KP.from_xy)
    adj = BundleAdjuster(spec["camera_matrix"], window_size=5, resident="verify")
    stub = _ResidentStub()
    adj._get_solver = lambda: stub
    before = {k: v.position.copy() for k, v in gmap.map_points.items()}
    with redirect_stdout(io.StringIO()) as out:
        adj.run(gmap)
    assert "LBA Complete. Initial Cost: 2.00, Final Cost: 1.00, Improvement: 50.00%" in out.getvalue()
    ids = stub.flat["ids"].tolist()
    assert adj.last_upload["points_set"] == len(gmap.map_points) and adj.last_upload["rows_appended"] == stub.flat["obs_cam"].shape[0]
    for k, mp in gmap.map_points.items():
        assert mp.position.shape == (3, 1)
        assert np.array_equal(mp.position, before[k] + 1.0 if k in ids else before[k])
        assert np.array_equal(stub.xyz[k], mp.position.ravel())
    with redirect_stdout(io.StringIO()):
        adj.run(gmap)
    assert adj.last_upload["points_set"] == 0 and adj.last_upload["rows_appended"] == 0
    assert np.array_equal(gmap.map_points[ids[0]].position, before[ids[0]] + 2.0)
    gmap.map_points[ids[0]].position = np.array([[1.0], [2.0], [3.0]])          # replaced behind the adjuster's back
    with redirect_stdout(io.StringIO()):
        adj.run(gmap)
    assert adj.last_upload["points_set"] == 1
    assert np.array_equal(gmap.map_points[ids[0]].position, np.array([[2.0], [3.0], [4.0]]))


def test_flat_import_through_the_shim():
    """``from bundle_adjuster import BundleAdjuster`` - the reference's own import line (src/pipeline.py:12) - resolves to the
    drop-in when bundle_adjustment_b200/shim is ahead on sys.path (run in a fresh interpreter)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import sys; sys.path.insert(0, %r); from bundle_adjuster import BundleAdjuster; import inspect; "
            "print(BundleAdjuster.__module__, list(inspect.signature(BundleAdjuster.__init__).parameters)[:3])"
            % os.path.join(root, "bundle_adjustment_b200", "shim"))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, check=True, cwd="/").stdout
    assert "bundle_adjustment_b200.bundle_adjuster" in out and "'camera_matrix', 'window_size'" in out
