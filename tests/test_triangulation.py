"""Two-view triangulation + cheirality: oracle vs the reference's own outputs (CPU), device vs both (GPU).

Tolerances.  OpenCV returns float32 when the pixel arrays are float32 (the pipeline's case), so the desk cases
agree to float32 rounding (1e-5 relative); the float64 case to 1e-9 for the homogeneous vector (up to its sign)
and to 3e-6 / |w| for the 3-D points, because the reference divides by (w + 1e-6) and the sign of the unit
singular vector - OpenCV's SVD's choice - is not part of the problem.
"""
import io
import os
from contextlib import redirect_stdout

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import triangulation as T

CASES = ["desk0", "desk1", "syn"]


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN_DIR, "pipeline_triangulation.npz"))


def _proj(K, R, t):
    return K @ np.hstack((np.eye(3), np.zeros((3, 1)))), K @ np.hstack((R.reshape(3, 3), t.reshape(3, 1)))


def _check_hom(hom, ref, rel):
    ref = ref.astype(np.float64)
    ref = ref / np.linalg.norm(ref, axis=0)
    sign = np.sign(np.sum(hom * ref, axis=0))
    assert np.all(sign != 0)
    assert np.abs(hom * sign - ref).max() <= rel


def _check_points(X, idx, g, name, rel):
    assert np.array_equal(idx, g[f"{name}_idx"])
    Xr = g[f"{name}_X"].astype(np.float64)
    assert X.shape == Xr.shape
    w = np.abs(g[f"{name}_hom"].astype(np.float64)[3, idx]) / np.linalg.norm(g[f"{name}_hom"].astype(np.float64)[:, idx], axis=0)
    tol = rel + 3e-6 / w                                    # relative, per point: (w + 1e-6) with either sign of w
    err = np.linalg.norm(X - Xr, axis=0) / np.linalg.norm(Xr, axis=0)
    assert np.all(err <= tol), float((err / tol).max())


@pytest.mark.parametrize("name", CASES)
def test_oracle_against_opencv_and_the_reference_function(gold, name):
    K = gold["camera_matrix"]
    P1, P2 = _proj(K, gold[f"{name}_R"], gold[f"{name}_t"])
    hom = T.triangulate_points(P1, P2, gold[f"{name}_pts1"], gold[f"{name}_pts2"])
    f32 = gold[f"{name}_hom"].dtype == np.float32
    _check_hom(hom, gold[f"{name}_hom"], 2e-6 if f32 else 1e-9)
    X, mask = T.triangulate_two_view(K, gold[f"{name}_R"], gold[f"{name}_t"], gold[f"{name}_pts1"], gold[f"{name}_pts2"])
    _check_points(X[:, mask], np.where(mask)[0], gold, name, 1e-5 if f32 else 1e-9)


def test_synthetic_case_exercises_the_cheirality_mask(gold):
    assert 0 < len(gold["syn_idx"]) < gold["syn_pts1"].shape[0]


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_device_triangulation_against_reference_outputs(gold, gpu_solver, name):
    from bundle_adjustment_b200.triangulation import Triangulator
    K = gold["camera_matrix"]
    tri = Triangulator(K, solver=gpu_solver)
    P1, P2 = _proj(K, gold[f"{name}_R"], gold[f"{name}_t"])
    p1, p2 = gold[f"{name}_pts1"], gold[f"{name}_pts2"]
    hom = tri.triangulate_points(P1, P2, p1.T, p2.T)                       # cv2 calling convention: 2xN in, 4xN out
    assert hom.shape == (4, p1.shape[0])
    assert np.abs(np.linalg.norm(hom, axis=0) - 1.0).max() <= 1e-12
    f32 = gold[f"{name}_hom"].dtype == np.float32
    _check_hom(hom, gold[f"{name}_hom"], 2e-6 if f32 else 1e-9)
    _check_hom(hom, T.triangulate_points(P1, P2, p1, p2), 1e-9)            # and the float64 restatement
    out = io.StringIO()
    with redirect_stdout(out):
        X, idx = tri.triangulate_two_view(gold[f"{name}_R"], gold[f"{name}_t"], p1, p2)
    assert out.getvalue().strip() == str(gold[f"{name}_line"])              # the pipeline's log line, word for word
    _check_points(X, idx, gold, name, 1e-5 if f32 else 1e-9)


@pytest.mark.gpu
def test_device_triangulation_edge_cases(gpu_solver):
    from bundle_adjustment_b200.triangulation import Triangulator
    K = np.array([[900.0, 0.0, 640.0], [0.0, 900.0, 360.0], [0.0, 0.0, 1.0]])
    tri = Triangulator(K, solver=gpu_solver)
    assert tri.triangulate_two_view(np.eye(3), np.array([1.0, 0.0, 0.0]), np.zeros((0, 2)), np.zeros((0, 2))) == (None, None)
    # exact correspondences of known points: recovered to rounding, all kept
    rng = np.random.default_rng(1)
    Xw = np.stack([rng.uniform(-2, 2, 1000), rng.uniform(-1, 1, 1000), rng.uniform(3, 9, 1000)])
    R, t = np.eye(3), np.array([[-0.5], [0.0], [0.0]])
    P1, P2 = _proj(K, R, t)
    h = np.vstack([Xw, np.ones((1, 1000))])
    p1 = ((P1 @ h)[:2] / (P1 @ h)[2]).T
    p2 = ((P2 @ h)[:2] / (P2 @ h)[2]).T
    with redirect_stdout(io.StringIO()):
        X, idx = tri.triangulate_two_view(R, t, p1, p2)
    assert len(idx) == 1000
    hom = tri.triangulate_points(P1, P2, p1, p2)
    Xh = hom[:3] / hom[3]
    assert np.abs(Xh - Xw).max() <= 1e-9
    with pytest.raises(ValueError):
        tri.triangulate_points(P1, P2, p1, p2[:-1])
