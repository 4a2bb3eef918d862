"""BAL text files <-> the flat problem.  CPU: reader / writer / model mapping against the slow
restatement in oracle/bal_model.py.  GPU: a BAL file read, solved on the device, written back."""
import numpy as np
import pytest

from bundle_adjustment_b200 import bal_io as B, synthetic as S
from bundle_adjustment_b200.geometry import matrix_to_rvec, rvec_to_matrix
from oracle import ba_oracle as O, bal_model as M


def synthetic_bal(n_cams=12, n_pts=600, n_obs=2600, seed=3, distortion=True, noise=0.5):
    """A BAL problem whose truth is one of the repo's synthetic scenes: per-camera focal lengths, radial terms."""
    fp = S.make_flat_problem(n_cams, n_pts, n_obs, seed)
    rng = np.random.default_rng(seed + 100)
    nc = fp.n_cams + 1
    flip = np.diag([1.0, -1.0, -1.0])
    rv0 = np.vstack([fp.fixed_rvec[None], fp.rvec])
    tv0 = np.vstack([fp.fixed_tvec[None], fp.tvec])
    cams = np.zeros((nc, 9))
    for c in range(nc):
        cams[c, 0:3] = matrix_to_rvec(flip @ rvec_to_matrix(rv0[c]))
        cams[c, 3:6] = flip @ tv0[c]
    cams[:, 6] = rng.uniform(800.0, 1000.0, nc)
    if distortion:
        cams[:, 7] = rng.normal(0.0, 1e-2, nc)
        cams[:, 8] = rng.normal(0.0, 1e-3, nc)
    ci = (fp.obs_cam + 1).astype(np.int32)
    # truth poses for the measurements: the scene's true cameras
    rvt = np.vstack([fp.fixed_rvec[None], fp.truth["rvec"]])
    tvt = np.vstack([fp.fixed_tvec[None], fp.truth["tvec"]])
    cams_true = cams.copy()
    for c in range(nc):
        cams_true[c, 0:3] = matrix_to_rvec(flip @ rvec_to_matrix(rvt[c]))
        cams_true[c, 3:6] = flip @ tvt[c]
    uv = B.bal_project(cams_true, fp.truth["pts"], ci, fp.obs_pt) + rng.normal(0.0, noise, (fp.n_obs, 2))
    return B.BalProblem(ci, fp.obs_pt.astype(np.int32), uv, cams, fp.pts.copy())


def _oracle_problem(fp):
    return O.Problem(fp.fx, fp.fy, fp.cx, fp.cy, fp.n_cams, fp.n_pts, fp.fixed_rvec, fp.fixed_tvec, fp.obs_cam, fp.obs_pt,
                     fp.obs_uv)


@pytest.mark.parametrize("suffix", ["", ".bz2", ".gz"])
def test_write_read_round_trip_is_bit_exact(tmp_path, suffix):
    bal = synthetic_bal()
    path = str(tmp_path / ("problem-13-600-pre.txt" + suffix))
    B.write_bal(path, bal)
    back = B.read_bal(path)
    for name in ("cam_index", "pt_index", "uv", "cameras", "points"):
        assert np.array_equal(getattr(bal, name), getattr(back, name)), name
    assert back.cam_index.dtype == np.int32 and back.pt_index.dtype == np.int32


def test_fast_reader_equals_token_parser(tmp_path):
    bal = synthetic_bal(n_cams=5, n_pts=60, n_obs=200)
    path = str(tmp_path / "p.txt")
    B.write_bal(path, bal)
    ci, pi, uv, cams, pts = M.parse_bal_slow(path)
    got = B.read_bal(path)
    assert np.array_equal(got.cam_index, ci) and np.array_equal(got.pt_index, pi)
    assert np.array_equal(got.uv, uv) and np.array_equal(got.cameras, cams) and np.array_equal(got.points, pts)


def test_reader_rejects_malformed_files(tmp_path):
    p = tmp_path / "bad.txt"
    p.write_text("2 3 1\n0 0 1.0 2.0\n" + "0.0\n" * 10)           # too few numbers
    with pytest.raises(ValueError, match="header says"):
        B.read_bal(str(p))
    p.write_text("1 1 1\n0 5 1.0 2.0\n" + "0.0\n" * 12)           # point index out of range
    with pytest.raises(ValueError, match="out of range"):
        B.read_bal(str(p))
    p.write_text("")
    with pytest.raises(ValueError, match="no header"):
        B.read_bal(str(p))


def test_empty_problem_round_trip(tmp_path):
    bal = B.BalProblem(np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros((0, 2)), np.zeros((0, 9)), np.zeros((0, 3)))
    path = str(tmp_path / "empty.txt")
    B.write_bal(path, bal)
    back = B.read_bal(path)
    assert back.n_cams == 0 and back.n_pts == 0 and back.n_obs == 0


def test_vectorised_bal_projection_equals_scalar_model():
    bal = synthetic_bal(n_cams=6, n_pts=80, n_obs=300)
    r_fast = bal.uv - B.bal_project(bal.cameras, bal.points, bal.cam_index, bal.pt_index)
    r_slow = M.residuals(bal.cam_index, bal.pt_index, bal.uv, bal.cameras, bal.points)
    assert np.abs(r_fast - r_slow).max() <= 1e-9 * np.abs(r_slow).max()


def test_undistort_inverts_the_radial_model():
    rng = np.random.default_rng(0)
    p = rng.uniform(-0.6, 0.6, (5000, 2))
    k1, k2 = rng.normal(0, 5e-2, 5000), rng.normal(0, 5e-3, 5000)
    n2 = np.sum(p * p, axis=1)
    d = p * (1.0 + k1 * n2 + k2 * n2 * n2)[:, None]
    assert np.abs(B.undistort(d, k1, k2) - p).max() <= 1e-13
    assert np.array_equal(B.undistort(np.zeros((3, 2)), np.ones(3), np.ones(3)), np.zeros((3, 2)))


@pytest.mark.parametrize("fixed", [0, 4])
def test_flat_residuals_are_the_scaled_bal_residuals_without_distortion(fixed):
    """k1 = k2 = 0: residual_flat = (f0 / f_c) * diag(1, -1) * residual_BAL, observation by observation."""
    bal = synthetic_bal(distortion=False)
    fp, mp = B.bal_to_flat(bal, fixed_camera=fixed)
    assert fp.n_cams == bal.n_cams - 1 and np.all(fp.obs_cam[bal.cam_index == fixed] == -1)
    r_flat = _oracle_problem(fp).residuals(fp.x0()).reshape(-1, 2)
    r_bal = M.residuals(bal.cam_index, bal.pt_index, bal.uv, bal.cameras, bal.points)
    scale = (mp.f0 / bal.cameras[bal.cam_index, 6])[:, None] * np.array([1.0, -1.0])
    assert np.abs(r_flat - r_bal * scale).max() <= 1e-9 * np.abs(r_bal).max()


def test_pose_mapping_round_trip():
    bal = synthetic_bal()
    fp, mp = B.bal_to_flat(bal, fixed_camera=2)
    back = B.flat_to_bal(bal, mp, fp.rvec, fp.tvec, fp.pts)
    assert np.abs(back.cameras - bal.cameras).max() <= 1e-12
    assert np.array_equal(back.points, bal.points) and back.uv is bal.uv


@pytest.mark.gpu
def test_bal_file_solved_on_the_device_and_written_back(tmp_path, gpu_solver):
    bal = synthetic_bal(n_cams=40, n_pts=6000, n_obs=28000, seed=9)
    path = str(tmp_path / "problem-41-6000-pre.txt.bz2")
    B.write_bal(path, bal)
    loaded = B.read_bal(path)
    fp, mp = B.bal_to_flat(loaded)
    gpu_solver.set_flat(fp)
    r0 = gpu_solver.residuals()
    ref0 = _oracle_problem(fp).residuals(fp.x0())
    assert np.linalg.norm(r0 - ref0) <= 1e-9 * np.linalg.norm(ref0)
    st = gpu_solver.solve(max_iters=40)
    assert st["final_cost"] < 0.05 * st["initial_cost"]
    rvec, tvec, pts = gpu_solver.get_params()
    solved = B.flat_to_bal(loaded, mp, rvec, tvec, pts)
    out = str(tmp_path / "solved.txt")
    B.write_bal(out, solved)
    again = B.read_bal(out)
    # the written file, evaluated with the BAL model itself, fits the measurements to the noise level
    rb = again.uv - B.bal_project(again.cameras, again.points, again.cam_index, again.pt_index)
    rms0 = np.sqrt(np.mean((bal.uv - B.bal_project(bal.cameras, bal.points, bal.cam_index, bal.pt_index)) ** 2))
    rms1 = np.sqrt(np.mean(rb ** 2))
    assert rms1 < 0.75 and rms1 < 0.2 * rms0, (rms0, rms1)
    assert np.array_equal(again.cameras[:, 6:], bal.cameras[:, 6:])            # intrinsics are held
    assert np.array_equal(again.cameras[mp.fixed_camera], bal.cameras[mp.fixed_camera])
