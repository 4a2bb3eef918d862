"""Bundle-Adjustment-in-the-Large (BAL) text files <-> the flat problem of ``ba_set_problem``.

The reference has no file format for bundle-adjustment problems; its BASELINE shapes only borrow the
*sizes* of the public BAL datasets (Ladybug 1723/156k/680k, Venice 1778/993k/5M).  This module lets the
real files be used when they are present.  File layout (``problem-C-P-pre.txt``, optionally ``.bz2``/``.gz``)::

    <n_cameras> <n_points> <n_observations>
    <camera index> <point index> <x> <y>              n_observations lines
    <one number per line>                              9 per camera: rotation vector (3), t (3), f, k1, k2
    <one number per line>                              3 per point

BAL camera model:  P = R X + t,  p = -P[:2] / P[2],  uv = f (1 + k1 |p|^2 + k2 |p|^4) p  (image centre = origin;
the camera looks down -z).  The reference's model (``src/bundle_adjuster.py:58-69``) is a pinhole with ONE
fixed intrinsic matrix, looking down +z, no distortion, six degrees of freedom per camera.  The mapping
(``bal_to_flat``) therefore holds every camera's (f, k1, k2) at the file's values and moves them into the
measurements:

* frame flip  R' = S R, t' = S t  with S = diag(1, -1, -1)  (a proper rotation, so rvec' exists);
* each measurement is undistorted with its camera's (k1, k2) and rescaled from its camera's focal length to
  a common ``f0``:  uv' = f0 * (p_x, -p_y),  p = undistort(uv / f);
* intrinsics handed to the solver: fx = fy = f0, cx = cy = 0;  camera ``fixed_camera`` (default 0) is the
  fixed one, as the reference always fixes the oldest keyframe of its window (``:141``).

Residuals of the flat problem are the BAL residuals measured in the undistorted image and scaled by f0 / f_c.
``flat_to_bal`` maps optimised poses / points back (observations and intrinsics are kept).
"""
from __future__ import annotations

import bz2
import gzip
import io
from dataclasses import dataclass, replace
from typing import Optional

import numpy as np

from .geometry import matrix_to_rvec, rvec_to_matrix
from .synthetic import FlatProblem, rodrigues

_S = np.diag([1.0, -1.0, -1.0])


@dataclass
class BalProblem:
    cam_index: np.ndarray   # (No,) int32
    pt_index: np.ndarray    # (No,) int32
    uv: np.ndarray          # (No, 2) float64, BAL image coordinates
    cameras: np.ndarray     # (Nc, 9) float64: rvec, t, f, k1, k2
    points: np.ndarray      # (Np, 3) float64

    @property
    def n_cams(self) -> int:
        return int(self.cameras.shape[0])

    @property
    def n_pts(self) -> int:
        return int(self.points.shape[0])

    @property
    def n_obs(self) -> int:
        return int(self.cam_index.shape[0])


def _open_text(path: str, mode: str):
    if path.endswith(".bz2"):
        return bz2.open(path, mode + "t")
    if path.endswith(".gz"):
        return gzip.open(path, mode + "t")
    return open(path, mode)


def read_bal(path: str) -> BalProblem:
    """Parse a BAL file.  All numbers are read in one C-speed pass (``numpy.fromstring`` in text mode), so a
    5 M-observation file takes seconds, not minutes."""
    with _open_text(path, "r") as f:
        text = f.read()
    a = np.fromstring(text, dtype=np.float64, sep=" ")
    if a.size < 3:
        raise ValueError(f"{path}: not a BAL file (no header)")
    nc, npt, no = int(a[0]), int(a[1]), int(a[2])
    need = 3 + 4 * no + 9 * nc + 3 * npt
    if nc < 0 or npt < 0 or no < 0 or a.size != need:
        raise ValueError(f"{path}: header says {nc} cameras / {npt} points / {no} observations = {need} numbers, "
                         f"file holds {a.size}")
    obs = a[3:3 + 4 * no].reshape(no, 4)
    ci, pi = obs[:, 0].astype(np.int32), obs[:, 1].astype(np.int32)
    if no and (np.any(obs[:, 0] != ci) or np.any(obs[:, 1] != pi) or ci.min() < 0 or ci.max() >= nc
               or pi.min() < 0 or pi.max() >= npt):
        raise ValueError(f"{path}: observation indices out of range")
    cams = a[3 + 4 * no:3 + 4 * no + 9 * nc].reshape(nc, 9).copy()
    pts = a[3 + 4 * no + 9 * nc:].reshape(npt, 3).copy()
    return BalProblem(ci, pi, np.ascontiguousarray(obs[:, 2:4]), cams, pts)


def write_bal(path: str, bal: BalProblem) -> None:
    """Write a BAL file; floats are printed with 17 significant digits, so ``read_bal`` returns the same bits."""
    buf = io.StringIO()
    buf.write(f"{bal.n_cams} {bal.n_pts} {bal.n_obs}\n")
    if bal.n_obs:
        rows = np.empty((bal.n_obs, 4))
        rows[:, 0], rows[:, 1], rows[:, 2:] = bal.cam_index, bal.pt_index, bal.uv
        np.savetxt(buf, rows, fmt=["%d", "%d", "%.16e", "%.16e"])
    if bal.n_cams:
        np.savetxt(buf, bal.cameras.reshape(-1, 1), fmt="%.16e")
    if bal.n_pts:
        np.savetxt(buf, bal.points.reshape(-1, 1), fmt="%.16e")
    with _open_text(path, "w") as f:
        f.write(buf.getvalue())


def bal_project(cameras: np.ndarray, points: np.ndarray, cam_index: np.ndarray, pt_index: np.ndarray) -> np.ndarray:
    """BAL camera model for every observation -> (No, 2)."""
    cam = cameras[cam_index]
    R = rodrigues(cam[:, 0:3])
    P = np.einsum("nij,nj->ni", R, points[pt_index]) + cam[:, 3:6]
    p = -P[:, :2] / P[:, 2:3]
    n2 = np.sum(p * p, axis=1)
    r = 1.0 + cam[:, 7] * n2 + cam[:, 8] * n2 * n2
    return cam[:, 6:7] * r[:, None] * p


def undistort(d: np.ndarray, k1: np.ndarray, k2: np.ndarray, iters: int = 25) -> np.ndarray:
    """Invert  d = (1 + k1 |p|^2 + k2 |p|^4) p  for p (Newton on the radius; d, p are (N, 2))."""
    rd = np.sqrt(np.sum(d * d, axis=1))
    rho = rd.copy()
    for _ in range(iters):
        r2 = rho * rho
        g = rho * (1.0 + k1 * r2 + k2 * r2 * r2) - rd
        dg = 1.0 + 3.0 * k1 * r2 + 5.0 * k2 * r2 * r2
        step = g / np.where(np.abs(dg) > 1e-12, dg, 1.0)
        rho = rho - step
        if np.max(np.abs(step), initial=0.0) <= 1e-15 * max(np.max(rd, initial=0.0), 1.0):
            break
    scale = np.where(rd > 0.0, rho / np.where(rd > 0.0, rd, 1.0), 1.0)
    return d * scale[:, None]


@dataclass
class BalMapping:
    """What ``flat_to_bal`` needs to undo ``bal_to_flat``."""
    f0: float
    fixed_camera: int
    slot_of_camera: np.ndarray   # (Nc,) int32: adjustable slot, -1 for the fixed camera


def bal_to_flat(bal: BalProblem, f0: Optional[float] = None, fixed_camera: int = 0):
    """-> (FlatProblem, BalMapping); see the module docstring for the mapping."""
    nc = bal.n_cams
    if not 0 <= fixed_camera < nc:
        raise ValueError("fixed_camera out of range")
    f = bal.cameras[:, 6]
    if f0 is None:
        f0 = float(np.median(f))
    R = rodrigues(bal.cameras[:, 0:3])
    Rf = np.einsum("ij,njk->nik", _S, R)
    rvec = np.array([matrix_to_rvec(Rf[i]) for i in range(nc)]).reshape(nc, 3)
    tvec = bal.cameras[:, 3:6] @ _S.T
    slot = np.arange(nc, dtype=np.int32) - (np.arange(nc) > fixed_camera)
    slot[fixed_camera] = -1
    ci = bal.cam_index
    p = undistort(bal.uv / f[ci, None], bal.cameras[ci, 7], bal.cameras[ci, 8])
    uv = np.stack([f0 * p[:, 0], -f0 * p[:, 1]], axis=1)
    adj = slot >= 0
    fp = FlatProblem(f0, f0, 0.0, 0.0, np.ascontiguousarray(rvec[adj]), np.ascontiguousarray(tvec[adj]),
                     rvec[fixed_camera].copy(), tvec[fixed_camera].copy(), bal.points.copy(),
                     slot[ci].astype(np.int32), bal.pt_index.astype(np.int32), np.ascontiguousarray(uv), None)
    return fp, BalMapping(float(f0), int(fixed_camera), slot)


def flat_to_bal(bal: BalProblem, mapping: BalMapping, rvec: np.ndarray, tvec: np.ndarray, pts: np.ndarray) -> BalProblem:
    """BAL problem with the adjustable cameras' poses and the points replaced by solver output (flat frame)."""
    cams = bal.cameras.copy()
    rvec = np.asarray(rvec, dtype=np.float64).reshape(-1, 3)
    tvec = np.asarray(tvec, dtype=np.float64).reshape(-1, 3)
    for c in range(bal.n_cams):
        s = int(mapping.slot_of_camera[c])
        if s < 0:
            continue
        cams[c, 0:3] = matrix_to_rvec(_S @ rvec_to_matrix(rvec[s]))
        cams[c, 3:6] = _S @ tvec[s]
    return replace(bal, cameras=cams, points=np.asarray(pts, dtype=np.float64).reshape(-1, 3).copy())
