"""Two-view triangulation of new map points on the device.

Mirrors what the reference's pipeline does when a keyframe is added (``src/pipeline.py:315-336``,
``VisualOdometryPipeline._triangulate_points``): ``cv2.triangulatePoints`` on the inlier matches of the
last keyframe and the new one, division by the homogeneous coordinate (+1e-6), and the cheirality test
(z > 0 in both views).  The arithmetic runs in ``libba_b200.so`` (``ba_triangulate``: one thread per
correspondence, one-sided Jacobi SVD of the 4x4 DLT matrix); there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np

from . import _capi as api
from .solver import FlatSolver


class Triangulator:
    """Owns (or borrows) a device handle; stateless otherwise."""

    def __init__(self, camera_matrix, *, device: int = 0, solver: Optional[FlatSolver] = None):
        self.camera_matrix = camera_matrix
        self._solver = solver if solver is not None else FlatSolver(device)

    def _call(self, P1, P2, R, t, pts1, pts2, want_hom, want_xyz, want_valid):
        pts1 = api.f64(pts1).reshape(-1, 2)
        pts2 = api.f64(pts2).reshape(-1, 2)
        if pts1.shape != pts2.shape:
            raise ValueError("pts1 and pts2 must hold the same number of points")
        n = pts1.shape[0]
        hom = np.zeros((n, 4)) if want_hom else None
        xyz = np.zeros((n, 3)) if want_xyz else None
        valid = np.zeros(n, dtype=np.uint8) if want_valid else None
        P1, P2 = api.f64(P1).reshape(12), api.f64(P2).reshape(12)
        Rr = api.f64(R).reshape(9) if R is not None else None
        tr = api.f64(t).reshape(3) if t is not None else None
        null_d = C.POINTER(C.c_double)()
        api.check(api.lib().ba_triangulate(
            self._solver._h, api.dptr(P1), api.dptr(P2), api.dptr(Rr) if Rr is not None else null_d,
            api.dptr(tr) if tr is not None else null_d, n, api.dptr(pts1), api.dptr(pts2),
            api.dptr(hom) if want_hom else null_d, api.dptr(xyz) if want_xyz else null_d,
            valid.ctypes.data_as(C.POINTER(C.c_uint8)) if want_valid else C.POINTER(C.c_uint8)()))
        return hom, xyz, valid

    def triangulate_points(self, P1, P2, pts1, pts2) -> np.ndarray:
        """``cv2.triangulatePoints`` look-alike: 3x4 projection matrices, 2xN (or Nx2) pixel arrays -> 4xN
        homogeneous points (unit columns; the sign of a column is arbitrary, as OpenCV's is)."""
        pts1, pts2 = np.asarray(pts1), np.asarray(pts2)
        if pts1.ndim == 2 and pts1.shape[0] == 2 and pts1.shape[1] != 2:
            pts1, pts2 = pts1.T, pts2.T
        hom, _, _ = self._call(P1, P2, None, None, pts1, pts2, True, False, False)
        return hom.T.copy()

    def triangulate_two_view(self, R_rel, t_rel, pts1, pts2) -> Tuple[Optional[np.ndarray], Optional[np.ndarray]]:
        """Same contract and the same printed line as the reference's ``_triangulate_points(R_rel, t_rel, pts1, pts2)``
        (``src/pipeline.py:315-336``): Nx2 pixel arrays of the two views, first camera at the origin ->
        (3xM points in the first camera's frame that lie in front of both cameras, their indices);
        ``(None, None)`` for an empty input."""
        pts1 = np.asarray(pts1)
        if pts1.shape[0] == 0:
            return None, None
        K = np.asarray(self.camera_matrix, dtype=np.float64)
        R_rel = np.asarray(R_rel, dtype=np.float64).reshape(3, 3)
        t_rel = np.asarray(t_rel, dtype=np.float64).reshape(3, 1)
        P1 = K @ np.hstack((np.eye(3), np.zeros((3, 1))))
        P2 = K @ np.hstack((R_rel, t_rel))
        _, xyz, valid = self._call(P1, P2, R_rel, t_rel, pts1, np.asarray(pts2), False, True, True)
        mask = valid.astype(bool)
        print(f"    -> Triangulation: Kept {int(mask.sum())} of {mask.shape[0]} points.")
        return xyz[mask].T.copy(), np.where(mask)[0]
