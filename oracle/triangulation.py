"""Two-view linear triangulation restated in numpy.  TEST INFRASTRUCTURE (checker for ``ba_triangulate``).

Reference: ``VisualOdometryPipeline._triangulate_points`` (``src/pipeline.py:315-336``); the arithmetic is
OpenCV's ``cv2.triangulatePoints`` (third-party, not under /root/reference, unpinned; 4.13.0 in this image),
whose published algorithm is the DLT: per correspondence the 4x4 matrix with rows
``x1 P1[2]-P1[0], y1 P1[2]-P1[1], x2 P2[2]-P2[0], y2 P2[2]-P2[1]`` and the right singular vector of the smallest
singular value.  Pinned against ``cv2.triangulatePoints`` itself and against the reference function run on
the desk frames (``tests/golden/pipeline_triangulation.npz``, ``oracle/gen_triangulation_golden.py``).
OpenCV returns float32 when the pixel arrays are float32 (as they are in the pipeline, ``src/pose_estimator.py:22-23``).
"""
from __future__ import annotations

import numpy as np


def triangulate_points(P1, P2, pts1, pts2) -> np.ndarray:
    """pts Nx2 -> 4xN homogeneous (unit columns, sign as numpy's SVD gives it)."""
    P1, P2 = np.asarray(P1, dtype=np.float64), np.asarray(P2, dtype=np.float64)
    pts1, pts2 = np.asarray(pts1, dtype=np.float64).reshape(-1, 2), np.asarray(pts2, dtype=np.float64).reshape(-1, 2)
    out = np.empty((4, pts1.shape[0]))
    for i in range(pts1.shape[0]):
        A = np.stack([pts1[i, 0] * P1[2] - P1[0], pts1[i, 1] * P1[2] - P1[1],
                      pts2[i, 0] * P2[2] - P2[0], pts2[i, 1] * P2[2] - P2[1]])
        out[:, i] = np.linalg.svd(A)[2][3]
    return out


def triangulate_two_view(K, R_rel, t_rel, pts1, pts2, hom=None):
    """The reference's post-processing (``:324-336``) of a 4xN homogeneous solution -> (3xN points, mask)."""
    K = np.asarray(K, dtype=np.float64)
    R_rel = np.asarray(R_rel, dtype=np.float64).reshape(3, 3)
    t_rel = np.asarray(t_rel, dtype=np.float64).reshape(3, 1)
    if hom is None:
        P1 = K @ np.hstack((np.eye(3), np.zeros((3, 1))))
        P2 = K @ np.hstack((R_rel, t_rel))
        hom = triangulate_points(P1, P2, pts1, pts2)
    X = hom / (hom[3] + 1e-6)
    in1 = X[2] > 0
    in2 = (R_rel @ X[:3] + t_rel)[2] > 0
    return X[:3], np.logical_and(in1, in2)
