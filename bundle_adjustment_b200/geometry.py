"""Host-side rotation helpers used when packing / unpacking poses at the ``BundleAdjuster`` boundary.

The reference converts between rotation matrices and rotation vectors with ``cv2.Rodrigues``
(``src/bundle_adjuster.py:157`` matrix -> vector when packing, ``:235`` vector -> matrix in
``_update_map``).  These are plain-numpy equivalents so the drop-in does not need OpenCV.
"""
from __future__ import annotations

import numpy as np

_EPS = float(np.finfo(np.float64).eps)


def rvec_to_matrix(rvec) -> np.ndarray:
    """Rotation vector (3,) -> 3x3 matrix (Rodrigues formula; identity below |r| = DBL_EPSILON)."""
    r = np.asarray(rvec, dtype=np.float64).reshape(3)
    th = float(np.sqrt(r @ r))
    if th < _EPS:
        return np.eye(3)
    k = r / th
    K = np.array([[0.0, -k[2], k[1]], [k[2], 0.0, -k[0]], [-k[1], k[0], 0.0]])
    c, s = np.cos(th), np.sin(th)
    return c * np.eye(3) + (1.0 - c) * np.outer(k, k) + s * K


def matrix_to_rvec(R) -> np.ndarray:
    """3x3 matrix -> rotation vector (3,).  The input is first projected onto SO(3) (polar factor via
    SVD), as OpenCV does, so slightly non-orthonormal pipeline poses are accepted."""
    M = np.asarray(R, dtype=np.float64).reshape(3, 3)
    U, _, Vt = np.linalg.svd(M)
    M = U @ Vt
    ax = np.array([M[2, 1] - M[1, 2], M[0, 2] - M[2, 0], M[1, 0] - M[0, 1]])
    s = 0.5 * float(np.sqrt(ax @ ax))
    c = min(max(0.5 * (M[0, 0] + M[1, 1] + M[2, 2] - 1.0), -1.0), 1.0)
    theta = float(np.arccos(c))
    if s >= 1e-5:
        return ax * (theta / (2.0 * s))
    if c > 0.0:
        return np.zeros(3)
    # angle ~ pi: the axis comes from the diagonal, signs from the off-diagonal terms
    v = np.sqrt(np.maximum(0.5 * (np.diag(M) + 1.0), 0.0))
    if M[0, 1] < 0.0:
        v[1] = -v[1]
    if M[0, 2] < 0.0:
        v[2] = -v[2]
    if abs(v[0]) < abs(v[1]) and abs(v[0]) < abs(v[2]) and ((M[1, 2] > 0.0) != (v[1] * v[2] > 0.0)):
        v[2] = -v[2]
    return v * (theta / float(np.sqrt(v @ v)))
