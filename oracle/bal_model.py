"""BAL ("Bundle Adjustment in the Large", Agarwal et al., ECCV 2010) file format and camera model, restated the
slow and obvious way.  TEST INFRASTRUCTURE: checker for ``bundle_adjustment_b200/bal_io.py``.

The reference has no BAL reader - parity is unpinned by the reference for this row; the format and the
projection are the published ones:  P = R X + t (Rodrigues rotation vector),  p = -P / P.z,
p' = f * (1 + k1 |p|^2 + k2 |p|^4) * p;  file = header, observation lines ``cam pt x y``, then 9 numbers per
camera and 3 per point, one number per line.
"""
from __future__ import annotations

import math

import numpy as np


def parse_bal_slow(path: str):
    """Token-by-token parser (pure Python) -> (cam_index, pt_index, uv, cameras (Nc,9), points (Np,3))."""
    with open(path) as f:
        tok = f.read().split()
    nc, npt, no = int(tok[0]), int(tok[1]), int(tok[2])
    k = 3
    ci, pi, uv = [], [], []
    for _ in range(no):
        ci.append(int(tok[k])); pi.append(int(tok[k + 1])); uv.append((float(tok[k + 2]), float(tok[k + 3])))
        k += 4
    cams = [[float(tok[k + 9 * c + j]) for j in range(9)] for c in range(nc)]
    k += 9 * nc
    pts = [[float(tok[k + 3 * p + j]) for j in range(3)] for p in range(npt)]
    assert k + 3 * npt == len(tok)
    return (np.array(ci, dtype=np.int64), np.array(pi, dtype=np.int64), np.array(uv, dtype=np.float64).reshape(-1, 2),
            np.array(cams, dtype=np.float64).reshape(-1, 9), np.array(pts, dtype=np.float64).reshape(-1, 3))


def project_one(cam, X):
    """One observation through the BAL camera (scalar arithmetic, angle-axis rotation of the point)."""
    r, t, f, k1, k2 = cam[0:3], cam[3:6], cam[6], cam[7], cam[8]
    th = math.sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2])
    if th > 0.0:
        k = [r[0] / th, r[1] / th, r[2] / th]
        c, s = math.cos(th), math.sin(th)
        kxX = [k[1] * X[2] - k[2] * X[1], k[2] * X[0] - k[0] * X[2], k[0] * X[1] - k[1] * X[0]]
        kdX = k[0] * X[0] + k[1] * X[1] + k[2] * X[2]
        P = [X[i] * c + kxX[i] * s + k[i] * kdX * (1.0 - c) + t[i] for i in range(3)]
    else:
        P = [X[i] + t[i] for i in range(3)]
    px, py = -P[0] / P[2], -P[1] / P[2]
    n2 = px * px + py * py
    d = 1.0 + k1 * n2 + k2 * n2 * n2
    return f * d * px, f * d * py


def residuals(cam_index, pt_index, uv, cameras, points) -> np.ndarray:
    """observed - projected, (No, 2), BAL image coordinates."""
    out = np.empty((len(cam_index), 2))
    for i in range(len(cam_index)):
        u, v = project_one(cameras[cam_index[i]], points[pt_index[i]])
        out[i] = (uv[i, 0] - u, uv[i, 1] - v)
    return out
