"""Golden vectors for the triangulation row, from the UNMODIFIED reference and OpenCV.

    python -m oracle.gen_triangulation_golden      # needs /root/reference and cv2

TEST INFRASTRUCTURE.  Case "desk": the reference's own front end (ORB, brute-force matching, ``estimate_pose``)
on its desk photographs 11 -> 22 and 22 -> 33, then the reference's ``_triangulate_points`` verbatim (float32
pixels, as in the pipeline).  Case "synthetic": exact two-view geometry with noise, float64 pixels straight into
``cv2.triangulatePoints`` - includes points behind a camera so that the cheirality mask has both values.
Stored: inputs, OpenCV's homogeneous output, the reference's 3-D points and kept indices.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle.reference_loader import load_reference  # noqa: E402

K = np.array([[912.7820434570312, 0.0, 650.2929077148438], [0.0, 913.0294189453125, 362.7241516113281], [0.0, 0.0, 1.0]])


def main():
    import cv2
    load_reference()
    import features
    import pipeline
    import pose_estimator

    class Shell:                        # the only attribute _triangulate_points reads from self
        camera_matrix = K

    gold = {"camera_matrix": K, "cv2_version": np.array(cv2.__version__)}
    ext, matcher = features.ORBExtractor(n_features=4000), features.BruteForceMatcher()
    imgs = {n: cv2.cvtColor(cv2.imread(f"/root/reference/desk_images/image{n}.png"), cv2.COLOR_BGR2GRAY) for n in (11, 22, 33)}
    feats = {n: ext.extract(im) for n, im in imgs.items()}
    for name, (a, b) in {"desk0": (11, 22), "desk1": (22, 33)}.items():
        (kp1, d1), (kp2, d2) = feats[a], feats[b]
        matches = matcher.match(d1, d2)
        with contextlib.redirect_stdout(io.StringIO()):
            R, t, p1, p2, _ = pose_estimator.estimate_pose(matches, kp1, kp2, K)
        out = io.StringIO()
        with contextlib.redirect_stdout(out):
            X, idx = pipeline.VisualOdometryPipeline._triangulate_points(Shell(), R, t, p1, p2)
        P1 = K @ np.hstack((np.eye(3), np.zeros((3, 1))))
        P2 = K @ np.hstack((R, t))
        hom = cv2.triangulatePoints(P1, P2, p1.T, p2.T)
        gold.update({f"{name}_R": R, f"{name}_t": t, f"{name}_pts1": p1, f"{name}_pts2": p2, f"{name}_hom": hom,
                     f"{name}_X": X, f"{name}_idx": idx, f"{name}_line": np.array(out.getvalue().strip())})
        print(name, p1.dtype, hom.dtype, X.shape, out.getvalue().strip())

    rng = np.random.default_rng(7)
    n = 400
    rv = np.array([0.02, -0.15, 0.03])
    R = cv2.Rodrigues(rv)[0]
    t = np.array([[-0.8], [0.05], [0.1]])
    Xw = np.stack([rng.uniform(-3, 3, n), rng.uniform(-2, 2, n), rng.uniform(2, 12, n)])
    Xw[2, :40] = rng.uniform(-6, -1, 40)                      # behind both cameras
    def proj(P, X):
        x = P @ np.vstack([X, np.ones((1, X.shape[1]))])
        return (x[:2] / x[2]).T
    P1 = K @ np.hstack((np.eye(3), np.zeros((3, 1))))
    P2 = K @ np.hstack((R, t))
    p1 = proj(P1, Xw) + rng.normal(0, 0.3, (n, 2))
    p2 = proj(P2, Xw) + rng.normal(0, 0.3, (n, 2))
    hom = cv2.triangulatePoints(P1, P2, p1.T.copy(), p2.T.copy())
    with contextlib.redirect_stdout(io.StringIO()) as out:
        X, idx = pipeline.VisualOdometryPipeline._triangulate_points(Shell(), R, t, p1, p2)
    gold.update({"syn_R": R, "syn_t": t, "syn_pts1": p1, "syn_pts2": p2, "syn_hom": hom, "syn_X": X, "syn_idx": idx,
                 "syn_line": np.array(out.getvalue().strip())})
    print("syn", p1.dtype, hom.dtype, X.shape, out.getvalue().strip())
    path = os.path.join(ROOT, "tests", "golden", "pipeline_triangulation.npz")
    np.savez_compressed(path, **gold)
    print("->", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
