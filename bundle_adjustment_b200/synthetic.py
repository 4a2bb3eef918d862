"""Synthetic bundle-adjustment problems of the shapes BASELINE.json names.

Two generators:

* :func:`make_flat_problem` - vectorised, straight to the flat SoA layout the C-ABI takes
  (``ba_set_problem`` in ``include/ba_b200.h``); used for the BAL/Ladybug/Venice-shaped configs
  where building Python ``Map`` objects (50 M dict entries) is not feasible.
* :func:`make_window_spec` / :func:`build_map` - a small sliding-window map built through
  ``Map``/``Keyframe``/``MapPoint`` objects (the reference's ``src/map_structures.py`` classes or
  the mirror in ``bundle_adjustment_b200.map_structures``), the layout ``BundleAdjuster.run``
  consumes.

Camera model and pose convention are the reference's (``src/bundle_adjuster.py:67``):
``x_cam = R(rvec) X + t``, ``u = fx x/z + cx``, ``v = fy y/z + cy``; intrinsics default to
``src/main.py:36-40``.  Observation order is keyframe-major, as ``_gather_local_data``
(``src/bundle_adjuster.py:195-218``) produces; camera index -1 marks the fixed (oldest) keyframe.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np

# intrinsics of the reference's demo video, src/main.py:36-40
FX, FY, CX, CY = 912.7820434570312, 913.0294189453125, 650.2929077148438, 362.7241516113281
CAMERA_MATRIX = np.array([[FX, 0.0, CX], [0.0, FY, CY], [0.0, 0.0, 1.0]])

# (cameras incl. the fixed one, points, observations) for BASELINE.json configs 2..5
SHAPES = {
    "bal50": (50, 20_000, 100_000),
    "ladybug": (1723, 156_000, 680_000),
    "venice": (1778, 993_000, 5_000_000),
    "shard50m": (10_000, 10_000_000, 50_000_000),
}
SEEDS = {"window": 1001, "bal50": 1002, "ladybug": 1003, "venice": 1004, "shard50m": 1005}


def rodrigues(rvec: np.ndarray) -> np.ndarray:
    """Rotation vectors (..., 3) -> matrices (..., 3, 3); same formula as ``cv2.Rodrigues``
    (R = cos(th) I + (1-cos(th)) k k^T + sin(th) [k]x, identity below th = DBL_EPSILON)."""
    r = np.asarray(rvec, dtype=np.float64)
    th = np.linalg.norm(r, axis=-1)
    small = th < np.finfo(np.float64).eps
    k = r / np.where(small, 1.0, th)[..., None]
    c, s = np.cos(th), np.sin(th)
    kx, ky, kz = k[..., 0], k[..., 1], k[..., 2]
    zero = np.zeros_like(kx)
    K = np.stack([np.stack([zero, -kz, ky], -1), np.stack([kz, zero, -kx], -1), np.stack([-ky, kx, zero], -1)], -2)
    eye = np.broadcast_to(np.eye(3), K.shape)
    R = c[..., None, None] * eye + (1.0 - c)[..., None, None] * (k[..., :, None] * k[..., None, :]) \
        + s[..., None, None] * K
    return np.where(small[..., None, None], eye, R)


def project(rvec, tvec, pts, fx=FX, fy=FY, cx=CX, cy=CY):
    """Pinhole projection of pts (N,3) by per-row poses (N,3)/(N,3) -> (N,2)."""
    R = rodrigues(rvec)
    xc = np.einsum("nij,nj->ni", R, pts) + tvec
    return np.stack([fx * xc[:, 0] / xc[:, 2] + cx, fy * xc[:, 1] / xc[:, 2] + cy], -1)


@dataclass
class FlatProblem:
    """Flat SoA bundle-adjustment problem; field meaning = ``ba_set_problem`` arguments."""
    fx: float
    fy: float
    cx: float
    cy: float
    rvec: np.ndarray        # (Nc,3) adjustable cameras, initial guess
    tvec: np.ndarray        # (Nc,3)
    fixed_rvec: np.ndarray  # (3,)
    fixed_tvec: np.ndarray  # (3,)
    pts: np.ndarray         # (Np,3) initial guess
    obs_cam: np.ndarray     # (No,) int32, -1 = fixed keyframe
    obs_pt: np.ndarray      # (No,) int32
    obs_uv: np.ndarray      # (No,2) float64 (float32-valued)
    truth: Optional[Dict[str, np.ndarray]] = None

    @property
    def n_cams(self) -> int:
        return int(self.rvec.shape[0])

    @property
    def n_pts(self) -> int:
        return int(self.pts.shape[0])

    @property
    def n_obs(self) -> int:
        return int(self.obs_cam.shape[0])

    def x0(self) -> np.ndarray:
        """Reference parameter vector [all rvecs | all tvecs | all points] (src/bundle_adjuster.py:162)."""
        return np.concatenate([self.rvec.ravel(), self.tvec.ravel(), self.pts.ravel()])

    def point_slice(self, lo: int, hi: int) -> "FlatProblem":
        """Sub-problem holding points [lo, hi) and their observations (all cameras kept);
        the unit of multi-GPU sharding."""
        keep = (self.obs_pt >= lo) & (self.obs_pt < hi)
        return FlatProblem(self.fx, self.fy, self.cx, self.cy, self.rvec, self.tvec, self.fixed_rvec,
                           self.fixed_tvec, np.ascontiguousarray(self.pts[lo:hi]),
                           np.ascontiguousarray(self.obs_cam[keep]),
                           np.ascontiguousarray(self.obs_pt[keep] - lo).astype(np.int32),
                           np.ascontiguousarray(self.obs_uv[keep]), None)


def _track_lengths(rng, n_pts, n_obs, max_len):
    """Per-point track lengths: 2 + geometric tail, mean n_obs/n_pts, summing exactly to n_obs."""
    mean = n_obs / n_pts
    if mean < 2.0:
        raise ValueError("need at least 2 observations per point on average")
    if mean == 2.0:
        L = np.full(n_pts, 2, dtype=np.int64)
    else:
        p = 1.0 / (mean - 1.0)
        L = 1 + rng.geometric(p, size=n_pts).astype(np.int64)   # geometric >= 1 -> L >= 2
    L = np.minimum(L, max_len)
    diff = n_obs - int(L.sum())
    while diff != 0:
        if diff > 0:
            cand = np.flatnonzero(L < max_len)
            pick = rng.choice(cand, size=min(diff, cand.size), replace=False)
            L[pick] += 1
        else:
            cand = np.flatnonzero(L > 2)
            pick = rng.choice(cand, size=min(-diff, cand.size), replace=False)
            L[pick] -= 1
        diff = n_obs - int(L.sum())
    return L


def make_flat_problem(n_cams: int, n_pts: int, n_obs: int, seed: int, *, outlier_frac: float = 0.05,
                      pixel_sigma: float = 0.5, keep_truth: bool = True, well_posed: bool = False) -> FlatProblem:
    """Cameras on a smooth trajectory along +x looking at a slab of points (depth 8-20).

    ``n_cams`` counts the fixed camera (camera 0), so ``n_cams - 1`` are adjustable.
    Noise model: pixel noise N(0, 0.5^2), 5 % outliers with uniform +-50 px error,
    uv rounded through float32; initial guess = truth + N(0, 0.003^2) rad on rvec, N(0, 0.02^2) on t,
    N(0, 0.05^2) on points; camera 0 fixed at truth.

    ``well_posed``: the same generator, but the outliers (still ``outlier_frac`` of all observations) are drawn so that
    every point keeps at least TWO inlier observations.  Under the Huber loss a two-view point with a gross outlier can
    have its optimum at infinite depth (the outlier asks for a negative disparity): the objective then has an infimum
    but no minimiser, no solver reaches a stationary point, and "converged cost" is undefined.  The converged-cost
    parity tests therefore use ``well_posed=True``; everything else (and the bench) keeps the default.
    """
    rng = np.random.default_rng(seed)
    dx = 0.42
    i = np.arange(n_cams, dtype=np.float64)
    centers = np.stack([i * dx, rng.normal(0, 0.05, n_cams), rng.normal(0, 0.05, n_cams)], -1)
    yaw = 0.08 * np.sin(2.0 * np.pi * 2.0 * i / max(n_cams, 2))
    rvec_true = rng.normal(0, 0.05, (n_cams, 3))
    rvec_true[:, 1] += yaw
    R_true = rodrigues(rvec_true)
    tvec_true = -np.einsum("nij,nj->ni", R_true, centers)

    Z = rng.uniform(8.0, 20.0, n_pts)
    half = 0.4 * Z                       # half-width (world x) of the strip of cameras that see the point
    x_lo, x_hi = -2.0, (n_cams - 1) * dx + 2.0
    X = rng.uniform(x_lo, x_hi, n_pts)
    Y = rng.uniform(-2.0, 2.0, n_pts)
    pts_true = np.stack([X, Y, Z], -1)

    lo = np.clip(np.ceil((X - half) / dx), 0, n_cams - 1).astype(np.int64)
    hi = np.clip(np.floor((X + half) / dx), 0, n_cams - 1).astype(np.int64) + 1
    W = hi - lo
    short = W < 2                        # only possible at the very ends of tiny trajectories
    lo = np.where(short, np.clip(lo - 1, 0, n_cams - 2), lo)
    W = np.where(short, 2, W)

    L = np.minimum(_track_lengths(rng, n_pts, n_obs, max_len=int(W.max())), W)
    diff = n_obs - int(L.sum())
    while diff > 0:                      # clipping by the window removed a few: give them back elsewhere
        cand = np.flatnonzero(L < W)
        pick = rng.choice(cand, size=min(diff, cand.size), replace=False)
        L[pick] += 1
        diff = n_obs - int(L.sum())

    start = np.cumsum(L) - L
    obs_pt = np.repeat(np.arange(n_pts, dtype=np.int64), L)
    j = np.arange(n_obs, dtype=np.int64) - np.repeat(start, L)
    Lr, Wr = np.repeat(L, L), np.repeat(W, L)
    # stratified jitter: one camera in each of L disjoint integer sub-strips of the window -> distinct cameras
    a, b = (j * Wr) // Lr, ((j + 1) * Wr) // Lr
    cam = np.repeat(lo, L) + a + np.minimum(np.floor(rng.random(n_obs) * (b - a)).astype(np.int64), b - a - 1)

    uv = project(rvec_true[cam], tvec_true[cam], pts_true[obs_pt])
    uv += rng.normal(0, pixel_sigma, uv.shape)
    if well_posed:
        # rank the observations of each point by a random key; the two smallest keys of a point stay inliers
        key = rng.random(n_obs)
        rank = np.empty(n_obs, dtype=np.int64)
        o = np.lexsort((key, obs_pt))
        rank[o] = np.arange(n_obs) - np.repeat(start, L)
        eligible = rank >= 2
        frac = min(1.0, outlier_frac * n_obs / max(int(eligible.sum()), 1))
        out = eligible & (rng.random(n_obs) < frac)
    else:
        out = rng.random(n_obs) < outlier_frac
    uv[out] += rng.uniform(-50.0, 50.0, (int(out.sum()), 2))
    uv = uv.astype(np.float32).astype(np.float64)

    order = np.argsort(cam, kind="stable")          # keyframe-major, points ascending inside a keyframe
    cam, obs_pt, uv = cam[order], obs_pt[order], uv[order]

    rvec0 = rvec_true + rng.normal(0, 0.003, rvec_true.shape)
    tvec0 = tvec_true + rng.normal(0, 0.02, tvec_true.shape)
    pts0 = pts_true + rng.normal(0, 0.05, pts_true.shape)
    truth = None
    if keep_truth:
        truth = {"rvec": rvec_true[1:].copy(), "tvec": tvec_true[1:].copy(), "pts": pts_true}
    return FlatProblem(FX, FY, CX, CY,
                       np.ascontiguousarray(rvec0[1:]), np.ascontiguousarray(tvec0[1:]),
                       rvec_true[0].copy(), tvec_true[0].copy(), np.ascontiguousarray(pts0),
                       (cam - 1).astype(np.int32), obs_pt.astype(np.int32), np.ascontiguousarray(uv), truth)


def make_named_problem(name: str, scale: float = 1.0, keep_truth: bool = True, well_posed: bool = False) -> FlatProblem:
    """One of the BASELINE.json shapes (``SHAPES``), optionally scaled down in points/observations."""
    n_cams, n_pts, n_obs = SHAPES[name]
    if scale != 1.0:
        n_pts = max(int(round(n_pts * scale)), 16)
        n_obs = max(int(round(n_obs * scale)), 2 * n_pts)
        n_cams = max(int(round(n_cams * max(scale, 0.05))), 4)
    return make_flat_problem(n_cams, n_pts, n_obs, SEEDS[name], keep_truth=keep_truth, well_posed=well_posed)


# ----------------------------------------------------------------------------------------------
# window-shaped maps through Map / Keyframe / MapPoint objects
# ----------------------------------------------------------------------------------------------

def make_window_spec(n_kfs: int = 10, n_pts: int = 3000, seed: int = SEEDS["window"], *, mean_track: float = 4.0,
                     extra_keypoints: int = 50, duplicates: int = 0, dangling: int = 0,
                     kf_id_stride: int = 1, outlier_frac: float = 0.05) -> Dict[str, np.ndarray]:
    """Plain-array description of a sliding-window map (so the same map can be built with the
    reference's classes and with the mirror classes, and stored as a golden fixture).

    ``duplicates``: number of (keyframe, map point) pairs listed twice in a keyframe with two different
    keypoints (reference quirk: the residual rows are both kept but share the LAST keypoint,
    ``src/bundle_adjuster.py:216``).  ``dangling``: observations whose map point id is not in the map
    (filtered by ``:213``).  ``kf_id_stride`` > 1 gives non-contiguous keyframe ids.
    """
    n_obs = int(round(n_pts * mean_track))
    fp = make_flat_problem(n_kfs, n_pts, n_obs, seed, outlier_frac=outlier_frac)
    rng = np.random.default_rng(seed + 7919)
    cam = fp.obs_cam.astype(np.int64) + 1                      # 0 .. n_kfs-1
    rvec = np.concatenate([fp.fixed_rvec[None], fp.rvec])
    tvec = np.concatenate([fp.fixed_tvec[None], fp.tvec])
    kf_ids = np.arange(n_kfs, dtype=np.int64) * kf_id_stride
    mp_ids = np.sort(rng.choice(np.arange(n_pts * 2), size=n_pts, replace=False)).astype(np.int64)

    kp_xy, kp_off, ob_mp, ob_kp, ob_off = [], [0], [], [], [0]
    for k in range(n_kfs):
        sel = np.flatnonzero(cam == k)
        n_real = sel.size
        xy = np.concatenate([fp.obs_uv[sel], rng.uniform([0, 0], [1280, 720], (extra_keypoints, 2))])
        perm = rng.permutation(xy.shape[0])                     # keypoint order is unrelated to map order
        inv = np.empty_like(perm)
        inv[perm] = np.arange(perm.size)
        xy = xy[perm].astype(np.float32)
        mp = mp_ids[fp.obs_pt[sel]]
        kp = inv[:n_real]
        if duplicates and n_real > 0:
            d = rng.choice(n_real, size=min(duplicates, n_real), replace=False)
            mp = np.concatenate([mp, mp[d]])
            kp = np.concatenate([kp, inv[n_real + rng.integers(0, extra_keypoints, d.size)]])
        if dangling:
            mp = np.concatenate([mp, mp_ids.max() + 1 + np.arange(dangling)])
            kp = np.concatenate([kp, rng.integers(0, xy.shape[0], dangling)])
        kp_xy.append(xy); kp_off.append(kp_off[-1] + xy.shape[0])
        ob_mp.append(mp); ob_kp.append(kp); ob_off.append(ob_off[-1] + mp.size)
    return {
        "camera_matrix": CAMERA_MATRIX.copy(),
        "kf_ids": kf_ids, "kf_R": rodrigues(rvec), "kf_t": tvec.reshape(n_kfs, 3, 1),
        "kp_xy": np.concatenate(kp_xy).astype(np.float32), "kp_off": np.array(kp_off, dtype=np.int64),
        "ob_mp": np.concatenate(ob_mp).astype(np.int64), "ob_kp": np.concatenate(ob_kp).astype(np.int64),
        "ob_off": np.array(ob_off, dtype=np.int64),
        "mp_ids": mp_ids, "mp_pos": fp.pts.reshape(n_pts, 3, 1).copy(),
    }


def build_map(spec: Dict[str, np.ndarray], MapCls, KeyframeCls, MapPointCls, keypoint_factory):
    """Instantiate ``spec`` with the given classes.  ``keypoint_factory(x, y)`` makes one keypoint
    object exposing ``.pt`` (``cv2.KeyPoint`` for the reference, ``KeyPoint.from_xy`` for the mirror)."""
    gmap = MapCls()
    for mp_id, pos in zip(spec["mp_ids"].tolist(), spec["mp_pos"]):
        gmap.add_map_point(MapPointCls(id=mp_id, position=pos.astype(np.float64).copy(), observations=[],
                                       color=np.zeros((3, 1))))
    for k, kf_id in enumerate(spec["kf_ids"].tolist()):
        a, b = int(spec["kp_off"][k]), int(spec["kp_off"][k + 1])
        kps = [keypoint_factory(float(x), float(y)) for x, y in spec["kp_xy"][a:b]]
        a, b = int(spec["ob_off"][k]), int(spec["ob_off"][k + 1])
        obs = list(zip(spec["ob_mp"][a:b].tolist(), spec["ob_kp"][a:b].tolist()))
        gmap.add_keyframe(KeyframeCls(id=kf_id, R=spec["kf_R"][k].copy(), t=spec["kf_t"][k].copy(), keypoints=kps,
                                      descriptors=None, observations=obs, img=None))
        for mp_id, kp_idx in obs:
            if mp_id in gmap.map_points:
                gmap.map_points[mp_id].observations.append((kf_id, kp_idx))
    return gmap
