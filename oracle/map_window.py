"""CPU statement of the device-resident map (``ba_map_*`` in include/ba_b200.h, kernels in csrc/ba_map.cuh).
TEST INFRASTRUCTURE ONLY (never shipped, never timed).

``MapStore`` keeps what the device keeps - the observation rows of every keyframe in append order, the positions of the map
points by id - and ``window`` builds from them the flat problem the reference's ``_gather_local_data``
(``/root/reference/src/bundle_adjuster.py:195-218``) + the row indexing of ``_cost_function`` (``:46-63``) define:

* rows of the window's keyframes, keyframe-major in window order, list order inside a keyframe (``:210-216``);
* a row whose map point is not in the map is dropped (``:213``);
* points numbered by the rank of their id among the window's ids (``sorted(...)`` at ``:218``);
* a (keyframe, map point) pair listed twice keeps both rows, both with the LAST keypoint (the dict at ``:216``).

It has the method names of ``FlatSolver``'s ``map_*`` so that it can stand in for the solver in CPU tests of the host logic,
and is compared row for row with the device's result in the GPU tests.
"""
from __future__ import annotations

import numpy as np


class MapStore:
    def __init__(self):
        self.counters = dict(resident_rows=0, rows_appended=0, points_set=0, windows=0, bytes_uploaded=0)   # cumulative
        self.map_clear()

    # -- the ba_map_* calls ---------------------------------------------------------------------------------
    def map_clear(self):
        self.row_kf = np.zeros(0, dtype=np.int64)
        self.row_mp = np.zeros(0, dtype=np.int64)
        self.row_uv = np.zeros((0, 2))
        self.xyz = {}                                   # id -> (3,) position; membership = "in the map"
        self.counters["resident_rows"] = 0
        self.flat = None

    def map_append_observations(self, kf_id, mp_ids, uv):
        mp_ids = np.asarray(mp_ids, dtype=np.int64).ravel()
        uv = np.asarray(uv, dtype=np.float64).reshape(-1, 2)
        assert uv.shape[0] == mp_ids.shape[0]
        self.row_kf = np.concatenate([self.row_kf, np.full(mp_ids.shape[0], int(kf_id), dtype=np.int64)])
        self.row_mp = np.concatenate([self.row_mp, mp_ids])
        self.row_uv = np.concatenate([self.row_uv, uv])
        self.counters["resident_rows"] = int(self.row_kf.shape[0])
        self.counters["rows_appended"] += int(mp_ids.shape[0])
        self.counters["bytes_uploaded"] += 20 * int(mp_ids.shape[0])

    def map_set_points(self, mp_ids, xyz):
        xyz = np.asarray(xyz, dtype=np.float64).reshape(-1, 3)
        for i, p in zip(np.asarray(mp_ids).ravel().tolist(), xyz):
            self.xyz[int(i)] = p.copy()
        self.counters["points_set"] += int(xyz.shape[0])
        self.counters["bytes_uploaded"] += 28 * int(xyz.shape[0])

    def map_forget_points(self, mp_ids):
        for i in np.asarray(mp_ids).ravel().tolist():
            self.xyz.pop(int(i), None)

    def map_window(self, fx, fy, cx, cy, kf_ids, rvec, tvec):
        kf_ids = [int(k) for k in np.asarray(kf_ids).ravel()]
        rvec, tvec = np.asarray(rvec, dtype=np.float64).reshape(len(kf_ids), 3), np.asarray(tvec, dtype=np.float64).reshape(len(kf_ids), 3)
        self.counters["windows"] += 1
        self.counters["bytes_uploaded"] += 52 * len(kf_ids)
        self.flat = None
        slot_of = {k: s for s, k in enumerate(kf_ids)}
        slot = np.array([slot_of.get(int(k), -1) for k in self.row_kf], dtype=np.int64)
        known = np.array([int(m) in self.xyz for m in self.row_mp], dtype=bool)
        sel = np.flatnonzero((slot >= 0) & known)
        if sel.size == 0:
            return 0, 0
        order = sel[np.argsort(slot[sel], kind="stable")]                       # keyframe-major, append order inside
        ids = np.unique(self.row_mp[order])
        obs_pt = np.searchsorted(ids, self.row_mp[order]).astype(np.int32)
        s = slot[order]
        last = {}
        for i, key in enumerate(zip(s.tolist(), obs_pt.tolist())):
            last[key] = i                                                        # later rows overwrite earlier ones
        src = np.array([last[key] for key in zip(s.tolist(), obs_pt.tolist())], dtype=np.int64)
        self.flat = dict(fx=fx, fy=fy, cx=cx, cy=cy, rvec=rvec[1:].copy(), tvec=tvec[1:].copy(), fixed_rvec=rvec[0].copy(),
                         fixed_tvec=tvec[0].copy(), pts=np.array([self.xyz[int(i)] for i in ids]).reshape(-1, 3),
                         obs_cam=(s - 1).astype(np.int32), obs_pt=obs_pt, obs_uv=self.row_uv[order][src], ids=ids.astype(np.int32))
        return int(ids.shape[0]), int(order.shape[0])

    def map_window_points(self):
        return self.flat["ids"].copy()

    def map_commit(self, pts=None):
        """The solver's current points -> resident positions (the stand-in is told them)."""
        pts = self.flat["pts"] if pts is None else np.asarray(pts, dtype=np.float64).reshape(-1, 3)
        for i, p in zip(self.flat["ids"].tolist(), pts):
            self.xyz[int(i)] = p.copy()

    def map_stats(self):
        return dict(self.counters)
