"""Drop-in replacement for the reference's ``BundleAdjuster`` (``src/bundle_adjuster.py:16-240``).

Same constructor and ``run(gmap)`` signature, same in-place mutation of ``gmap.keyframes[id].R/.t`` and
``gmap.map_points[id].position``, same skip / divergence behaviour and the same stdout lines (the
``LBA Complete`` line is parsed by the reference's ``src/analyze_log.py:42-45``).  The optimisation
itself - the reference's ``scipy.optimize.least_squares`` call at ``:170-174`` - runs on the GPU through
``libba_b200.so`` (Levenberg-Marquardt on the Schur complement, same Huber objective).  The private
helpers keep the reference's names and argument lists so they can be called side by side with the
reference's in tests.

Import as ``from bundle_adjustment_b200 import BundleAdjuster``; for the reference's flat
``from bundle_adjuster import BundleAdjuster`` (``src/pipeline.py:12``) put ``bundle_adjustment_b200/shim`` ahead of the
reference's ``src/`` on ``sys.path`` (``shim/bundle_adjuster.py`` re-exports this class).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Tuple

import numpy as np

from . import _capi as api
from .geometry import matrix_to_rvec, rvec_to_matrix
from .solver import FlatSolver


class _NotResident(Exception):
    """The map cannot be mirrored on the device (ids that are no small non-negative integers, an oversized window):
    this call flattens on the host instead."""


class _ResidentMap:
    """Host-side bookkeeping of the device-resident map (``ba_map_*``): what has been uploaded so far."""

    def __init__(self, mpd):
        self.mpd = mpd                      # the map_points dict the device copy mirrors
        self.kfs: Dict[int, tuple] = {}     # kf_id -> (observation list, keypoint list, rows uploaded)
        self.known: set = set()             # map point ids whose position is on the device
        self.last_key = None                # last key of mpd when it was scanned (dicts keep insertion order)


class BundleAdjuster:
    """Local bundle adjustment over the most recent window of keyframes, on a B200."""

    def __init__(self, camera_matrix, window_size=5, *, device: int = 0, max_iters: int = 50, ftol: float = 1e-5,
                 xtol: float = 1e-5, gtol: float = 1e-8, huber_delta: float = 1.0, robust_mode: str = "irls",
                 verbose: int = 0, lba_steps_dir: Optional[str] = None, solver_options: Optional[dict] = None,
                 resident=True):
        # public, mutable attributes exactly as in the reference (src/main.py:83-86 mutates window_size); the tolerances default
        # to the reference's call (xtol = ftol = 1e-5, max_nfev = 50: src/bundle_adjuster.py:172), gtol to SciPy's default
        self.camera_matrix = camera_matrix
        self.window_size = window_size
        self.device = device
        self.solver_options = dict(max_iters=max_iters, ftol=ftol, xtol=xtol, gtol=gtol, huber_delta=huber_delta,
                                   robust_mode=robust_mode, verbose=verbose)
        if solver_options:
            self.solver_options.update(solver_options)
        # the reference dumps a .pcd after every successful LBA (src/bundle_adjuster.py:187-193) through
        # Open3D, which is unavailable on this box; set lba_steps_dir to get the same file as ASCII PCD.
        self.lba_steps_dir = lba_steps_dir
        self.last_stats: Optional[dict] = None
        self._solver: Optional[FlatSolver] = None
        # incremental flattening: per keyframe, the arrays of its observation list as last seen
        self._kf_cache: Dict[int, tuple] = {}
        # device-resident map: observation rows and point positions stay on the GPU between run() calls, a call
        # uploads what the pipeline appended since the last one and the device builds the window.  True: positions the
        # device already holds are trusted (only BA rewrites them: the reference assigns `position` at creation and in
        # _update_map, nowhere else); "verify": they are also compared with the host's every call and re-uploaded where they
        # differ; False: flatten on the host every call (the round-1 path).
        self.resident = resident
        self._res: Optional[_ResidentMap] = None
        self.last_upload: Optional[dict] = None

    # ------------------------------------------------------------------------------------------
    def _get_solver(self) -> FlatSolver:
        if self._solver is None:
            self._solver = FlatSolver(self.device)   # raises without a GPU / built library: no CPU path
        return self._solver

    def _intrinsics(self):
        K = np.asarray(self.camera_matrix, dtype=np.float64)
        return K[0, 0], K[1, 1], K[0, 2], K[1, 2]

    @staticmethod
    def _index_observations(adjustable_kf_ids, map_point_ids, observations, keypoints_2d, fixed_kf_id=None):
        """Observation rows -> (obs_cam, obs_pt, obs_uv) in the reference's row order.  Rows whose map
        point or keyframe is unknown are dropped, as ``_cost_function`` drops them (``:54, 63``)."""
        kf_idx = {kf_id: i for i, kf_id in enumerate(adjustable_kf_ids)}
        mp_idx = {mp_id: i for i, mp_id in enumerate(map_point_ids)}
        cams, pts, uvs = [], [], []
        for kf_id, mp_id in observations:
            p = mp_idx.get(mp_id)
            if p is None:
                continue
            c = kf_idx.get(kf_id)
            if c is None:
                if fixed_kf_id is not None and kf_id != fixed_kf_id:
                    continue
                c = -1
            cams.append(c)
            pts.append(p)
            uvs.append(keypoints_2d[(kf_id, mp_id)])
        return (np.asarray(cams, dtype=np.int32), np.asarray(pts, dtype=np.int32),
                np.asarray(uvs, dtype=np.float64).reshape(-1, 2))

    # -- reference-named helpers -----------------------------------------------------------------
    def _gather_local_data(self, gmap, local_kf_ids: list):
        """Same contract as the reference's (``:195-218``): sorted unique map point ids of the window,
        ``(kf_id, mp_id)`` rows in keyframe-major order, and the ``(kf_id, mp_id) -> (u, v)`` dict (a pair
        listed twice keeps both rows but only the last keypoint)."""
        seen = set()
        observations: List[Tuple[int, int]] = []
        keypoints_2d: Dict[Tuple[int, int], Tuple[float, float]] = {}
        map_points = gmap.map_points
        for kf_id in local_kf_ids:
            kf = gmap.keyframes[kf_id]
            kps = kf.keypoints
            for mp_id, kp_idx in kf.observations:
                if mp_id in map_points:
                    seen.add(mp_id)
                    observations.append((kf_id, mp_id))
                    keypoints_2d[(kf_id, mp_id)] = kps[kp_idx].pt
        return sorted(seen), observations, keypoints_2d

    # -- incremental, vectorised flattening of a window (what run() uses) ------------------------------
    def _keyframe_arrays(self, kf_id, kf):
        """(map point ids, uv) of one keyframe's observation list, cached.  The pipeline only ever APPENDS to a
        keyframe's list (``src/pipeline.py:305-306``), so a cached prefix stays valid and only the tail is
        converted; any other change of the list (shorter, another list object) rebuilds the entry.  A (keyframe,
        map point) pair listed twice keeps both rows with the LAST keypoint, as the reference's dict does (``:216``)."""
        obs = kf.observations
        n = len(obs)
        ent = self._kf_cache.get(kf_id)
        if ent is not None and ent[0] is obs and ent[1] is kf.keypoints and ent[2] == n:
            return ent[5], ent[6]
        if ent is not None and ent[0] is obs and ent[1] is kf.keypoints and ent[2] < n:
            start, mp_raw, uv_raw = ent[2], ent[3], ent[4]
        else:
            start, mp_raw, uv_raw = 0, np.zeros(0, dtype=np.int64), np.zeros((0, 2))
        kps = kf.keypoints
        tail = obs[start:]
        mp_new = np.fromiter((o[0] for o in tail), dtype=np.int64, count=len(tail))
        uv_new = np.array([kps[o[1]].pt for o in tail], dtype=np.float64).reshape(-1, 2)
        mp_raw = np.concatenate([mp_raw, mp_new])
        uv_raw = np.concatenate([uv_raw, uv_new])
        # duplicates of a map point inside this keyframe: every row takes the uv of the last one
        uniq, inv = np.unique(mp_raw, return_inverse=True)
        uv_eff = uv_raw
        if uniq.shape[0] != mp_raw.shape[0]:
            last = np.zeros(uniq.shape[0], dtype=np.int64)
            last[inv] = np.arange(mp_raw.shape[0])          # later rows overwrite earlier ones
            uv_eff = uv_raw[last[inv]]
        self._kf_cache[kf_id] = (obs, kps, n, mp_raw, uv_raw, mp_raw, uv_eff)
        return mp_raw, uv_eff

    def _gather_flat(self, gmap, local_kf_ids: list, fixed_kf_id):
        """Vectorised equivalent of ``_gather_local_data`` + ``_index_observations``:
        -> (sorted map point ids, obs_cam, obs_pt, obs_uv) with identical contents and row order."""
        mps, uvs, cams = [], [], []
        slot = {kf_id: i - 1 for i, kf_id in enumerate(local_kf_ids)}      # fixed keyframe (first) -> -1
        for kf_id in local_kf_ids:
            mp, uv = self._keyframe_arrays(kf_id, gmap.keyframes[kf_id])
            mps.append(mp); uvs.append(uv)
            cams.append(np.full(mp.shape[0], -1 if kf_id == fixed_kf_id else slot[kf_id], dtype=np.int32))
        mp = np.concatenate(mps) if mps else np.zeros(0, dtype=np.int64)
        uv = np.concatenate(uvs) if uvs else np.zeros((0, 2))
        cam = np.concatenate(cams) if cams else np.zeros(0, dtype=np.int32)
        ids = np.unique(mp)
        # `mp_id in gmap.map_points` (:213), once per distinct id
        mpd = gmap.map_points
        known = np.fromiter((m in mpd for m in ids.tolist()), dtype=bool, count=ids.shape[0])
        if not known.all():
            keep = known[np.searchsorted(ids, mp)]
            mp, uv, cam = mp[keep], uv[keep], cam[keep]
            ids = ids[known]
        obs_pt = np.searchsorted(ids, mp).astype(np.int32)
        return ids.tolist(), cam, obs_pt, np.ascontiguousarray(uv)

    # -- device-resident window (what run() uses when self.resident) ---------------------------------------
    @staticmethod
    def _small_ids(ids) -> np.ndarray:
        a = np.asarray(ids)
        if a.size and (a.dtype.kind not in "iu" or int(a.min()) < 0 or int(a.max()) >= api.MAP_MAX_ID):
            raise _NotResident
        return a.astype(np.int32)

    def _window_resident(self, gmap, local_kf_ids: list):
        """Upload what is new since the last call and let the device build the window's flat problem (``ba_map_window``):
        -> sorted map point ids of the window, or None when it holds no observation of a known point.  Equivalent to
        ``_gather_flat`` + ``set_problem`` (same rows, same order, same point numbering)."""
        solver = self._get_solver()
        mpd = gmap.map_points
        self._small_ids(local_kf_ids)
        if self._res is None or self._res.mpd is not mpd:
            solver.map_clear()
            self._res = _ResidentMap(mpd)
        for attempt in (0, 1):
            st, ok = self._res, True
            for kf_id in local_kf_ids:
                kf = gmap.keyframes[kf_id]
                obs, kps = kf.observations, kf.keypoints
                n, ent = len(obs), st.kfs.get(kf_id)
                if ent is None:
                    start = 0
                elif ent[0] is obs and ent[1] is kps and ent[2] <= n:
                    start = ent[2]                      # the pipeline only appends (src/pipeline.py:172-173, 306-307)
                else:
                    ok = False                          # another list, or a shorter one: rebuild the resident copy
                    break
                if start < n:
                    tail = obs[start:]
                    mp = self._small_ids(np.fromiter((o[0] for o in tail), dtype=np.int64, count=len(tail)))
                    uv = np.array([kps[o[1]].pt for o in tail], dtype=np.float64).reshape(-1, 2)
                    solver.map_append_observations(kf_id, mp, uv)
                st.kfs[kf_id] = (obs, kps, n)
            if ok:
                break
            solver.map_clear()
            self._res = _ResidentMap(mpd)
        st = self._res
        # new map points: the tail of the dict (insertion order); anything else that changed triggers a full comparison
        known = st.known
        last_key = next(reversed(mpd)) if mpd else None
        if len(mpd) != len(known) or last_key != st.last_key:
            new = []
            for mp_id in reversed(mpd):
                if mp_id in known:
                    break
                new.append(mp_id)
            if len(known) + len(new) != len(mpd):
                gone = [i for i in known if i not in mpd]
                if gone:
                    solver.map_forget_points(np.asarray(gone, dtype=np.int32))
                    known.difference_update(gone)
                new = [i for i in mpd if i not in known]
            if new:
                ids = self._small_ids(new)
                xyz = np.array([np.asarray(mpd[i].position, dtype=np.float64).ravel() for i in new]).reshape(-1, 3)
                solver.map_set_points(ids, xyz)
                known.update(new)
            st.last_key = last_key
        rvecs = np.array([matrix_to_rvec(gmap.keyframes[i].R) for i in local_kf_ids]).reshape(-1, 3)
        tvecs = np.array([np.asarray(gmap.keyframes[i].t, dtype=np.float64).ravel() for i in local_kf_ids]).reshape(-1, 3)
        fx, fy, cx, cy = self._intrinsics()
        for attempt in (0, 1, 2):
            try:
                n_pts, n_obs = solver.map_window(fx, fy, cx, cy, local_kf_ids, rvecs, tvecs)
            except api.BaError as e:
                if "pair table" in str(e):
                    raise _NotResident from e
                raise
            if n_obs == 0:
                return None
            ids = solver.map_window_points().tolist()
            # `mp_id in gmap.map_points` (:213) for ids that were removed behind our back
            missing = [i for i in ids if i not in mpd]
            if missing:
                solver.map_forget_points(np.asarray(missing, dtype=np.int32))
                known.difference_update(missing)
                continue
            if self.resident == "verify" and attempt < 2:
                host = np.array([np.asarray(mpd[i].position, dtype=np.float64).ravel() for i in ids]).reshape(-1, 3)
                dev = solver.get_params()[2]
                changed = np.flatnonzero((host != dev).any(axis=1))
                if changed.size:
                    solver.map_set_points(np.asarray(ids, dtype=np.int32)[changed], host[changed])
                    continue
            return ids
        raise _NotResident

    def _cost_function(self, params, fixed_kf_pose, fixed_kf_id, adjustable_kf_ids, map_point_ids, observations,
                       keypoints_2d):
        """Re-projection residuals (observed - projected, x then y per row) for the parameter vector
        ``[rvecs | tvecs | points]``; same signature as the reference's (``:24-72``), evaluated by the
        device residual kernel."""
        nc, npt = len(adjustable_kf_ids), len(map_point_ids)
        params = np.asarray(params, dtype=np.float64)
        fixed_R, fixed_t = fixed_kf_pose
        obs_cam, obs_pt, obs_uv = self._index_observations(adjustable_kf_ids, map_point_ids, observations,
                                                           keypoints_2d, fixed_kf_id)
        if obs_cam.size == 0:
            return np.array([])
        fx, fy, cx, cy = self._intrinsics()
        s = self._get_solver()
        s.set_problem(fx, fy, cx, cy, params[:3 * nc], params[3 * nc:6 * nc], matrix_to_rvec(fixed_R),
                      np.asarray(fixed_t, dtype=np.float64).ravel(), params[6 * nc:6 * nc + 3 * npt],
                      obs_cam, obs_pt, obs_uv)
        return s.residuals()

    def _prepare_sparsity_matrix(self, num_adj_kfs, num_mps, adj_kf_ids, mp_ids, observations):
        """0/1 Jacobian structure, shape (2 * n_obs, 6 * n_kfs + 3 * n_mps); same signature as the
        reference's (``:74-120``).  Built by the device (``ba_sparsity_csr``) and returned as a SciPy CSR
        matrix whose ``indptr``/``indices`` equal ``reference_matrix.tocsr()``'s.  The GPU solver does not
        need it; it exists for callers and tests that do."""
        from scipy.sparse import csr_matrix
        kp = {o: (0.0, 0.0) for o in observations}
        obs_cam, obs_pt, obs_uv = self._index_observations(adj_kf_ids, mp_ids, observations, kp, None)
        if obs_cam.size != len(observations):
            raise ValueError("observation of a map point outside mp_ids")      # the reference would leave empty rows
        s = self._get_solver()
        s.set_problem(1.0, 1.0, 0.0, 0.0, np.zeros((num_adj_kfs, 3)), np.zeros((num_adj_kfs, 3)), np.zeros(3), np.zeros(3),
                      np.tile([0.0, 0.0, 1.0], (num_mps, 1)), obs_cam, obs_pt, obs_uv)
        indptr, indices, shape = s.sparsity_csr()
        return csr_matrix((np.ones(indices.shape[0], dtype=np.int64), indices.astype(np.int32), indptr.astype(np.int32)),
                          shape=shape)

    def _update_map(self, gmap, optimized_params, adjustable_kf_ids, local_map_point_ids):
        """Write the optimised vector back (reference ``:220-240``): R from the rotation vector, t as 3x1,
        positions as 3x1."""
        nc = len(adjustable_kf_ids)
        x = np.asarray(optimized_params, dtype=np.float64)
        rvecs = x[:3 * nc].reshape(nc, 3)
        tvecs = x[3 * nc:6 * nc].reshape(nc, 3)
        pts = x[6 * nc:].reshape(len(local_map_point_ids), 3)
        for i, kf_id in enumerate(adjustable_kf_ids):
            kf = gmap.keyframes[kf_id]
            kf.R = rvec_to_matrix(rvecs[i])
            kf.t = tvecs[i].reshape(3, 1).copy()
        # one (n, 3, 1) block, every position a view of its row: the reference hands out views of res.x as well (:237-240)
        block = np.ascontiguousarray(pts).reshape(-1, 3, 1).copy()
        mpd = gmap.map_points
        for i, mp_id in enumerate(local_map_point_ids):
            mpd[mp_id].position = block[i]

    # -- the entry point ---------------------------------------------------------------------------
    def run(self, gmap):
        """Run LBA on the most recent window of keyframes (reference ``run``, ``:122-193``)."""
        print("    --- Running Local Bundle Adjustment ---")
        self.last_stats = None

        all_kf_ids = sorted(gmap.keyframes.keys())
        if len(all_kf_ids) < self.window_size:
            print("    -> LBA Skipped: Not enough keyframes.")
            return

        local_kf_ids = all_kf_ids[-(self.window_size + 1): -1]      # the newest keyframe is not optimised
        fixed_kf_id = local_kf_ids[0]
        adjustable_kf_ids = local_kf_ids[1:]
        if not adjustable_kf_ids:
            print("    -> LBA Skipped: No adjustable keyframes.")
            return

        # the third skip needs no device (and is decided before one is asked for, as in the reference): is there ANY
        # observation of a known map point in the window?  Stops at the first one.
        mpd = gmap.map_points
        if not any(o[0] in mpd for kf_id in local_kf_ids for o in gmap.keyframes[kf_id].observations):
            print("    -> LBA Skipped: No points in the local window.")
            return

        solver = self._get_solver()
        on_device = False
        self.last_upload = None
        if self.resident:
            before = solver.map_stats()
            try:
                local_map_point_ids = self._window_resident(gmap, local_kf_ids)
                on_device = True
                after = solver.map_stats()
                self.last_upload = {k: after[k] - before[k] for k in ("rows_appended", "points_set", "bytes_uploaded")}
                self.last_upload["resident_rows"] = after["resident_rows"]
            except _NotResident:
                self._res = None
        if on_device:
            if local_map_point_ids is None:
                print("    -> LBA Skipped: No points in the local window.")
                return
        else:
            local_map_point_ids, obs_cam, obs_pt, obs_uv = self._gather_flat(gmap, local_kf_ids, fixed_kf_id)
            if not local_map_point_ids:
                print("    -> LBA Skipped: No points in the local window.")
                return
            for stale in [k for k in self._kf_cache if k not in gmap.keyframes]:
                del self._kf_cache[stale]

            rvecs = np.array([matrix_to_rvec(gmap.keyframes[i].R) for i in adjustable_kf_ids]).reshape(-1, 3)
            tvecs = np.array([np.asarray(gmap.keyframes[i].t, dtype=np.float64).ravel() for i in adjustable_kf_ids]).reshape(-1, 3)
            mpd = gmap.map_points
            try:    # positions are (3, 1) float64 arrays in the pipeline: one concatenate instead of a conversion per point
                pts = np.concatenate([mpd[i].position for i in local_map_point_ids], dtype=np.float64).reshape(-1, 3)
                if pts.shape[0] != len(local_map_point_ids):
                    raise ValueError
            except (ValueError, TypeError):
                pts = np.array([np.asarray(mpd[i].position, dtype=np.float64).ravel() for i in local_map_point_ids]).reshape(-1, 3)
            fixed_kf = gmap.keyframes[fixed_kf_id]

            fx, fy, cx, cy = self._intrinsics()
            solver.set_problem(fx, fy, cx, cy, rvecs, tvecs, matrix_to_rvec(fixed_kf.R),
                               np.asarray(fixed_kf.t, dtype=np.float64).ravel(), pts, obs_cam, obs_pt, obs_uv)
        stats = solver.solve(**self.solver_options)
        self.last_stats = stats

        # the reference reports and guards on the RAW sum of squares, not the Huber cost (:165, :176-179)
        initial_cost, final_cost = stats["initial_raw_cost"], stats["final_raw_cost"]
        if final_cost >= initial_cost:
            print(f"    -> LBA Diverged! Cost increased from {initial_cost:.2f} to {final_cost:.2f}. Discarding results.")
            return

        self._update_map(gmap, solver.get_x(), adjustable_kf_ids, local_map_point_ids)
        if on_device:
            solver.map_commit()      # the resident positions follow the map (a discarded solve leaves them alone)

        improvement = 100.0 * (initial_cost - final_cost) / (initial_cost + 1e-8)
        print(f"    -> LBA Complete. Initial Cost: {initial_cost:.2f}, Final Cost: {final_cost:.2f}, "
              f"Improvement: {improvement:.2f}%")

        if self.lba_steps_dir is not None:
            path = os.path.join(self.lba_steps_dir, f"map_after_lba_kf_{fixed_kf_id}.pcd")
            if _write_pcd(path, gmap):
                print(f"    -> Saved intermediate map to {path}")


def _write_pcd(path: str, gmap) -> bool:
    """ASCII .pcd of all map points (what the reference saves through Open3D at ``:187-193``)."""
    pos = [np.asarray(p.position, dtype=np.float64).reshape(3) for p in gmap.map_points.values()]
    if not pos:
        return False
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        f.write("# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z\nSIZE 8 8 8\nTYPE F F F\n"
                f"COUNT 1 1 1\nWIDTH {len(pos)}\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS {len(pos)}\nDATA ascii\n")
        for p in pos:
            f.write(f"{p[0]:.17g} {p[1]:.17g} {p[2]:.17g}\n")
    return True
