"""Thin Python owner of one ``ba_handle`` (one GPU, one CUDA stream).

``FlatSolver`` works on flat arrays in the reference's own layout (the arguments of
``ba_set_problem``); ``BundleAdjuster`` (``bundle_adjuster.py``) sits on top of it for ``Map`` objects.
All arithmetic happens in ``libba_b200.so``; nothing here computes residuals or steps on the host.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Tuple

import numpy as np

from . import _capi as api


def make_options(**kw) -> api.Options:
    """Defaults from ``ba_default_options`` overridden by keyword (names = ``ba_options`` fields;
    ``robust_mode`` / ``precond`` also accept their string names)."""
    o = api.Options()
    api.lib().ba_default_options(C.byref(o))
    for k, v in kw.items():
        if v is None:
            continue
        if k == "robust_mode" and isinstance(v, str):
            v = api.ROBUST[v]
        if k == "precond" and isinstance(v, str):
            v = api.PRECOND[v]
        if k == "schur_path" and isinstance(v, str):
            v = api.SCHUR_PATH[v]
        if not hasattr(o, k) or k.startswith("reserved"):
            raise TypeError(f"unknown solver option {k!r}")
        setattr(o, k, v)
    return o


class FlatSolver:
    def __init__(self, device: int = 0, *, rank: int = 0, nranks: int = 1, nccl_id: Optional[bytes] = None):
        self._h = C.c_void_p()
        self._shape: Optional[Tuple[int, int, int]] = None
        L = api.lib()
        if nranks > 1:
            if nccl_id is None or len(nccl_id) != 128:
                raise ValueError("distributed solver needs the 128-byte NCCL id from nccl_unique_id()")
            api.check(L.ba_create_distributed(device, rank, nranks, nccl_id, C.byref(self._h)))
        else:
            api.check(L.ba_create(device, C.byref(self._h)))
        self.rank, self.nranks = rank, nranks

    @staticmethod
    def nccl_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        api.check(api.lib().ba_nccl_unique_id(buf))
        return buf.raw

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            api.lib().ba_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- problem ---------------------------------------------------------------------------------
    def set_problem(self, fx, fy, cx, cy, rvec, tvec, fixed_rvec, fixed_tvec, pts, obs_cam, obs_pt, obs_uv):
        rvec, tvec = api.f64(rvec).reshape(-1, 3), api.f64(tvec).reshape(-1, 3)
        pts = api.f64(pts).reshape(-1, 3)
        obs_cam, obs_pt = api.i32(obs_cam).ravel(), api.i32(obs_pt).ravel()
        obs_uv = api.f64(obs_uv).reshape(-1, 2)
        frv, ftv = api.f64(fixed_rvec).reshape(3), api.f64(fixed_tvec).reshape(3)
        if tvec.shape != rvec.shape or obs_pt.shape != obs_cam.shape or obs_uv.shape[0] != obs_cam.shape[0]:
            raise ValueError("inconsistent array shapes")
        nc, npt, nobs = rvec.shape[0], pts.shape[0], obs_cam.shape[0]
        api.check(api.lib().ba_set_problem(self._h, float(fx), float(fy), float(cx), float(cy), nc, npt, nobs,
                                           api.dptr(rvec), api.dptr(tvec), api.dptr(frv), api.dptr(ftv), api.dptr(pts),
                                           api.iptr(obs_cam), api.iptr(obs_pt), api.dptr(obs_uv)))
        self._shape = (nc, npt, nobs)

    def set_flat(self, fp):
        """Upload a ``synthetic.FlatProblem`` (or anything with the same attributes)."""
        self.set_problem(fp.fx, fp.fy, fp.cx, fp.cy, fp.rvec, fp.tvec, fp.fixed_rvec, fp.fixed_tvec, fp.pts,
                         fp.obs_cam, fp.obs_pt, fp.obs_uv)

    # -- device-resident map (``ba_map_*``) --------------------------------------------------
    def map_clear(self):
        api.check(api.lib().ba_map_clear(self._h))

    def map_append_observations(self, kf_id: int, mp_ids, uv):
        """Rows of ONE keyframe, in the order of its observation list."""
        mp_ids, uv = api.i32(mp_ids).ravel(), api.f64(uv).reshape(-1, 2)
        if uv.shape[0] != mp_ids.shape[0]:
            raise ValueError("inconsistent array shapes")
        api.check(api.lib().ba_map_append_observations(self._h, int(kf_id), mp_ids.shape[0], api.iptr(mp_ids), api.dptr(uv)))

    def map_set_points(self, mp_ids, xyz):
        mp_ids, xyz = api.i32(mp_ids).ravel(), api.f64(xyz).reshape(-1, 3)
        if xyz.shape[0] != mp_ids.shape[0]:
            raise ValueError("inconsistent array shapes")
        api.check(api.lib().ba_map_set_points(self._h, mp_ids.shape[0], api.iptr(mp_ids), api.dptr(xyz)))

    def map_forget_points(self, mp_ids):
        mp_ids = api.i32(mp_ids).ravel()
        api.check(api.lib().ba_map_forget_points(self._h, mp_ids.shape[0], api.iptr(mp_ids)))

    def map_window(self, fx, fy, cx, cy, kf_ids, rvec, tvec) -> Tuple[int, int]:
        """Build and upload the flat problem of a window from the resident rows: ``kf_ids[0]`` is the fixed keyframe.
        -> (n_pts, n_obs); (0, 0) and no problem when the window is empty."""
        kf_ids = api.i32(kf_ids).ravel()
        n = kf_ids.shape[0]
        rvec, tvec = api.f64(rvec).reshape(n, 3), api.f64(tvec).reshape(n, 3)
        n_pts, n_obs = C.c_int32(), C.c_int32()
        self._shape = None
        api.check(api.lib().ba_map_window(self._h, float(fx), float(fy), float(cx), float(cy), n, api.iptr(kf_ids), api.dptr(rvec),
                                          api.dptr(tvec), C.byref(n_pts), C.byref(n_obs)))
        if n_obs.value > 0:
            self._shape = (n - 1, n_pts.value, n_obs.value)
        return n_pts.value, n_obs.value

    def map_window_points(self) -> np.ndarray:
        _, npt, _ = self._need()
        ids = np.empty(npt, dtype=np.int32)
        api.check(api.lib().ba_map_window_points(self._h, api.iptr(ids)))
        return ids

    def map_commit(self):
        self._need()
        api.check(api.lib().ba_map_commit(self._h))

    def map_stats(self) -> Dict[str, int]:
        out = np.zeros(8, dtype=np.int64)
        api.check(api.lib().ba_map_stats(self._h, api.lptr(out)))
        keys = ("resident_rows", "rows_appended", "points_set", "windows", "bytes_uploaded")
        return {k: int(v) for k, v in zip(keys, out)}

    def _need(self):
        if self._shape is None:
            raise api.BaError(-4, "no problem uploaded")
        return self._shape

    def set_params(self, rvec, tvec, pts):
        nc, npt, _ = self._need()
        rvec, tvec, pts = api.f64(rvec).reshape(nc, 3), api.f64(tvec).reshape(nc, 3), api.f64(pts).reshape(npt, 3)
        api.check(api.lib().ba_set_params(self._h, api.dptr(rvec), api.dptr(tvec), api.dptr(pts)))

    def set_x(self, x):
        """Parameters as the reference's vector [rvecs | tvecs | points] (src/bundle_adjuster.py:162)."""
        nc, npt, _ = self._need()
        x = api.f64(x).ravel()
        self.set_params(x[:3 * nc], x[3 * nc:6 * nc], x[6 * nc:])

    def restore_initial(self):
        """Device-side reset to the parameters of the last ``set_problem``."""
        self._need()
        api.check(api.lib().ba_restore_initial(self._h))

    def timer_start(self):
        api.check(api.lib().ba_timer_start(self._h))

    def timer_stop(self) -> float:
        """Milliseconds of device time since ``timer_start`` on the solver's stream (CUDA events)."""
        ms = C.c_double()
        api.check(api.lib().ba_timer_stop(self._h, C.byref(ms)))
        return ms.value

    def fp64_peak(self) -> float:
        """Measured FP64 FMA rate of the device in TFLOP/s (``ba_fp64_peak``)."""
        v = C.c_double()
        api.check(api.lib().ba_fp64_peak(self._h, C.byref(v)))
        return v.value

    def get_params(self):
        nc, npt, _ = self._need()
        rvec, tvec, pts = np.empty((nc, 3)), np.empty((nc, 3)), np.empty((npt, 3))
        api.check(api.lib().ba_get_params(self._h, api.dptr(rvec), api.dptr(tvec), api.dptr(pts)))
        return rvec, tvec, pts

    def get_x(self) -> np.ndarray:
        rvec, tvec, pts = self.get_params()
        return np.concatenate([rvec.ravel(), tvec.ravel(), pts.ravel()])

    # -- evaluation ------------------------------------------------------------------------------
    def residuals(self) -> np.ndarray:
        _, _, nobs = self._need()
        r = np.empty(2 * nobs)
        api.check(api.lib().ba_residuals(self._h, api.dptr(r)))
        return r

    def cost(self, huber_delta: float = 1.0) -> Tuple[float, float]:
        """(sum r^2, 0.5 * sum huber) at the current parameters (summed over ranks when distributed)."""
        raw, rob = C.c_double(), C.c_double()
        api.check(api.lib().ba_cost(self._h, float(huber_delta), C.byref(raw), C.byref(rob)))
        return raw.value, rob.value

    def jacobian_blocks(self) -> Tuple[np.ndarray, np.ndarray]:
        _, _, nobs = self._need()
        jc, jp = np.empty((nobs, 2, 6)), np.empty((nobs, 2, 3))
        api.check(api.lib().ba_jacobian_blocks(self._h, api.dptr(jc), api.dptr(jp)))
        return jc, jp

    def sparsity_csr(self):
        """(indptr, indices, shape) of the Jacobian pattern, int64."""
        nc, npt, nobs = self._need()
        nnz = C.c_int64()
        api.check(api.lib().ba_sparsity_csr(self._h, None, None, C.byref(nnz)))
        indptr, indices = np.empty(2 * nobs + 1, dtype=np.int64), np.empty(nnz.value, dtype=np.int64)
        api.check(api.lib().ba_sparsity_csr(self._h, api.lptr(indptr), api.lptr(indices), C.byref(nnz)))
        return indptr, indices, (2 * nobs, 6 * nc + 3 * npt)

    def normal_blocks(self, **opts) -> Dict[str, np.ndarray]:
        nc, npt, _ = self._need()
        o = make_options(**opts)
        gc, gp, U, V = np.zeros((nc, 6)), np.zeros((npt, 3)), np.zeros((nc, 6, 6)), np.zeros((npt, 3, 3))
        api.check(api.lib().ba_normal_blocks(self._h, C.byref(o), api.dptr(gc), api.dptr(gp), api.dptr(U), api.dptr(V)))
        return {"gc": gc, "gp": gp, "U": U, "V": V}

    def schur_matvec(self, u, lam: float, **opts) -> np.ndarray:
        """Damped Schur complement at the current parameters applied to u (n_adj, 6); test hook."""
        nc, _, _ = self._need()
        o = make_options(**opts)
        u = api.f64(u).reshape(nc, 6)
        w = np.zeros((nc, 6))
        api.check(api.lib().ba_schur_matvec(self._h, C.byref(o), float(lam), api.dptr(u), api.dptr(w)))
        return w

    # -- solve -----------------------------------------------------------------------------------
    def solve(self, **opts) -> Dict[str, float]:
        self._need()
        o = make_options(**opts)
        st = api.Stats()
        api.check(api.lib().ba_solve(self._h, C.byref(o), C.byref(st)))
        return st.as_dict()

    def solve_trace(self) -> Dict[str, np.ndarray]:
        """Per-LM-iteration record of the last solve, stamped on the device (``ba_solve_trace``): entry 0 = initial point."""
        n = C.c_int32()
        api.check(api.lib().ba_solve_trace(self._h, 0, None, None, None, None, None, None, C.byref(n)))
        k = n.value
        cost, raw, ms, lam = np.zeros(k), np.zeros(k), np.zeros(k), np.zeros(k)
        pcg, acc = np.zeros(k, dtype=np.int32), np.zeros(k, dtype=np.int32)
        api.check(api.lib().ba_solve_trace(self._h, k, api.dptr(cost), api.dptr(raw), api.dptr(ms), api.iptr(pcg), api.iptr(acc),
                                           api.dptr(lam), C.byref(n)))
        return {"cost": cost, "raw_cost": raw, "ms": ms, "pcg_total": pcg, "accepted": acc, "lambda": lam}

    def layout_info(self) -> Dict[str, int]:
        out = np.zeros(12, dtype=np.int64)
        api.check(api.lib().ba_layout_info(self._h, api.lptr(out)))
        keys = ("n_obs", "n_obs_adj", "n_pts", "n_cams", "point_tiles", "segments", "pm_windows", "cm_tiles", "matvec_rows",
                "epochs", "max_epoch_slots", "epoch_tiles")
        return {k: int(v) for k, v in zip(keys, out)}

    def tile_layout(self) -> Dict[str, np.ndarray]:
        """The point-tile structure (test hook, ``ba_tile_layout``)."""
        info = self.layout_info()
        T, Sg, no, nc, npt = info["point_tiles"], info["segments"], info["n_obs"], info["n_cams"], info["n_pts"]
        i32 = lambda n: np.zeros(n, dtype=np.int32)
        out = {"ft_pt": i32(T + 1), "ft_obs": i32(T + 1), "ft_seg": i32(T + 1), "tc_meta": np.zeros((no, 2), dtype=np.int32),
               "tc_uv": np.zeros((no, 2)), "seg_cam": i32(Sg), "seg_info": i32(Sg), "seg_row": i32(Sg),
               "cam_row_off": i32(nc + 2), "pt_perm": i32(npt)}
        api.check(api.lib().ba_tile_layout(
            self._h, api.iptr(out["ft_pt"]), api.iptr(out["ft_obs"]), api.iptr(out["ft_seg"]),
            api.iptr(out["tc_meta"]), api.dptr(out["tc_uv"]), api.iptr(out["seg_cam"]),
            api.iptr(out["seg_info"]), api.iptr(out["seg_row"]), api.iptr(out["cam_row_off"]), api.iptr(out["pt_perm"])))
        return out

    def tile_items(self) -> Dict[str, np.ndarray]:
        """The packed work items of the accumulator mat-vec pass and its tile headers (test hook, ``ba_tile_items``)."""
        n = np.zeros(1, dtype=np.int64)
        api.check(api.lib().ba_tile_items(self._h, api.lptr(n), None, None))
        T = self.layout_info()["point_tiles"]
        items, hdr = np.zeros(int(n[0]), dtype=np.int32), np.zeros((T, 8), dtype=np.int32)
        if n[0] > 0:
            api.check(api.lib().ba_tile_items(self._h, api.lptr(n), api.iptr(items), api.iptr(hdr)))
        return {"items": items, "hdr": hdr}

    def kernel_stats(self) -> Dict[str, Tuple[int, float]]:
        L = api.lib()
        n = L.ba_kernel_count()
        launches, ms = np.zeros(n, dtype=np.int64), np.zeros(n)
        api.check(L.ba_kernel_stats(self._h, api.lptr(launches), api.dptr(ms)))
        return {L.ba_kernel_name(k).decode(): (int(launches[k]), float(ms[k])) for k in range(n)}
