"""Tight converged-cost reference for mid-size problems.  TEST INFRASTRUCTURE ONLY (never shipped, never timed).

Levenberg-Marquardt on the reference's objective (0.5 * sum huber(r_k^2) per scalar residual, 1 px:
``scipy.optimize.least_squares(loss='huber', f_scale=1)`` at ``/root/reference/src/bundle_adjuster.py:170-174``)
with EXACT solves of the reduced camera system: the Schur complement is assembled as a sparse matrix (vectorised
over the pairs of observations that share a point) and factorised with SuperLU, so the iteration can be driven to
``max |g| <= gtol`` on the BASELINE.json shapes (bal50: 294 unknowns in the reduced system, ladybug: 10 332,
venice: 10 662) - the "T1-tight" tier of SURVEY.md section 8c, which SciPy's own trf/lsmr does not reach
(DESIGN.md section 2).  ``oracle/gen_converged_golden.py`` runs it in the build container and commits the converged
costs under ``tests/golden/``; on bal50 the result is cross-checked against SciPy (``solve_scipy_tight``).

Same parameterisation as ``oracle/lm_schur.py`` (world-frame twist per camera, additive points); the stationary
points and the cost are those of the reference's (rvec, t, X) vector.
"""
from __future__ import annotations

import time
from typing import Callable, Optional

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spl

from . import ba_oracle as O
from . import lm_schur as L


def reduced_system(prob: O.Problem, lin: L.Linearization, lam: float, dmin=1e-6, dmax=1e32, chunk_pairs=4_000_000):
    """-> (S csc, rhs (6 Nc,), backsub closure).  S = U + lam D_c - E (V + lam D_p)^-1 E^T as a sparse matrix."""
    adj = prob.obs_cam >= 0
    nc = prob.n_cams
    Dp = np.clip(np.einsum("nii->ni", lin.V), dmin, dmax)
    Dc = np.clip(np.einsum("nii->ni", lin.U), dmin, dmax)
    Vd = lin.V.copy()
    Vd[:, [0, 1, 2], [0, 1, 2]] += lam * Dp
    Vinv = np.linalg.inv(Vd)
    ca, pa = prob.obs_cam[adj], prob.obs_pt[adj]
    E = np.einsum("nki,nk,nkj->nij", lin.Jc[adj], lin.wH[adj], lin.Jp[adj])          # (Na,6,3)
    EV = np.einsum("nij,njk->nik", E, Vinv[pa])
    rhs = -lin.gc.copy()
    np.add.at(rhs, ca, np.einsum("nij,nj->ni", EV, lin.gp[pa]))
    order = np.argsort(pa, kind="stable")
    pa_s, ca_s, E_s, EV_s = pa[order], ca[order], E[order], EV[order]
    cnt = np.bincount(pa_s, minlength=prob.n_pts)
    start = np.concatenate([[0], np.cumsum(cnt)])
    # all ordered pairs (i, j) of observations of one point, chunked over observations i
    n_pairs_of = cnt[pa_s]
    cum = np.concatenate([[0], np.cumsum(n_pairs_of)])
    Sm = sp.csr_matrix((6 * nc, 6 * nc))
    a = 0
    r6, c6 = np.arange(6)[None, :, None], np.arange(6)[None, None, :]
    while a < pa_s.size:
        b = int(np.searchsorted(cum, cum[a] + chunk_pairs, side="right")) - 1
        b = min(max(b, a + 1), pa_s.size)
        li = np.repeat(np.arange(a, b), n_pairs_of[a:b])
        off = np.arange(li.size) - np.repeat(cum[a:b] - cum[a], n_pairs_of[a:b])
        lj = start[pa_s[li]] + off
        blk = -np.einsum("nij,nkj->nik", EV_s[li], E_s[lj])
        rr = (6 * ca_s[li][:, None, None] + r6 + 0 * c6).ravel()
        cc = (6 * ca_s[lj][:, None, None] + 0 * r6 + c6).ravel()
        Sm = Sm + sp.coo_matrix((blk.ravel(), (rr, cc)), shape=(6 * nc, 6 * nc)).tocsr()
        a = b
    Ub = lin.U.copy()
    Ub[:, np.arange(6), np.arange(6)] += lam * Dc
    Ud = sp.block_diag(list(Ub), format="csr") if nc else sp.csr_matrix((0, 0))
    S = (Sm + Ud).tocsc()

    def backsub(dc):
        w = lin.gp.copy()
        np.add.at(w, pa, np.einsum("nij,ni->nj", E, dc[ca]))
        dp = -np.einsum("nij,nj->ni", Vinv, w)
        model = -(np.sum(lin.gc * dc) + np.sum(lin.gp * dp)) - 0.5 * (
            np.einsum("ni,nij,nj->", dc, lin.U, dc) + np.einsum("ni,nij,nj->", dp, lin.V, dp)
            + 2.0 * np.einsum("ni,nij,nj->", dc[ca], E, dp[pa]))
        return dp, float(model)
    return S, rhs.ravel(), backsub


def gradient_max(lin: L.Linearization) -> float:
    return float(max(np.abs(lin.gc).max() if lin.gc.size else 0.0, np.abs(lin.gp).max()))


def solve(prob: O.Problem, rvec, tvec, pts, *, mode="irls", max_iters=200, lam0=1e-4, lam_min=1e-14, lam_max=1e16,
          gtol=1e-8, ftol=0.0, verbose=False, callback: Optional[Callable] = None, linesearch=True) -> L.LMResult:
    """Exact-step LM (Nielsen damping, per-point line search as on the device) until max|g| < gtol (or a relative cost
    decrease below ftol)."""
    rvec, tvec, pts = rvec.copy(), tvec.copy(), pts.copy()
    lam, nu = lam0, 2.0
    lin = L.linearize(prob, rvec, tvec, pts, mode, True)
    hist, status, it = [], 0, 0
    for it in range(1, max_iters + 1):
        gmax = gradient_max(lin)
        if gmax < gtol:
            status = 1
            break
        t0 = time.perf_counter()
        S, rhs, backsub = reduced_system(prob, lin, lam)
        dc = spl.splu(S).solve(rhs).reshape(prob.n_cams, 6) if prob.n_cams else np.zeros((0, 6))
        dp, model = backsub(dc)
        rv, tv, pn = L.retract(rvec, tvec, pts, dc, dp)
        if linesearch:
            pn, _ = L.point_linesearch(prob, rv, tv, pts, pn, True)
        lin_new = L.linearize(prob, rv, tv, pn, mode, True)
        rho = (lin.cost - lin_new.cost) / model if model > 0 else -1.0
        accepted = bool(np.isfinite(lin_new.cost) and rho > 0)
        rec = dict(it=it, lam=lam, cost=lin.cost, cost_new=lin_new.cost, model=model, rho=rho, accepted=accepted, gmax=gmax,
                   step=float(np.sqrt((dc * dc).sum() + (dp * dp).sum())), seconds=time.perf_counter() - t0)
        hist.append(rec)
        if verbose:
            print(f"it {it:3d} {'acc' if accepted else 'REJ'} cost {lin.cost:.12e} -> {lin_new.cost:.12e} rho {rho:+.3e} "
                  f"lam {lam:.1e} |g| {gmax:.3e} |dx| {rec['step']:.3e} ({rec['seconds']:.1f} s)", flush=True)
        if callback:
            callback(rec)
        if accepted:
            dcost = lin.cost - lin_new.cost
            rvec, tvec, pts, lin = rv, tv, pn, lin_new
            lam = max(lam * max(1.0 / 3.0, 1.0 - (2.0 * rho - 1.0) ** 3), lam_min)
            nu = 2.0
            if dcost <= ftol * lin.cost:
                status = 2
                break
        else:
            lam *= nu
            nu *= 2.0
            if lam > lam_max:
                status = 5
                break
    return L.LMResult(rvec, tvec, pts, lin.cost, lin.raw, it, status, hist)


def analytic_jacobian_csr(prob: O.Problem, x: np.ndarray) -> sp.csr_matrix:
    """d r / d x in the REFERENCE's parameter vector [rvecs | tvecs | points] (``src/bundle_adjuster.py:157-162``) from the
    oracle's analytic blocks, with exactly the sparsity pattern of ``_prepare_sparsity_matrix`` (``:74-120``)."""
    Jc, Jp = prob.jacobian_blocks(x)                  # (No,2,6) w.r.t. (rvec, t); (No,2,3)
    no, nc = prob.n_obs, prob.n_cams
    adj = np.flatnonzero(prob.obs_cam >= 0)
    rows_p = (2 * np.arange(no)[:, None, None] + np.arange(2)[None, :, None] + 0 * np.arange(3)[None, None, :])
    cols_p = (6 * nc + 3 * prob.obs_pt[:, None, None] + 0 * np.arange(2)[None, :, None] + np.arange(3)[None, None, :])
    c = prob.obs_cam[adj]
    rows_c = (2 * adj[:, None, None] + np.arange(2)[None, :, None] + 0 * np.arange(6)[None, None, :])
    k = np.arange(6)[None, None, :]
    cols_c = np.where(k < 3, 3 * c[:, None, None] + k, 3 * nc + 3 * c[:, None, None] + (k - 3)) + 0 * np.arange(2)[None, :, None]
    data = np.concatenate([Jp.ravel(), Jc[adj].ravel()])
    rows = np.concatenate([rows_p.ravel(), rows_c.ravel()])
    cols = np.concatenate([cols_p.ravel(), cols_c.ravel()])
    return sp.csr_matrix((data, (rows, cols)), shape=(2 * no, 6 * nc + 3 * prob.n_pts))


def solve_scipy_tight(prob: O.Problem, x0: np.ndarray, *, tol=1e-12, max_nfev=2000):
    """The reference's call with the analytic sparse Jacobian instead of 2-point differences and tight tolerances:
    ``least_squares(fun, x0, jac=<CSR>, loss='huber', tr_solver='lsmr', xtol=ftol=gtol=tol)`` - the SciPy-side
    cross-check of the converged cost (feasible on bal50)."""
    from scipy.optimize import least_squares
    return least_squares(prob.residuals, x0, jac=lambda x: analytic_jacobian_csr(prob, x), loss="huber", f_scale=1.0,
                         tr_solver="lsmr", xtol=tol, ftol=tol, gtol=tol, max_nfev=max_nfev, method="trf")
