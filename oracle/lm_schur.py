"""numpy restatement of the DEVICE algorithm (LM + Schur complement), TEST INFRASTRUCTURE ONLY.

The reference solves the local-BA objective with SciPy's trust-region-reflective method
(``src/bundle_adjuster.py:170-174``).  The CUDA path minimises the SAME objective
(0.5 * sum(huber(r_k^2)) per scalar residual, threshold 1 px) with Levenberg-Marquardt on the Schur
complement, as BASELINE.json's north_star prescribes.  This file restates that LM loop on the CPU with
exact (dense Cholesky) solves of the reduced camera system, so the GPU tests can check the kernels'
intermediate quantities (gradient, Hessian blocks, Schur right-hand side, step, cost trajectory) and not
only the final cost.  It is not used for timing and not shipped.

Parameterisation (matches ``bundle_adjustment_b200/csrc/ba_kernels.cuh``): the camera increment is the
world-frame twist (w, v) with  x_cam = R (X + w x X + v) + t,  i.e.  R <- R Exp(w),  t <- t + R v;  the point
increment is additive.  The stationary points and the cost are those of the reference's
(rvec, t, X) parameterisation because the map between the two increments is a smooth bijection
(dw = J_r(rvec) d rvec, v = R^T dt).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List

import numpy as np

from . import ba_oracle as O


def skew(v):
    z = np.zeros_like(v[..., 0])
    return np.stack([np.stack([z, -v[..., 2], v[..., 1]], -1),
                     np.stack([v[..., 2], z, -v[..., 0]], -1),
                     np.stack([-v[..., 1], v[..., 0], z], -1)], -2)


def log_so3(R):
    """Rotation matrix (3,3) -> rotation vector via the unit quaternion (stable for all angles < pi)."""
    tr = R[0, 0] + R[1, 1] + R[2, 2]
    if tr > 0:
        s = np.sqrt(tr + 1.0) * 2
        q = np.array([0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s])
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = np.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = np.array([(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s])
    elif R[1, 1] > R[2, 2]:
        s = np.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = np.array([(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s])
    else:
        s = np.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = np.array([(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s])
    if q[0] < 0:
        q = -q
    n = np.linalg.norm(q[1:])
    if n < 1e-12:
        return 2.0 * q[1:]
    return q[1:] * (2.0 * np.arctan2(n, q[0]) / n)


@dataclass
class Linearization:
    cost: float                 # 0.5 * sum rho
    raw: float                  # sum r^2
    r: np.ndarray               # (No,2) residuals
    wH: np.ndarray              # (No,2) curvature weights
    wg: np.ndarray              # (No,2) gradient weights
    Jp: np.ndarray              # (No,2,3)
    Jc: np.ndarray              # (No,2,6) local (world-twist) camera Jacobian, zero for fixed-keyframe rows
    V: np.ndarray               # (Np,3,3)
    gp: np.ndarray              # (Np,3)
    U: np.ndarray               # (Nc,6,6)
    gc: np.ndarray              # (Nc,6)


@dataclass
class LMResult:
    rvec: np.ndarray
    tvec: np.ndarray
    pts: np.ndarray
    cost: float
    raw: float
    iterations: int
    status: int
    history: List[dict] = field(default_factory=list)

    def x(self):
        return np.concatenate([self.rvec.ravel(), self.tvec.ravel(), self.pts.ravel()])


def weights(r, mode: str):
    """(curvature weight, gradient weight) per scalar residual.  'irls': both rho'(r^2);
    'triggs': SciPy's scale_for_robust_loss_function (max(rho' + 2 rho'' r^2, EPS), rho')."""
    _, rho1, rho2 = O.huber_rho(r * r)
    if mode == "irls":
        return rho1, rho1
    if mode == "triggs":
        return np.maximum(rho1 + 2.0 * rho2 * r * r, O.DBL_EPS), rho1
    if mode == "none":
        return np.ones_like(r), np.ones_like(r)
    if mode.startswith("hybrid:"):
        # curvature weight kappa * rho' in the linear regime (kappa = 1: IRLS, kappa = 0: the true second derivative)
        kappa = float(mode.split(":")[1])
        return np.where(rho2 < 0.0, np.maximum(kappa * rho1, O.DBL_EPS), rho1), rho1
    raise ValueError(mode)


def linearize(prob: O.Problem, rvec, tvec, pts, mode="irls", huber=True) -> Linearization:
    x = np.concatenate([rvec.ravel(), tvec.ravel(), pts.ravel()])
    xc, Rg, P = prob.camera_points(x)
    r = prob.residuals(x).reshape(-1, 2)
    iz = 1.0 / xc[:, 2]
    dpi = np.zeros((prob.n_obs, 2, 3))
    dpi[:, 0, 0] = prob.fx * iz
    dpi[:, 0, 2] = -prob.fx * xc[:, 0] * iz * iz
    dpi[:, 1, 1] = prob.fy * iz
    dpi[:, 1, 2] = -prob.fy * xc[:, 1] * iz * iz
    Jp = -np.einsum("nij,njk->nik", dpi, Rg)
    G = np.concatenate([-skew(P), np.broadcast_to(np.eye(3), (prob.n_obs, 3, 3))], axis=2)   # (No,3,6)
    Jc = np.einsum("nij,njk->nik", Jp, G)
    adj = prob.obs_cam >= 0
    Jc[~adj] = 0.0
    if huber and mode.startswith("guarded"):
        # Per-point guarded second-order weights: a point whose TRUE curvature (rows in the quadratic regime only, the
        # Triggs / SciPy weights) is at least tau times its IRLS curvature in every direction uses the true weights for all
        # its rows; any other point (too few inlier rows at the current iterate) keeps the IRLS majoriser.
        tau = float(mode.split(":")[1]) if ":" in mode else 0.1
        wT, wg = weights(r, "triggs")
        wI, _ = weights(r, "irls")
        Vt = np.zeros((prob.n_pts, 3, 3)); Vi = np.zeros((prob.n_pts, 3, 3))
        np.add.at(Vt, prob.obs_pt, np.einsum("nki,nk,nkj->nij", Jp, wT, Jp))
        np.add.at(Vi, prob.obs_pt, np.einsum("nki,nk,nkj->nij", Jp, wI, Jp))
        # min generalised eigenvalue of (Vt, Vi) through the Cholesky factor of Vi
        Lc = np.linalg.cholesky(Vi + 1e-300 * np.eye(3))
        Li = np.linalg.inv(Lc)
        M = np.einsum("nij,njk,nlk->nil", Li, Vt, Li)
        safe = np.linalg.eigvalsh(M)[:, 0] >= tau
        wH = np.where(safe[prob.obs_pt][:, None], wT, wI)
    else:
        wH, wg = weights(r, mode if huber else "none")
    cost = O.huber_cost(r.ravel()) if huber else 0.5 * float((r * r).sum())
    V = np.zeros((prob.n_pts, 3, 3)); gp = np.zeros((prob.n_pts, 3))
    U = np.zeros((prob.n_cams, 6, 6)); gc = np.zeros((prob.n_cams, 6))
    np.add.at(V, prob.obs_pt, np.einsum("nki,nk,nkj->nij", Jp, wH, Jp))
    np.add.at(gp, prob.obs_pt, np.einsum("nki,nk->ni", Jp, wg * r))
    np.add.at(U, prob.obs_cam[adj], np.einsum("nki,nk,nkj->nij", Jc[adj], wH[adj], Jc[adj]))
    np.add.at(gc, prob.obs_cam[adj], np.einsum("nki,nk->ni", Jc[adj], (wg * r)[adj]))
    return Linearization(cost, float((r * r).sum()), r, wH, wg, Jp, Jc, V, gp, U, gc)


def schur_step(prob: O.Problem, lin: Linearization, lam: float, dmin=1e-6, dmax=1e32):
    """Solve (H + lam D) delta = -g exactly through the Schur complement.  D = clamp(diag(H))."""
    adj = prob.obs_cam >= 0
    nc = prob.n_cams
    Dp = np.clip(np.einsum("nii->ni", lin.V), dmin, dmax)
    Dc = np.clip(np.einsum("nii->ni", lin.U), dmin, dmax)
    Vd = lin.V + lam * np.einsum("ni,ij->nij", Dp, np.eye(3))
    Vinv = np.linalg.inv(Vd)
    E = np.einsum("nki,nk,nkj->nij", lin.Jc[adj], lin.wH[adj], lin.Jp[adj])          # (Na,6,3)
    ca, pa = prob.obs_cam[adj], prob.obs_pt[adj]
    S = np.zeros((6 * nc, 6 * nc))
    for c in range(nc):
        S[6 * c:6 * c + 6, 6 * c:6 * c + 6] = lin.U[c] + lam * np.diag(Dc[c])
    EV = np.einsum("nij,njk->nik", E, Vinv[pa])                                      # (Na,6,3)
    rhs = -lin.gc.copy()
    np.add.at(rhs, ca, np.einsum("nij,nj->ni", EV, lin.gp[pa]))
    # S -= sum_p E_cp Vinv E_c'p^T over camera pairs sharing a point
    order = np.argsort(pa, kind="stable")
    pa_s, ca_s, E_s, EV_s = pa[order], ca[order], E[order], EV[order]
    bounds = np.flatnonzero(np.diff(pa_s)) + 1
    starts = np.concatenate([[0], bounds]); ends = np.concatenate([bounds, [pa_s.size]])
    for a, b in zip(starts, ends):
        for i in range(a, b):
            for j in range(a, b):
                S[6 * ca_s[i]:6 * ca_s[i] + 6, 6 * ca_s[j]:6 * ca_s[j] + 6] -= EV_s[i] @ E_s[j].T
    dc = np.linalg.solve(S, rhs.ravel()).reshape(nc, 6)
    w = lin.gp.copy()
    np.add.at(w, pa, np.einsum("nij,ni->nj", E, dc[ca]))
    dp = -np.einsum("nij,nj->ni", Vinv, w)
    model = -(np.sum(lin.gc * dc) + np.sum(lin.gp * dp)) - 0.5 * (
        np.einsum("ni,nij,nj->", dc, lin.U, dc) + np.einsum("ni,nij,nj->", dp, lin.V, dp)
        + 2.0 * np.einsum("ni,nij,nj->", dc[ca], E, dp[pa]))
    return dc, dp, float(model), {"rhs": rhs, "S": S, "Dp": Dp, "Dc": Dc, "Vinv": Vinv}


def retract(rvec, tvec, pts, dc, dp):
    R = O.rodrigues(rvec)
    Rn = np.einsum("nij,njk->nik", R, O.rodrigues(dc[:, :3]))
    rv = np.array([log_so3(Rn[i]) for i in range(Rn.shape[0])]).reshape(-1, 3)
    tv = tvec + np.einsum("nij,nj->ni", R, dc[:, 3:])
    return rv, tv, pts + dp


LS_ALPHAS = (0.5, 2.0, 4.0, 8.0, 16.0)   # tried after alpha = 1, in this order (k_tile_linesearch: kLsAlpha)


def point_costs(prob: O.Problem, rvec, tvec, pts, huber=True) -> np.ndarray:
    """0.5 * sum rho over the observations of each point (n_pts,); non-finite -> 1e300 per observation, as on the device."""
    x = np.concatenate([np.asarray(rvec).ravel(), np.asarray(tvec).ravel(), np.asarray(pts).ravel()])
    with np.errstate(all="ignore"):
        r = prob.residuals(x).reshape(-1, 2)
        c = 0.5 * (O.huber_rho(r * r)[0] if huber else r * r).sum(1)
    c = np.where(np.isfinite(c), c, 1e300)
    out = np.zeros(prob.n_pts)
    np.add.at(out, prob.obs_pt, c)
    return out


def point_linesearch(prob: O.Problem, rvec_new, tvec_new, pts_cur, pts_new, huber=True):
    """CPU statement of k_tile_linesearch (csrc/ba_tiles.cuh): with the cameras at their trial poses every point scales its
    own step by the factor out of {1, 1/2, 2, 4, 8, 16} with the smallest cost over its observations; a factor other than 1
    must win by more than 1e-9 relative.  -> (points, chosen factors)."""
    dp = pts_new - pts_cur
    c1 = point_costs(prob, rvec_new, tvec_new, pts_new, huber)
    thr = c1 - 1e-9 * np.abs(c1)
    best, alpha = c1.copy(), np.ones(prob.n_pts)
    for a in LS_ALPHAS:
        c = point_costs(prob, rvec_new, tvec_new, pts_cur + a * dp, huber)
        m = (c < thr) & (c < best)
        best[m] = c[m]
        alpha[m] = a
    return pts_cur + alpha[:, None] * dp, alpha


def solve(prob: O.Problem, rvec, tvec, pts, *, mode="irls", huber=True, max_iters=100, lam0=1e-4, ftol=1e-12,
          gtol=1e-10, lam_min=1e-12, lam_max=1e16, verbose=False, linesearch=True) -> LMResult:
    """Levenberg-Marquardt with Nielsen's damping update and the per-point line search (the rules the device loop uses)."""
    rvec, tvec, pts = rvec.copy(), tvec.copy(), pts.copy()
    lam, nu = lam0, 2.0
    lin = linearize(prob, rvec, tvec, pts, mode, huber)
    hist, status, it = [], 0, 0
    for it in range(1, max_iters + 1):
        gmax = max(np.abs(lin.gc).max() if lin.gc.size else 0.0, np.abs(lin.gp).max())
        if gmax < gtol:
            status = 1
            break
        dc, dp, model, _ = schur_step(prob, lin, lam)
        rv, tv, pn = retract(rvec, tvec, pts, dc, dp)
        if linesearch:
            pn, _ = point_linesearch(prob, rv, tv, pts, pn, huber)
        lin_new = linearize(prob, rv, tv, pn, mode, huber)
        rho = (lin.cost - lin_new.cost) / model if model > 0 else -1.0
        accepted = bool(np.isfinite(lin_new.cost) and rho > 0)
        hist.append(dict(it=it, lam=lam, cost=lin.cost, cost_new=lin_new.cost, model=model, rho=rho,
                         accepted=accepted, gmax=gmax))
        if verbose:
            print(hist[-1])
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
    return LMResult(rvec, tvec, pts, lin.cost, lin.raw, it, status, hist)


# ------------------------------------------------------------------------------------------------------------------
# Preconditioned CG on the reduced camera system, as the device runs it, and the two-level preconditioner planned for
# round 2.  TEST INFRASTRUCTURE: the CPU statement the device kernels will be checked
# against; the study behind the choice is scripts/precond_study.py.
# ------------------------------------------------------------------------------------------------------------------
class SchurOperator:
    """Implicit S u = (U + lam D) u - E V^-1 E^T u of one linearisation - what the tile pass (+ lam D u in the tail)
    computes - and its 6x6 diagonal blocks (the Schur-Jacobi preconditioner)."""

    def __init__(self, prob: O.Problem, lin: Linearization, lam: float, dmin=1e-6, dmax=1e32):
        adj = prob.obs_cam >= 0
        self.nc = prob.n_cams
        self.ca, self.pa = prob.obs_cam[adj], prob.obs_pt[adj]
        Dp = np.clip(np.einsum("nii->ni", lin.V), dmin, dmax)
        self.Dc = np.clip(np.einsum("nii->ni", lin.U), dmin, dmax)
        self.Vinv = np.linalg.inv(lin.V + lam * np.einsum("ni,ij->nij", Dp, np.eye(3)))
        self.E = np.einsum("nki,nk,nkj->nij", lin.Jc[adj], lin.wH[adj], lin.Jp[adj])
        self.U, self.lam, self.n_pts = lin.U, lam, prob.n_pts
        self.rhs = -lin.gc.copy()
        np.add.at(self.rhs, self.ca, np.einsum("nij,njk,nk->ni", self.E, self.Vinv[self.pa], lin.gp[self.pa]))

    def matvec(self, u):
        u = u.reshape(self.nc, 6)
        s = np.zeros((self.n_pts, 3))
        np.add.at(s, self.pa, np.einsum("nij,ni->nj", self.E, u[self.ca]))
        z = np.einsum("nij,nj->ni", self.Vinv, s)
        w = np.einsum("nij,nj->ni", self.U, u) + self.lam * self.Dc * u
        np.subtract.at(w, self.ca, np.einsum("nij,nj->ni", self.E, z[self.pa]))
        return w.ravel()

    def jacobi_blocks(self):
        M = self.U + self.lam * np.einsum("ni,ij->nij", self.Dc, np.eye(6))
        EV = np.einsum("nij,njk->nik", self.E, self.Vinv[self.pa])
        np.subtract.at(M, self.ca, np.einsum("nij,nkj->nik", EV, self.E))
        return np.linalg.inv(M)


def coarse_prolongation(n_cams: int, agg: int) -> np.ndarray:
    """Z (6 n_cams x 6 n_agg): one rigid 6-vector per aggregate of `agg` consecutive cameras (piecewise constant)."""
    n_agg = (n_cams + agg - 1) // agg
    Z = np.zeros((6 * n_cams, 6 * n_agg))
    c = np.arange(n_cams)
    for d in range(6):
        Z[6 * c + d, 6 * (c // agg) + d] = 1.0
    return Z


def coarse_matrix_by_probing(matvec, Z: np.ndarray) -> np.ndarray:
    """Z^T S Z column by column: one mat-vec per coarse unknown (what the device does with its tile pass)."""
    A0 = np.stack([Z.T @ matvec(Z[:, j]) for j in range(Z.shape[1])], axis=1)
    return 0.5 * (A0 + A0.T)


def coarse_matrix_direct(prob: O.Problem, lin: Linearization, pts, lam: float, agg: int, dmin=1e-6, dmax=1e32,
                         integer_bits=58):
    """Z^T S Z without a mat-vec - the CPU statement of the device's k_coarse_direct (csrc/ba_peer.cuh).  With local camera
    coordinates Jc = Jp G_p, G_p = [-[X_p]x | I], so the E blocks of point p summed over the cameras of aggregate A are
    F_{A,p} = W_{A,p} G_p, W_{A,p} = sum of Jp^T wH Jp over the point's observations from A, and
        Z^T S Z = Z^T U Z + lam Z^T D Z - sum_p F_p^T (V_p + lam D_p)^-1 F_p.
    integer_bits: the point part is accumulated like on the device, as integers after scaling entry (i, j) by
    2^bits / sqrt(d_i d_j), d = diag(Z^T U Z) (python integers: exact; the test checks the bound |sum| <= 2^bits that the
    device's 64-bit accumulators rely on).  None = plain float64 sums."""
    nc, n_agg = prob.n_cams, (prob.n_cams + agg - 1) // agg
    nz = 6 * n_agg
    adj = prob.obs_cam >= 0
    ca, pa = prob.obs_cam[adj], prob.obs_pt[adj]
    aa = ca // agg
    W = np.einsum("nki,nk,nkj->nij", lin.Jp[adj], lin.wH[adj], lin.Jp[adj])                     # (Na,3,3)
    Dp = np.clip(np.einsum("nii->ni", lin.V), dmin, dmax)
    Vinv = np.linalg.inv(lin.V + lam * np.einsum("ni,ij->nij", Dp, np.eye(3)))
    G = np.concatenate([-skew(pts), np.broadcast_to(np.eye(3), (prob.n_pts, 3, 3))], axis=2)   # (Np,3,6)
    # one entry per (point, aggregate): F = W_{A,p} G_p
    key = pa.astype(np.int64) * n_agg + aa
    uniq, inv = np.unique(key, return_inverse=True)
    WA = np.zeros((len(uniq), 3, 3))
    np.add.at(WA, inv, W)
    ep, ea = uniq // n_agg, uniq % n_agg
    F = np.einsum("nij,njk->nik", WA, G[ep])                                                   # (Ne,3,6)
    ZUZ = np.zeros((nz, nz))
    for c in range(nc):
        a = c // agg
        ZUZ[6 * a:6 * a + 6, 6 * a:6 * a + 6] += lin.U[c]
    d = np.diag(ZUZ).copy()
    rscale = np.where(d > 0.0, 1.0 / np.sqrt(np.where(d > 0.0, d, 1.0)), 0.0)
    point_part = np.zeros((nz, nz))
    acc = np.zeros((nz, nz), dtype=object) if integer_bits is not None else None
    if acc is not None:
        acc[:] = 0
    order = np.argsort(ep, kind="stable")
    start = np.searchsorted(ep[order], np.arange(prob.n_pts + 1))
    for p in range(prob.n_pts):
        es = order[start[p]:start[p + 1]]
        for e in es:
            HeT = F[e].T @ Vinv[p]                                                             # (6,3)
            for f in es:
                T = HeT @ F[f]                                                                 # block (A_e, A_f)
                i0, j0 = 6 * ea[e], 6 * ea[f]
                if acc is None:
                    point_part[i0:i0 + 6, j0:j0 + 6] += T
                else:
                    sc = np.outer(rscale[i0:i0 + 6], rscale[j0:j0 + 6]) * 2.0 ** integer_bits
                    q = np.rint(T * sc)
                    for i in range(6):
                        for j in range(6):
                            acc[i0 + i, j0 + j] += int(q[i, j])
    if acc is not None:
        big = max(abs(int(v)) for v in acc.ravel())
        assert big <= 2 ** integer_bits, "the Cauchy-Schwarz bound of the scaled sums"
        sc = np.outer(rscale, rscale) * 2.0 ** integer_bits
        point_part = np.where(sc > 0.0, acc.astype(float) / np.where(sc > 0.0, sc, 1.0), 0.0)
    Dc = np.clip(np.einsum("nii->ni", lin.U), dmin, dmax)
    ZDZ = np.zeros(nz)
    np.add.at(ZDZ, (6 * (np.arange(nc) // agg))[:, None] + np.arange(6)[None, :], Dc)
    A0 = ZUZ + lam * np.diag(ZDZ) - point_part
    return 0.5 * (A0 + A0.T)


def pcg(matvec, b, Minv_blocks, *, tol=0.1, max_iters=100, Z=None, A0_inv=None):
    """CG preconditioned with the 6x6 blocks `Minv_blocks` (n_cams,6,6) and, if Z / A0_inv are given, the additive coarse
    correction  M^-1 r = blockdiag r + Z A0^-1 Z^T r.  Stops like the device: r^T M^-1 r <= tol^2 r0^T M^-1 r0.
    -> (x, iterations, final r^T M^-1 r / initial)."""
    nc = Minv_blocks.shape[0]

    def prec(r):
        u = np.einsum("nij,nj->ni", Minv_blocks, r.reshape(nc, 6)).ravel()
        if Z is not None:
            u = u + Z @ (A0_inv @ (Z.T @ r))
        return u

    x = np.zeros_like(b)
    r = b.copy()
    u = prec(r)
    p = u.copy()
    g0 = g = float(r @ u)
    it = 0
    if not g0 > 0.0:
        return x, 0, 0.0
    for it in range(1, max_iters + 1):
        Ap = matvec(p)
        alpha = g / float(p @ Ap)
        x += alpha * p
        r -= alpha * Ap
        u = prec(r)
        gn = float(r @ u)
        if not gn > tol * tol * g0:
            g = gn
            break
        p = u + (gn / g) * p
        g = gn
    return x, it, g / g0
