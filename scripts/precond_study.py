"""CPU study for DESIGN.md section 8 item 2: how many PCG iterations would other preconditioners of the reduced camera
system need?  Builds S = U + lambda D - E V^-1 E^T of a venice-like synthetic problem (same generator, all 1778
cameras, fewer points) with the CPU restatement of the device algorithm, and counts PCG iterations to the solver's
stopping rule (r^T M^-1 r <= 0.01 r0^T M^-1 r0) for

    SJ      block-Jacobi on S, 6x6 (what the device does)
    CJ-k    block-Jacobi over clusters of k consecutive cameras (6k x 6k dense blocks of S)
    AS-k    additive Schwarz over clusters of k cameras overlapping by k/2
    CJ-k+C  CJ-k plus a coarse correction (one rigid 6-vector per cluster, Galerkin coarse matrix)
    BT      block-tridiagonal part of S, exact (what a chain-local preconditioner can reach)

TEST / ANALYSIS INFRASTRUCTURE (imports oracle/); nothing here is shipped or timed.

    python scripts/precond_study.py [n_cams] [n_pts] [n_obs]
"""
import os
import sys
import time

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spl

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import synthetic as S_  # noqa: E402
from oracle import ba_oracle as O, lm_schur as L  # noqa: E402

n_cams = int(sys.argv[1]) if len(sys.argv) > 1 else 1778
n_pts = int(sys.argv[2]) if len(sys.argv) > 2 else 30000
n_obs = int(sys.argv[3]) if len(sys.argv) > 3 else 150000


def build_S(prob, lin, lam, dmin=1e-6, dmax=1e32):
    adj = prob.obs_cam >= 0
    nc = prob.n_cams
    Dp = np.clip(np.einsum("nii->ni", lin.V), dmin, dmax)
    Dc = np.clip(np.einsum("nii->ni", lin.U), dmin, dmax)
    Vinv = np.linalg.inv(lin.V + lam * np.einsum("ni,ij->nij", Dp, np.eye(3)))
    E = np.einsum("nki,nk,nkj->nij", lin.Jc[adj], lin.wH[adj], lin.Jp[adj])
    ca, pa = prob.obs_cam[adj], prob.obs_pt[adj]
    EV = np.einsum("nij,njk->nik", E, Vinv[pa])
    rhs = -lin.gc.copy()
    np.add.at(rhs, ca, np.einsum("nij,nj->ni", EV, lin.gp[pa]))
    order = np.argsort(pa, kind="stable")
    pa_s, ca_s, E_s, EV_s = pa[order], ca[order], E[order], EV[order]
    cnt = np.bincount(pa_s, minlength=prob.n_pts)
    start = np.concatenate([[0], np.cumsum(cnt)])[:-1]
    # all ordered pairs (i, j) of observations of the same point
    li = np.repeat(np.arange(pa_s.size), cnt[pa_s])
    off = np.arange(li.size) - np.repeat(np.cumsum(cnt[pa_s]) - cnt[pa_s], cnt[pa_s])
    lj = start[pa_s[li]] + off
    blk = -np.einsum("nij,nkj->nik", EV_s[li], E_s[lj])                      # (pairs, 6, 6)
    bi, bj = ca_s[li], ca_s[lj]
    rr = (6 * bi[:, None, None] + np.arange(6)[None, :, None] + 0 * np.arange(6)[None, None, :]).ravel()
    cc = (6 * bj[:, None, None] + 0 * np.arange(6)[None, :, None] + np.arange(6)[None, None, :]).ravel()
    Sm = sp.coo_matrix((blk.ravel(), (rr, cc)), shape=(6 * nc, 6 * nc)).tocsr()
    Ud = sp.block_diag([lin.U[c] + lam * np.diag(Dc[c]) for c in range(nc)], format="csr")
    return (Sm + Ud).tocsr(), rhs.ravel()


def pcg(A, b, Minv, tol=0.1, maxit=1000, want_x=False):
    x = np.zeros_like(b); r = b.copy(); u = Minv(r); p = u.copy()
    g0 = g = r @ u
    it = 0
    for it in range(1, maxit + 1):
        Ap = A @ p
        alpha = g / (p @ Ap)
        x += alpha * p; r -= alpha * Ap
        u = Minv(r); gn = r @ u
        if gn <= tol * tol * g0:
            break
        p = u + (gn / g) * p; g = gn
    return (it, x) if want_x else it


def cluster_precond(A, nc, k, overlap=0, coarse=False):
    n = 6 * nc
    step = k - overlap
    blocks = []
    c0 = 0
    while c0 < nc:
        c1 = min(c0 + k, nc)
        blocks.append((6 * c0, 6 * c1))
        if c1 == nc:
            break
        c0 += step
    inv = [np.linalg.inv(A[a:b, a:b].toarray()) for a, b in blocks]
    Z = A0 = None
    if coarse:
        rows = np.arange(n)
        cols = 6 * np.minimum(rows // (6 * k), (nc - 1) // k) + rows % 6
        Z = sp.csr_matrix((np.ones(n), (rows, cols)))
        A0 = spl.splu((Z.T @ A @ Z).tocsc())

    def apply(r):
        out = np.zeros_like(r)
        for (a, b), Bi in zip(blocks, inv):
            out[a:b] += Bi @ r[a:b]
        if coarse:
            out += Z @ A0.solve(Z.T @ r)
        return out
    return apply


def coarse_space(nc, k, dofs=6):
    """Piecewise-constant prolongation: aggregate a of k consecutive cameras, one value per aggregate and camera dof."""
    n = 6 * nc
    rows = np.arange(n)
    keep = (rows % 6) >= 6 - dofs if dofs < 6 else np.ones(n, dtype=bool)      # dofs = 3: the translation part (v) only
    cols = dofs * np.minimum(rows // (6 * k), (nc - 1) // k) + (rows % 6) - (6 - dofs)
    return sp.csr_matrix((np.ones(keep.sum()), (rows[keep], cols[keep])), shape=(n, dofs * ((nc - 1) // k + 1)))


def two_level(A, nc, k, dofs=6, A0_from=None):
    """Block-Jacobi (6x6) plus an additive coarse correction Z (Z^T A Z)^-1 Z^T; A0_from: build the coarse matrix from
    another (earlier) system matrix - a lagged coarse operator."""
    fine = cluster_precond(A, nc, 1)
    Z = coarse_space(nc, k, dofs)
    A0 = spl.splu((Z.T @ (A if A0_from is None else A0_from) @ Z).tocsc())
    return lambda r: fine(r) + Z @ A0.solve(Z.T @ r)


def band_precond(A, nc, bw):
    coo = A.tocoo()
    keep = np.abs(coo.row // 6 - coo.col // 6) <= bw
    B = sp.csc_matrix((coo.data[keep], (coo.row[keep], coo.col[keep])), shape=A.shape)
    lu = spl.splu(B)
    return lu.solve


fp = S_.make_flat_problem(n_cams, n_pts, n_obs, seed=1004)
prob = O.Problem.from_flat(fp)
rvec, tvec, pts = fp.rvec.copy(), fp.tvec.copy(), fp.pts.copy()
print(f"{n_cams} cameras / {n_pts} points / {fp.n_obs} observations")
lam = 1e-4
n_lm = int(os.environ.get("LM_ITERS", "6"))
inexact = os.environ.get("INEXACT", "0") == "1"     # take the device's steps: PCG with SJ, tol 0.1, at most 100 iterations
every = int(os.environ.get("EVERY", "1"))
for lm_it in range(n_lm):
    lin = L.linearize(prob, rvec, tvec, pts)
    t0 = time.perf_counter()
    A, b = build_S(prob, lin, lam)
    nc = prob.n_cams
    bw = int(np.abs(A.tocoo().row // 6 - A.tocoo().col // 6).max())
    res = {}
    sj = cluster_precond(A, nc, 1)
    res["SJ"], x_sj = pcg(A, b, sj, maxit=100 if inexact else 1000, want_x=True)
    if os.environ.get("STEP", "sj") == "two_level" and lm_it >= 5:
        # the planned device behaviour: Schur-Jacobi while the damping falls, then the two-level preconditioner with the
        # coarse matrix of THAT iteration kept to the end; the LM step is this solve's result
        if "A_step" not in globals():
            A_step = A.copy()
        res["step: SJ+C-32 lagged"], x_sj = pcg(A, b, two_level(A, nc, 32, A0_from=A_step), maxit=100, want_x=True)
    if lm_it % every:
        res = {k: v for k, v in res.items() if k == "SJ" or k.startswith("step")}
    else:
      if True:
        if os.environ.get("CLUSTERS", "0") == "1":
            for k in (4, 8, 16):
                res[f"CJ-{k}"] = pcg(A, b, cluster_precond(A, nc, k))
            for k in (8, 16):
                res[f"AS-{k}"] = pcg(A, b, cluster_precond(A, nc, k, overlap=k // 2))
        for k in (8, 16):
            res[f"CJ-{k}+C"] = pcg(A, b, cluster_precond(A, nc, k, coarse=True))
        res["BT"] = pcg(A, b, band_precond(A, nc, 1))
        res["B8"] = pcg(A, b, band_precond(A, nc, 8))
        for k in (8, 16, 32, 64):
            res[f"SJ+C-{k}"] = pcg(A, b, two_level(A, nc, k))
        res["SJ+Ct-16"] = pcg(A, b, two_level(A, nc, 16, dofs=3))
        if lm_it >= 5 and "A_lag" not in globals():
            A_lag = A.copy()
        if "A_lag" in globals():
            res["SJ+C-16 lagged"] = pcg(A, b, two_level(A, nc, 16, A0_from=A_lag))
            res["SJ+C-32 lagged"] = pcg(A, b, two_level(A, nc, 32, A0_from=A_lag))
    print(f"LM it {lm_it} cost {lin.cost:.6e} lambda {lam:.1e} camera bandwidth {bw} nnz blocks {A.nnz // 36}: "
          + "  ".join(f"{k} {v}" for k, v in res.items()) + f"   ({time.perf_counter() - t0:.0f} s)", flush=True)
    # exact LM step (sparse LU), accepted unconditionally; damping shrinks like after a good step
    dc = (x_sj if inexact else spl.splu(A.tocsc()).solve(b)).reshape(nc, 6)
    adj = prob.obs_cam >= 0
    Dp = np.clip(np.einsum("nii->ni", lin.V), 1e-6, 1e32)
    Vinv = np.linalg.inv(lin.V + lam * np.einsum("ni,ij->nij", Dp, np.eye(3)))
    E = np.einsum("nki,nk,nkj->nij", lin.Jc[adj], lin.wH[adj], lin.Jp[adj])
    w = lin.gp.copy()
    np.add.at(w, prob.obs_pt[adj], np.einsum("nij,ni->nj", E, dc[prob.obs_cam[adj]]))
    dp = -np.einsum("nij,nj->ni", Vinv, w)
    rvec, tvec, pts = L.retract(rvec, tvec, pts, dc, dp)
    lam = max(lam / 3.0, 1e-6)
