"""GPU debug: compare the two Schur mat-vec paths entry by entry."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bundle_adjustment_b200 import FlatSolver, synthetic as S
from oracle import ba_oracle as O, lm_schur as L
fp = S.make_flat_problem(6, 300, 1200, seed=3)
prob = O.Problem.from_flat(fp)
lin = L.linearize(prob, fp.rvec, fp.tvec, fp.pts)
u = np.random.default_rng(0).normal(size=(fp.n_cams, 6))
np.set_printoptions(linewidth=200, precision=6)
with FlatSolver(0) as s:
    s.set_flat(fp)
    for lam in (1e-3, 1e9):
        _, _, _, aux = L.schur_step(prob, lin, lam)
        w_ref = (aux["S"] @ u.ravel()).reshape(-1, 6)
        a = s.schur_matvec(u, lam, schur_path="tiles")
        b = s.schur_matvec(u, lam, schur_path="two_pass")
        print("lam", lam, "tiles err", np.abs(a-w_ref).max()/np.abs(w_ref).max(), "two_pass err", np.abs(b-w_ref).max()/np.abs(w_ref).max())
        Uu = np.einsum("nij,nj->ni", lin.U, u) + lam*aux["Dc"]*u
        print(" tiles vs Uu", np.abs(a-Uu).max()/np.abs(Uu).max())
        print((a-w_ref)/np.abs(w_ref).max())
