import sys, time, os
sys.path.insert(0, os.getcwd())
import numpy as np
from bundle_adjustment_b200 import FlatSolver, synthetic as S
from oracle import ba_oracle as O, lm_schur as L
fp = S.make_flat_problem(6, 300, 1200, seed=3)
P = O.Problem.from_flat(fp)
s = FlatSolver(0)
s.set_flat(fp)
x0 = fp.x0()
r = s.residuals(); r0 = P.residuals(x0)
print("resid rel", np.linalg.norm(r-r0)/np.linalg.norm(r0))
jc, jp = s.jacobian_blocks(); Jc, Jp = P.jacobian_blocks(x0)
print("Jc rel", np.abs(jc-Jc).max()/np.abs(Jc).max(), "Jp rel", np.abs(jp-Jp).max()/np.abs(Jp).max())
raw, rob = s.cost()
print("cost", raw, O.raw_cost(r0), rob, O.huber_cost(r0))
nb = s.normal_blocks()
lin = L.linearize(P, fp.rvec, fp.tvec, fp.pts)
for k, ref in [("gc", lin.gc), ("gp", lin.gp), ("U", lin.U), ("V", lin.V)]:
    print(k, np.abs(nb[k]-ref).max()/np.abs(ref).max())
ip, ii, shp = s.sparsity_csr(); A = O.sparsity_csr(P.n_cams, P.n_pts, P.obs_cam, P.obs_pt)
print("sparsity", (ip==A.indptr).all(), (ii==A.indices).all(), shp==A.shape)
st = s.solve(max_iters=5, verbose=1, pcg_tol=1e-10, pcg_max_iters=200, ftol=0, xtol=0)
print(st)
res = L.solve(P, fp.rvec, fp.tvec, fp.pts, max_iters=5, ftol=0)
print([ (h['cost_new'], h['accepted']) for h in res.history])
x = s.get_x(); print("x vs oracle LM", np.abs(x - res.x()).max())
st = s.solve(max_iters=200, verbose=0)
print(st)
for name in ["bal50", "ladybug", "venice"]:
    fp = S.make_named_problem(name, keep_truth=False)
    t = time.time(); s.set_flat(fp); t1 = time.time()-t
    st = s.solve(max_iters=30, verbose=1 if name=="bal50" else 0)
    print(name, "upload+sort s", round(t1,3), st)
    s2 = s.solve(max_iters=3, verbose=2)
    ks = s.kernel_stats()
    for k,(n,ms) in ks.items():
        if n: print(f"   {k:26s} n={n:5d} ms={ms:9.3f} avg_us={1e3*ms/n:9.2f}")
