"""One LM iteration on a BASELINE shape with the coarse matrix built at once (for ncu on k_coarse_direct / k_spd_inverse6)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S
fp = S.make_named_problem(sys.argv[1] if len(sys.argv) > 1 else "venice", keep_truth=False)
s = FlatSolver(0)
s.set_flat(fp)
for rep in range(2):
    s.restore_initial()
    st = s.solve(max_iters=2, precond="two_level", coarse_after=-1, pcg_max_iters=8, verbose=2)
ks = s.kernel_stats()
print({k: (n, round(1e3 * ms / max(n, 1), 1)) for k, (n, ms) in ks.items() if n})
print(st["coarse_builds"], st["pcg_iterations"], st["final_cost"])
