import os, sys
sys.path.insert(0, "/root/repo")
from bundle_adjustment_b200 import FlatSolver, synthetic as S
fp = S.make_flat_problem(5, 2800, 6000, seed=S.SEEDS["window"])
s = FlatSolver(0)
s.set_flat(fp)
st = s.solve(device_loop=0, max_iters=6)
print(st["iterations"], st["solve_ms"])
