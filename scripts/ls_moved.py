"""How many points the per-point line search moves (factor != 1) per LM iteration (experiment build, see window_stamps.py).

    python scripts/ls_moved.py venice
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bundle_adjustment_b200 import _capi  # noqa: E402

_capi.LIB_PATH = os.path.join(ROOT, "scripts", "exp", "libba_stamps.so")
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
fp = S.make_named_problem(shape, keep_truth=False)
s = FlatSolver(0)
s.set_flat(fp)
out = (C.c_ulonglong * 8)()
prev = [0] * 6
print("factors 1, 1/2, 2, 4, 8, 16: points choosing each, per LM iteration")
for it in range(1, 31):
    s.restore_initial()
    _capi.lib().ba_ls_moved(out)
    st = s.solve(max_iters=it)
    _capi.lib().ba_ls_moved(out)
    cum = list(out[:6])
    print(f"  iteration {it:2d}: " + " ".join(f"{c - p:8d}" for c, p in zip(cum, prev)) + f"   ({st['status_text']})")
    prev = cum
    if st["iterations"] < it:
        break
