"""Stage times inside k_dense_schur on a window-sized problem (experiment build with -DBA_DN_STAMPS, see scripts/README.md):

    nvcc ... -DBA_DN_STAMPS -o scripts/exp/libba_stamps.so bundle_adjustment_b200/csrc/ba_solver.cu -ldl
    python scripts/window_stamps.py
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bundle_adjustment_b200 import _capi  # noqa: E402

_capi.LIB_PATH = os.path.join(ROOT, "scripts", "exp", "libba_stamps.so")
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

fp = S.make_flat_problem(5, 2800, 6000, seed=S.SEEDS["window"])
s = FlatSolver(0)
s.set_flat(fp)
names = ["start", "pairs done", "block_sum", "ticket (last CTA)", "extras", "assemble", "factorisation + solves", "write x", "(pose update: in the back-substitution launch)"]
rows = []
for it in (3, 5, 8, 12):
    s.restore_initial()
    s.solve(max_iters=it)
    out = (C.c_ulonglong * 16)()
    assert _capi.lib().ba_dn_stamps(out) == 0
    t = np.array(out[:9], dtype=np.int64)
    rows.append(np.diff(t))
    u = np.array(out[:12], dtype=np.int64)
    print(f"  inside the factorisation: columns {u[9] - u[5]} ns, forward {u[10] - u[9]} ns, backward {u[11] - u[10]} ns")
rows = np.array(rows)
print("ns per stage (four samples, %globaltimer; stamp 0..2 by CTA 0, 3.. by the last CTA):")
for i, n in enumerate(names[1:]):
    print(f"  {n:20s} " + " ".join(f"{v:7d}" for v in rows[:, i]))
print(f"  {'total':20s} " + " ".join(f"{v:7d}" for v in rows.sum(axis=1)))
