import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bundle_adjustment_b200 import FlatSolver, synthetic as S
fp = S.make_flat_problem(6, 300, 1200, seed=3)
np.set_printoptions(linewidth=220)
with FlatSolver(0) as s:
    s.set_flat(fp)
    info = s.layout_info(); print(info)
    L = s.tile_layout()
    for k in ("ft_pt", "ft_obs", "ft_seg", "seg_cam", "seg_start", "seg_row", "cam_row_off"):
        print(k, L[k])
    w = L["tc_word"]
    print("pl", (w & 1023)[:40]); print("pos", ((w >> 10) & 1023)[:40]); print("sl", (w >> 20)[:40])
    # checks
    T = info["point_tiles"]
    for t in range(T):
        o0, o1 = L["ft_obs"][t], L["ft_obs"][t+1]
        pos = np.sort(((w[o0:o1] >> 10) & 1023))
        assert np.array_equal(pos, np.arange(o1 - o0)), ("pos not a permutation", t)
        sl = (w[o0:o1] >> 20).astype(int)
        assert np.all(np.diff(sl) >= 0) and sl[0] == 0 and sl[-1] == L["ft_seg"][t+1]-L["ft_seg"][t]-1, ("sl", t, sl)
    print("ok")
    from oracle import ba_oracle as O, lm_schur as Lm

