"""Copy the tile structure of the first tiles of a BASELINE shape to gpurun_out/ (input of scripts/bank_model.py)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bundle_adjustment_b200 import FlatSolver, synthetic as S  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
n_tiles = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
fp = S.make_named_problem(shape, keep_truth=False)
s = FlatSolver(0)
s.set_flat(fp)
L = s.tile_layout()
T = min(n_tiles, len(L["ft_pt"]) - 1)
o1, s1 = L["ft_obs"][T], L["ft_seg"][T]
os.makedirs("gpurun_out", exist_ok=True)
np.savez_compressed(f"gpurun_out/tiles_{shape}.npz", ft_pt=L["ft_pt"][:T + 1], ft_obs=L["ft_obs"][:T + 1], ft_seg=L["ft_seg"][:T + 1],
                    tc_meta=L["tc_meta"][:o1], seg_info=L["seg_info"][:s1], seg_cam=L["seg_cam"][:s1])
print(s.layout_info(), T, o1, s1)
