"""Shared-memory wavefront model of the point-indexed accesses of the tile pass (phases A and C) on a dumped
tile layout (scripts/dump_tiles.py), and what a conflict-aware order of the observations inside a camera run
would buy.  CPU only.

    python scripts/bank_model.py gpurun_out/tiles_venice.npz [tiles]
"""
import sys
from collections import Counter

import numpy as np

d = np.load(sys.argv[1])
n_tiles = int(sys.argv[2]) if len(sys.argv) > 2 else 400
ft_obs, ft_pt, meta = d["ft_obs"], d["ft_pt"], d["tc_meta"]


def wf_groups(keys, addrs, group):
    """Sum over groups of `group` lanes of the largest number of distinct addresses that share a bank class."""
    tot = 0
    for g0 in range(0, len(keys), group):
        cls = {}
        for k, a in zip(keys[g0:g0 + group], addrs[g0:g0 + group]):
            cls.setdefault(k, set()).add(a)
        tot += max(len(v) for v in cls.values())
    return tot


def cost(pl, pos, p0):
    n = len(pl)
    c = {}
    c["X01/z01 16B (quarter)"] = 2 * wf_groups(pl % 8, pl, 8)
    c["X2/z2 8B (half)"] = wf_groups((pl + (p0 & 1)) % 16, pl, 16) + wf_groups(pl % 16, pl, 16)
    c["sA scatter 3x8B (half)"] = 3 * wf_groups(pos % 16, pos, 16)
    return c


def ideal(n):
    q, h = -(-n // 8), -(-n // 16)
    return {"X01/z01 16B (quarter)": 2 * q, "X2/z2 8B (half)": 2 * h, "sA scatter 3x8B (half)": 3 * h}


def greedy(cam, pl, pos, p0):
    """Re-place the observations of every camera run inside the run's slots: slot by slot take the remaining
    observation of the run that collides least with what the slot's quarter / half warp already holds."""
    n = len(cam)
    out = np.arange(n)
    remaining = {}
    for i in range(n):
        remaining.setdefault(cam[i], []).append(i)
    half_pl, half_pos, quarter_pl = Counter(), Counter(), Counter()
    for s in range(n):
        if s % 16 == 0:
            half_pl.clear(); half_pos.clear()
        if s % 8 == 0:
            quarter_pl.clear()
        cand = remaining[cam[s]]
        best, bc = None, None
        for j in cand:
            cst = 3 * half_pos[pos[j] % 16] + 2 * half_pl[pl[j] % 16] + 2 * quarter_pl[pl[j] % 8]
            if bc is None or cst < bc:
                best, bc = j, cst
                if cst == 0:
                    break
        cand.remove(best)
        out[s] = best
        half_pl[pl[best] % 16] += 1; half_pos[pos[best] % 16] += 1; quarter_pl[pl[best] % 8] += 1
    return out


tot0, tot1, toti = Counter(), Counter(), Counter()
nobs = 0
for t in range(0, min(n_tiles, len(ft_obs) - 1)):
    o0, o1 = ft_obs[t], ft_obs[t + 1]
    if o1 <= o0:
        continue
    cam = meta[o0:o1, 0]
    pl = meta[o0:o1, 1] & 1023
    pos = (meta[o0:o1, 1] >> 10) & 1023
    p0 = int(ft_pt[t])
    nobs += o1 - o0
    tot0.update(cost(pl, pos, p0))
    toti.update(ideal(o1 - o0))
    perm = greedy(cam, pl, pos, p0)
    tot1.update(cost(pl[perm], pos[perm], p0))
print(f"{nobs} observations in {n_tiles} tiles; wavefronts per 32 observations")
for k in tot0:
    print(f"  {k:28s} now {32 * tot0[k] / nobs:6.2f}   greedy {32 * tot1[k] / nobs:6.2f}   ideal {32 * toti[k] / nobs:6.2f}")
s0, s1, si = sum(tot0.values()), sum(tot1.values()), sum(toti.values())
print(f"  {'sum':28s} now {32 * s0 / nobs:6.2f}   greedy {32 * s1 / nobs:6.2f}   ideal {32 * si / nobs:6.2f}")
