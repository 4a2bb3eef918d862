"""Point sharding for the multi-GPU solve.

Points are split into ``nranks`` contiguous ranges balanced by observation count; every rank keeps all
cameras.  Each point's 3x3 block, its observations and its Schur contribution stay on one GPU, so the
only cross-GPU traffic is the all-reduce of camera-sized sums inside ``libba_b200.so``.
"""
from __future__ import annotations

import numpy as np


def partition_points(obs_pt: np.ndarray, n_pts: int, nranks: int) -> np.ndarray:
    """Boundaries b[0..nranks] (b[0] = 0, b[-1] = n_pts): rank r owns points [b[r], b[r+1]).
    Chosen so every rank holds close to n_obs / nranks observations."""
    if nranks < 1:
        raise ValueError("nranks must be >= 1")
    if n_pts < nranks:
        # the same ValueError on every rank (each rank computes the partition of the whole problem): an empty shard would fail
        # ba_set_problem on one rank only
        raise ValueError(f"cannot shard {n_pts} points over {nranks} ranks: every rank needs at least one point")
    counts = np.bincount(np.asarray(obs_pt, dtype=np.int64), minlength=n_pts)
    csum = np.concatenate([[0], np.cumsum(counts)])
    total = int(csum[-1])
    targets = (np.arange(1, nranks, dtype=np.float64) * total / nranks)
    cuts = np.searchsorted(csum, targets, side="left")
    bounds = np.concatenate([[0], np.clip(cuts, 0, n_pts), [n_pts]]).astype(np.int64)
    bounds = np.maximum.accumulate(bounds)
    # at least one point per rank (skewed track lengths can put two cuts on the same point)
    for r in range(1, nranks):
        bounds[r] = max(bounds[r], bounds[r - 1] + 1)
    for r in range(nranks - 1, 0, -1):
        bounds[r] = min(bounds[r], bounds[r + 1] - 1)
    return bounds


def shard_problem(fp, rank: int, nranks: int):
    """The ``FlatProblem`` slice rank ``rank`` uploads: its points (renumbered from 0), their observations
    (original relative order kept), all cameras."""
    b = partition_points(fp.obs_pt, fp.n_pts, nranks)
    return fp.point_slice(int(b[rank]), int(b[rank + 1])), (int(b[rank]), int(b[rank + 1]))
