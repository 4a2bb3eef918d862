"""CPU, world_size 2 over gloo: the point-sharded reductions of the multi-GPU solve.

Each rank linearises only its shard (its points + their observations, all cameras) with the CPU
restatement of the device algorithm; camera-side sums are all-reduced exactly as ``libba_b200.so`` does
over NCCL (U, g_c, Schur right-hand side, mat-vec partials, cost scalars).  The result must equal the
unsharded computation up to summation order.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from bundle_adjustment_b200 import sharding, synthetic as S
from oracle import ba_oracle as O, lm_schur as L


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _camera_sums(fp, lam=1e-3):
    """Per-camera sums one rank contributes: U, g_c, E Vinv g_p, S-matvec partial for a fixed probe vector."""
    prob = O.Problem.from_flat(fp)
    lin = L.linearize(prob, fp.rvec, fp.tvec, fp.pts)
    adj = prob.obs_cam >= 0
    ca, pa = prob.obs_cam[adj], prob.obs_pt[adj]
    Dp = np.clip(np.einsum("nii->ni", lin.V), 1e-6, 1e32)
    Vinv = np.linalg.inv(lin.V + lam * np.einsum("ni,ij->nij", Dp, np.eye(3)))
    E = np.einsum("nki,nk,nkj->nij", lin.Jc[adj], lin.wH[adj], lin.Jp[adj])
    rhs = np.zeros((prob.n_cams, 6))
    np.add.at(rhs, ca, np.einsum("nij,njk,nk->ni", E, Vinv[pa], lin.gp[pa]))
    x = np.sin(np.arange(prob.n_cams * 6, dtype=np.float64)).reshape(prob.n_cams, 6)      # probe vector
    w = np.zeros((prob.n_pts, 3))
    np.add.at(w, pa, np.einsum("nij,ni->nj", E, x[ca]))
    z = np.einsum("nij,nj->ni", Vinv, w)
    y = np.zeros((prob.n_cams, 6))
    np.add.at(y, ca, -np.einsum("nij,nj->ni", E, z[pa]))         # the E Vinv E^T x part; U x is added from the summed U
    return lin.U, lin.gc, rhs, y, np.array([lin.cost, lin.raw])


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    fp = S.make_flat_problem(12, 900, 4000, seed=21)
    shard, (lo, hi) = sharding.shard_problem(fp, rank, world)
    parts = [torch.from_numpy(np.ascontiguousarray(a)) for a in _camera_sums(shard)]
    for t in parts:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    sizes = torch.tensor([shard.n_obs, hi - lo], dtype=torch.int64)
    dist.all_reduce(sizes)
    if rank == 0:
        np.savez(os.path.join(out_dir, "reduced.npz"), U=parts[0].numpy(), gc=parts[1].numpy(), rhs=parts[2].numpy(),
                 y=parts[3].numpy(), cost=parts[4].numpy(), sizes=sizes.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_sharded_camera_sums_equal_unsharded(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    red = np.load(tmp_path / "reduced.npz")
    fp = S.make_flat_problem(12, 900, 4000, seed=21)
    U, gc, rhs, y, cost = _camera_sums(fp)
    assert tuple(red["sizes"]) == (fp.n_obs, fp.n_pts)
    for name, full in (("U", U), ("gc", gc), ("rhs", rhs), ("y", y), ("cost", cost)):
        assert np.abs(red[name] - full).max() <= 1e-12 * np.abs(full).max(), name
