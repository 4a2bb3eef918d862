"""One rank of the multi-GPU test (launched by tests/test_gpu_multi.py through torch.distributed.run)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bundle_adjustment_b200 import FlatSolver, sharding, synthetic as S  # noqa: E402


PROBLEMS = {"chain60": ((60, 20000, 100000, 21), dict(max_iters=8)),
            # long camera chain, two-level preconditioner from the first LM iteration on (coarse matrix summed over the ranks); the
            # damping starts at its floor and the PCG runs to 1e-8, so that the LM steps are exact and do not hinge on a stopping
            # threshold crossed one iteration earlier or later
            "two_level": ((300, 12000, 60000, 31), dict(max_iters=4, precond="two_level", coarse_after=-1, ftol=0.0, xtol=0.0, gtol=0.0,
                                                        initial_lambda=1e-6, pcg_tol=1e-8, pcg_max_iters=2000))}


def main(out_path, pcg_tail, which="chain60"):
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idt = torch.tensor(list(FlatSolver.nccl_unique_id()), dtype=torch.uint8, device="cuda")
    dist.broadcast(idt, 0)
    (nc, npt, nobs, seed), kw = PROBLEMS[which]
    fp = S.make_flat_problem(nc, npt, nobs, seed=seed)
    shard, (lo, hi) = sharding.shard_problem(fp, rank, world)
    with FlatSolver(local, rank=rank, nranks=world, nccl_id=bytes(idt.cpu().tolist())) as s:
        s.set_flat(shard)
        st = s.solve(pcg_tail=pcg_tail, **kw)
        rvec, tvec, pts = s.get_params()
    np.savez(f"{out_path}.rank{rank}.npz", rvec=rvec, tvec=tvec, pts=pts, lo=lo, hi=hi, cost=st["final_cost"],
             pcg=st["pcg_iterations"], iters=st["iterations"], status=st["status"], precond=st["precond_used"],
             coarse_builds=st["coarse_builds"])
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]), *(sys.argv[3:4]))
