"""Two GPUs of one box: the point-sharded solve with the camera partials exchanged through peer memory inside the PCG
tail (csrc/ba_peer.cuh) against the single-GPU solve of the same problem.  Skipped on a box with one GPU."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from bundle_adjustment_b200 import FlatSolver, synthetic as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _gpus():
    import torch
    return torch.cuda.device_count()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run_two_ranks(tmp_path, tag, env_extra, pcg_tail=1, which="chain60"):
    out = str(tmp_path / tag)
    env = dict(os.environ, **env_extra)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "_multi_gpu_worker.py"), out, str(pcg_tail), which]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=int(os.environ.get("BA_TEST_RANK_TIMEOUT", "120")))
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    return [np.load(f"{out}.rank{r}.npz") for r in range(2)]


def _single_gpu():
    fp = S.make_flat_problem(60, 20000, 100000, seed=21)
    with FlatSolver(0) as s:
        s.set_flat(fp)
        st = s.solve(max_iters=8)
        rvec, tvec, pts = s.get_params()
    return st, rvec, tvec, pts


def test_two_ranks_match_one_gpu_and_the_protocols_agree_bitwise(tmp_path):
    if _gpus() < 2:
        pytest.skip("needs two GPUs")
    st1, rvec1, tvec1, pts1 = _single_gpu()
    variants = {"ll": ({"BA_B200_XCHG": "ll"}, 1), "push": ({"BA_B200_XCHG": "push"}, 1), "pull": ({"BA_B200_XCHG": "pull"}, 1),
                "kernels": ({"BA_B200_XCHG": "ll"}, 0),        # rows_finish_xchg + pcg_dots_peer + pcg_update instead of the tail
                "kernels_push": ({"BA_B200_XCHG": "push"}, 0),
                "nccl": ({"BA_B200_NO_PEER": "1"}, 1)}
    want = os.environ.get("BA_TEST_XCHG_VARIANTS", "ll,push,pull,kernels").split(",")
    runs = {tag: _run_two_ranks(tmp_path, tag, *variants[tag]) for tag in want}
    r0, r1 = runs["ll"]
    # ranks agree exactly with each other (replicated cameras, identical sums in rank order)
    assert np.array_equal(r0["rvec"], r1["rvec"]) and np.array_equal(r0["tvec"], r1["tvec"])
    assert float(r0["cost"]) == float(r1["cost"]) and int(r0["pcg"]) == int(r1["pcg"])
    # ... and with the single-GPU solve up to the order of the per-camera sums
    assert abs(float(r0["cost"]) - st1["final_cost"]) <= 1e-8 * st1["final_cost"]
    assert int(r0["iters"]) == st1["iterations"]
    pts = np.concatenate([r0["pts"], r1["pts"]])
    assert int(r0["lo"]) == 0 and int(r1["hi"]) == len(pts1) and int(r0["hi"]) == int(r1["lo"])
    assert np.abs(pts - pts1).max() <= 1e-6 * np.abs(pts1).max()
    assert np.abs(r0["tvec"] - tvec1).max() <= 1e-6 * max(np.abs(tvec1).max(), 1.0)
    # the three peer-memory protocols move the same numbers: identical bits; the separate kernels too
    for tag in ("kernels", "kernels_push"):  # other row-sum order than the tail: same solution up to summation order
        if tag in runs:
            assert abs(float(runs[tag][0]["cost"]) - st1["final_cost"]) <= 1e-8 * st1["final_cost"], tag
            assert np.array_equal(runs[tag][0]["rvec"], runs[tag][1]["rvec"]), tag
    for tag in ("push", "pull"):
        if tag not in runs:
            continue
        for a, b in zip(runs["ll"], runs[tag]):
            assert np.array_equal(a["rvec"], b["rvec"]) and np.array_equal(a["pts"], b["pts"]), tag
            assert float(a["cost"]) == float(b["cost"]) and int(a["pcg"]) == int(b["pcg"]), tag
    # NCCL all-reduce per iteration (ring / tree order): same solution up to summation order
    if "nccl" in runs:
        assert abs(float(runs["nccl"][0]["cost"]) - st1["final_cost"]) <= 1e-8 * st1["final_cost"]


def test_two_level_preconditioner_on_two_ranks(tmp_path):
    """The coarse matrix is the sum of the ranks' probed parts (one all-reduce per build); the sharded solve must follow the
    single-GPU solve: same iteration counts, same cost up to summation order, ranks bit-identical."""
    if _gpus() < 2:
        pytest.skip("needs two GPUs")
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _multi_gpu_worker import PROBLEMS
    (nc, npt, nobs, seed), kw = PROBLEMS["two_level"]
    fp = S.make_flat_problem(nc, npt, nobs, seed=seed)
    with FlatSolver(0) as s:
        s.set_flat(fp)
        st1 = s.solve(**kw)
    assert st1["precond_text"] == "two_level" and st1["coarse_builds"] >= 1
    r0, r1 = _run_two_ranks(tmp_path, "two_level", {}, 1, "two_level")
    assert int(r0["precond"]) == 2 and int(r0["coarse_builds"]) == st1["coarse_builds"]
    assert np.array_equal(r0["rvec"], r1["rvec"]) and float(r0["cost"]) == float(r1["cost"])
    assert int(r0["iters"]) == st1["iterations"]
    assert abs(float(r0["cost"]) - st1["final_cost"]) <= 1e-9 * st1["final_cost"]
    assert abs(int(r0["pcg"]) - st1["pcg_iterations"]) <= 0.05 * st1["pcg_iterations"] + 2, (int(r0["pcg"]), st1["pcg_iterations"])
