#!/usr/bin/env python
"""Benchmark of the local/global BA hot path (BASELINE.json metric: LM iterations/s and observations/s
on the 5 M-observation Venice-shaped problem, next to the reference's SciPy path on the host cores).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload venice]

A "step" is one complete device solve (``ba_solve``: LM iterations until ftol/xtol/gtol or 50 iterations)
of the workload from the same initial guess.  ``value`` = observations x LM iterations / second with the
problem resident in HBM (device-side parameter reset between steps); ``e2e`` = the same through the
C ABI with HOST buffers: ``ba_set_problem`` (H2D + device-side sort) + ``ba_solve`` + ``ba_get_params``
(D2H) inside the timed region.  N > 1: one process per GPU (torchrun), points sharded by rank, cameras
replicated, NCCL all-reduce of camera-sized sums inside the library; weak scaling (what is sharded grows
with N: N x 993 k points / N x 5 M observations; the 1778 replicated cameras stay), `--scaling strong` keeps
the N = 1 problem.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "ba_observations_x_lm_iterations_per_second"
UNIT = "obs*iter/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="venice", choices=["bal50", "ladybug", "venice", "shard50m"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--cpu-sample-obs", type=int, default=1_000_000, help="observations in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """nvidia-smi clock / throttle-reason samples every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self._stop_evt = index, [], threading.Event()

    def run(self):
        try:
            proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                     "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE, text=True)
        except OSError:
            return
        try:
            while not self._stop_evt.is_set():
                line = proc.stdout.readline()
                if not line:
                    break
                self.samples.append([x.strip() for x in line.split(",")])
        finally:
            proc.kill()

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        sm = [float(s[0]) for s in self.samples if s and s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if len(s) > 1 and s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for s in self.samples if len(s) >= 6 for i in range(4) if s[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def make_workload(name, n_gpus, scaling):
    """Weak scaling grows what is sharded - points and their observations - with the number of GPUs and keeps
    the cameras (replicated on every rank) fixed, so the work per GPU and the size of the reduced camera
    system stay what they are at N = 1.  Strong scaling solves the N = 1 problem on N GPUs."""
    from bundle_adjustment_b200 import synthetic as S
    n_cams, n_pts, n_obs = S.SHAPES[name]
    mult = n_gpus if scaling == "weak" else 1
    fp = S.make_flat_problem(n_cams, n_pts * mult, n_obs * mult, S.SEEDS[name], keep_truth=False)
    label = f"{name}-shape synthetic: {n_cams} cameras / {n_pts * mult} points / {n_obs * mult} observations"
    return fp, label


# ------------------------------------------------------------------------------------------------
# CPU legs (the only places this file touches oracle/)
# ------------------------------------------------------------------------------------------------
def cpu_reference_solve(sample_obs, workload):
    """One run of the reference's least_squares call (vectorised residual, tier T1) on a bounded sample
    of the workload's shape.  -> (obs*iterations/s, description, nfev)"""
    from bundle_adjustment_b200 import synthetic as S
    from oracle import ba_oracle as O
    n_cams, n_pts, n_obs = S.SHAPES[workload]
    scale = min(1.0, sample_obs / n_obs)
    fp = S.make_named_problem(workload, scale=scale, keep_truth=False)
    prob = O.Problem.from_flat(fp)
    t0 = time.perf_counter()
    res = O.solve_reference(prob, fp.x0())
    dt = time.perf_counter() - t0
    desc = (f"scipy.optimize.least_squares(jac_sparsity, loss='huber', xtol=ftol=1e-5, max_nfev=50) with the "
            f"vectorised residual on a {fp.n_cams + 1}-camera / {fp.n_pts}-point / {fp.n_obs}-observation sample of the "
            f"{workload} shape: {res.nfev} evaluations in {dt:.1f} s (status {res.status})")
    return fp.n_obs * res.nfev / dt, desc, res.nfev, dt


def run_reference(args, rank):
    """--impl reference: the reference's CPU path (oracle port, tier T1) on the host cores, bounded sample."""
    if rank != 0:
        return
    vals, desc, nfev, dts = [], "", 0, []
    for i in range(args.warmup + args.steps):
        v, desc, nfev, dt = cpu_reference_solve(args.cpu_sample_obs, args.workload)
        if i >= args.warmup:
            vals.append(v)
            dts.append(dt)
    value = float(np.mean(vals))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(dts)), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.workload}-shape, bounded CPU sample ({args.cpu_sample_obs} observations)",
                   "solver": "scipy least_squares trf/lsmr, 2-point FD over jac_sparsity, huber"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port", "sample": desc,
                         "host_cpus": os.cpu_count()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "lm_iters_per_step": nfev,
    }
    print(json.dumps(line))


def algorithmic_bytes(kernel, n_obs, n_obs_adj, n_pts, n_cams, n_seg=0, n_rows=None):
    """Bytes one launch must move.  Tile passes stream the Jacobian cache (48 B) and the
    observation record (8 B) per observation; two-pass / linearisation kernels recompute from the 24 B
    (point-major) or 20 B (camera-major) observation record.  Each point / camera record is counted once."""
    n_rows = n_seg if n_rows is None else n_rows
    if kernel == "schur_matvec_tiles":       # per point X (32) + Vinv (48) + track offset (4); camera vector; 8 B table entry
        return 56 * n_obs + 84 * n_pts + 48 * (n_cams + 1) + 8 * n_seg + 48 * n_rows   # per segment; one 48 B row per accumulator slot
    if kernel == "pcg_tail":                 # rows read; u, p, s, x, r, D, w streamed once or twice; Minv (288 B per camera)
        return 48 * n_rows + (48 * 12 + 288) * n_cams
    if kernel == "backsub_tiles":            # + h, g, D read (32 each), X_trial written (32)
        return 56 * n_obs + (84 + 96 + 32) * n_pts + 48 * (n_cams + 1)
    if kernel == "rows_finish":
        return 48 * n_rows + 144 * n_cams
    if kernel == "schur_matvec_points":      # uv+idx; per point X (32) + Vinv (48) read, z (32) written; camera R|t|vec
        return 24 * n_obs + 112 * n_pts + 144 * (n_cams + 1)
    if kernel == "schur_matvec_cameras":     # uv+pt; per point X (32) + z (32) read; camera R|t|vec; 6-double partial per tile
        return 20 * n_obs_adj + 64 * n_pts + 144 * n_cams
    if kernel == "lin_points":               # uv+idx; X read, V (48) + g (32) written; camera R|t; Jacobian cache written
        return (24 + 4 + 48) * n_obs + 112 * n_pts + 96 * (n_cams + 1)
    if kernel == "lin_cameras":
        return 20 * n_obs_adj + 32 * n_pts + 96 * n_cams
    if kernel == "schur_rhs_cameras":        # + h (32) + Vinv (48) per point
        return 20 * n_obs_adj + 112 * n_pts + 96 * n_cams
    if kernel == "backsub_points":           # as the mat-vec point pass + h, g, D read and X_trial written
        return 24 * n_obs + 208 * n_pts + 144 * (n_cams + 1)
    return None


def algorithmic_flops(kernel, n_obs, n_pts):
    """FP64 flops one launch must do (FMA = 2).  Tile mat-vec per observation: d = w x X + v (12), a = Jt d (10),
    Jt^T a (12), track sum (3), b = a - Jt z (12), m = Jt^T b (12), X x m (9), segment sums (6) = 76; per point
    z = Vinv s (15)."""
    if kernel == "schur_matvec_tiles":
        return 76 * n_obs + 15 * n_pts
    if kernel == "backsub_tiles":
        return 43 * n_obs + 40 * n_pts
    return None


def run_ours(args, rank, world, local_rank):
    # libraries (NCCL) may write to fd 1; keep the JSON line the only thing on stdout
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    from bundle_adjustment_b200 import FlatSolver, sharding

    dist = None
    nccl_id = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.tensor(list(FlatSolver.nccl_unique_id()), dtype=torch.uint8, device="cuda")
        dist.broadcast(idt, 0)
        nccl_id = bytes(idt.cpu().tolist())

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    fp_full, label = make_workload(args.workload, world, args.scaling)
    n_obs_total, n_pts_total, n_cams = fp_full.n_obs, fp_full.n_pts, fp_full.n_cams
    fp = fp_full if world == 1 else sharding.shard_problem(fp_full, rank, world)[0]
    del fp_full

    # pinned host copies: what the e2e leg uploads every step
    def pinned(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        return t, t.numpy()
    keep, host = [], {}
    for k in ("rvec", "tvec", "pts", "obs_cam", "obs_pt", "obs_uv"):
        t, host[k] = pinned(getattr(fp, k))
        keep.append(t)
    h2d = sum(host[k].nbytes for k in host) + 48
    d2h = host["rvec"].nbytes + host["tvec"].nbytes + host["pts"].nbytes

    solver = FlatSolver(local_rank, rank=rank, nranks=world, nccl_id=nccl_id)

    def upload():
        solver.set_problem(fp.fx, fp.fy, fp.cx, fp.cy, host["rvec"], host["tvec"], fp.fixed_rvec, fp.fixed_tvec,
                           host["pts"], host["obs_cam"], host["obs_pt"], host["obs_uv"])

    upload()
    opts = dict()
    stats = None
    for _ in range(max(args.warmup, 3)):
        solver.restore_initial()
        stats = solver.solve(**opts)

    # ---- timed region 1: device-resident ---------------------------------------------------------
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    barrier()
    solver.timer_start()
    launches = 0
    for _ in range(args.steps):
        solver.restore_initial()
        stats = solver.solve(**opts)
        launches += stats["launches"] + 2
    ms = solver.timer_stop()
    barrier()
    ms = max_over_ranks(ms)
    clocks = sampler.stop() if sampler else None
    iters = stats["iterations"]
    value = n_obs_total * iters * args.steps / (ms * 1e-3)

    # ---- timed region 2: end to end through the C ABI with host buffers ---------------------------
    for _ in range(1):
        upload(); solver.solve(**opts); solver.get_params()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        upload()
        st_e = solver.solve(**opts)
        solver.get_params()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e_value = n_obs_total * st_e["iterations"] * args.steps / e2e_s

    # ---- per-kernel device times (pooled CUDA events on the solver stream, same solve) --------------
    solver.restore_initial()
    # pcg_check_every=1: the host looks at the convergence flag after every PCG iteration, so no launch of this
    # solve is an early exit and launches == iterations (same arithmetic, same iteration counts)
    st_p = solver.solve(verbose=2, **{**opts, "pcg_check_every": 1})
    ks = solver.kernel_stats()
    total_kernel_ms = sum(v[1] for v in ks.values())
    top = max((k for k in ks if algorithmic_bytes(k, 1, 1, 1, 1, 1, 1) is not None), key=lambda k: ks[k][1])
    n_obs_adj = int((fp.obs_cam >= 0).sum())
    layout = solver.layout_info()
    n_seg = layout["segments"]
    peak, peak_src = peaks()
    n_rows = layout["matvec_rows"]
    ab = algorithmic_bytes(top, fp.n_obs, n_obs_adj, fp.n_pts, fp.n_cams, n_seg, n_rows)
    avg_ms = ks[top][1] / max(ks[top][0], 1)
    achieved = ab / (avg_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.isfile(tpath) and world == 1:
        with open(tpath) as f:
            traffic = json.load(f).get(args.workload, {}).get(top)
    # FP64 pipe (north_star: reported beside the HBM fraction): measured DFMA rate of this device against the
    # arithmetic of the dominant kernel (FMA = 2 flops)
    fp64_peak = solver.fp64_peak()
    fl = algorithmic_flops(top, fp.n_obs, fp.n_pts)
    fp64 = {"peak_tflops": fp64_peak, "kernel": top, "flops_per_launch": fl,
            "achieved_tflops": (fl / (avg_ms * 1e-3) / 1e12) if fl else None,
            "util": (fl / (avg_ms * 1e-3) / 1e12 / fp64_peak) if fl and fp64_peak > 0 else None,
            "how": "ba_fp64_peak: 16 independent DFMA chains per thread, 8 CTAs x 256 threads per SM, best of 3 (CUDA events)"}
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "frac_of_nominal_8000": achieved / 8000.0, "fp64": fp64,
                "traffic": traffic, "algorithmic_bytes_per_launch": ab, "avg_launch_us": 1e3 * avg_ms,
                "launches_per_solve": ks[top][0], "share_of_kernel_time": ks[top][1] / max(total_kernel_ms, 1e-9),
                "peak_source": peak_src, "layout": layout,
                "kernels": {k: {"launches": v[0], "ms": round(v[1], 3),
                                "GBps": (algorithmic_bytes(k, fp.n_obs, n_obs_adj, fp.n_pts, fp.n_cams, n_seg, n_rows) or 0) * v[0] / max(v[1], 1e-9) / 1e6}
                            for k, v in ks.items() if v[0] and v[1] > 0.0}}

    if rank != 0:
        solver.close()
        if dist is not None:
            dist.barrier(); dist.destroy_process_group()
        return

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        v, desc, nfev, dt = cpu_reference_solve(args.cpu_sample_obs, args.workload)
        cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": "port", "sample": desc, "host_cpus": os.cpu_count()}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": label, "step": "one ba_solve (LM to ftol 1e-6 / 50 iterations) from the same initial guess",
                   "parallelism": f"points sharded over {world} GPU(s), cameras replicated",
                   "l2": "inputs larger than L2 (observation streams 2 x 24 B x n_obs + 128 B point records)",
                   "robust": "huber 1 px per coordinate (IRLS)", "pcg": "Schur-Jacobi, tol 0.1"},
        "lm_iters_per_step": iters, "lm_iters_per_s": iters * args.steps / (ms * 1e-3),
        "pcg_iters_per_step": stats["pcg_iterations"], "final_cost": stats["final_cost"], "initial_cost": stats["initial_cost"],
        "status": stats["status_text"],
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": 1e3 * e2e_s / args.steps},
        "gpu_launches": int(launches),
        "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
    }
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    print(json.dumps(line), flush=True)
    os.dup2(2, 1)
    solver.close()
    if dist is not None:
        dist.barrier(); dist.destroy_process_group()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit(f"--gpus {args.gpus} needs torchrun: python -m torch.distributed.run --nnodes=1 "
                         f"--nproc-per-node {args.gpus} --master-addr 127.0.0.1 --master-port 29500 bench.py --gpus {args.gpus}")
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
