#!/usr/bin/env python
"""Benchmark of the local/global BA hot path (BASELINE.json metric: LM iterations/s and observations/s on the
5 M-observation Venice-shaped problem at 1/2/4/8 B200, next to the reference's SciPy path on the host cores).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload venice] [--scaling strong|weak]

A "step" is one complete solve of the workload from the generator's initial guess with the reference's own stopping rule
(``least_squares(..., loss='huber', xtol=1e-5, ftol=1e-5, max_nfev=50)``, src/bundle_adjuster.py:170-174): for our arm one
``ba_solve`` (LM iterations until ftol / xtol / gtol or 50 iterations), for the reference arm one SciPy call.
``value`` = observations x iterations / second with the problem resident in HBM (device-side parameter reset between steps);
``e2e`` = the same through the C ABI with HOST buffers: ``ba_set_problem`` (H2D + device-side sorts) + ``ba_solve`` +
``ba_get_params`` (D2H) inside the timed region.

N > 1 (torchrun, one process per GPU): STRONG scaling of the same 5 M problem (BASELINE config 4) - points sharded by rank,
cameras replicated, the per-iteration exchange inside the library (peer memory / NCCL).  Extra keys carry the weak-scaled
figure (N x 5 M), at N = 1 and N = 8 the 50 M-observation problem (config 5), the per-PCG-iteration kernel times and
``parity_vs_1gpu`` (the N-rank solve against a single-GPU solve of the same problem on rank 0).

Reference arm (``--impl reference``, rank 0 only, CPU): the oracle port of the reference's call (tier T1: the identical
``least_squares`` call with the vectorised residual).  The FULL workload is solved once (first warm-up step) and reported under
``full_config``; if W + K such solves fit the time budget every step is a full solve, otherwise the remaining steps run on the
largest sample of the same shape that fits (stated in ``cpu_baseline.sample``).  ``config`` is the workload's, identical in both
arms.
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
CONFIG_NO = {"bal50": 2, "ladybug": 3, "venice": 4, "shard50m": 5}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="venice", choices=["bal50", "ladybug", "venice", "shard50m"])
    ap.add_argument("--scaling", default="strong", choices=["weak", "strong"])
    ap.add_argument("--extras", default="auto", choices=["auto", "none"],
                    help="auto: also the weak-scaled figure (N > 1), the 50 M-observation problem (N = 1, 8), parity against 1 GPU")
    ap.add_argument("--cpu-sample-obs", type=int, default=5_000_000,
                    help="observations in the cpu_baseline sample of our arm (default: the whole 5 M venice workload, one SciPy solve of ~35 s)")
    ap.add_argument("--ref-budget-s", type=float, default=540.0, help="wall-clock budget of the whole reference arm")
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


# ------------------------------------------------------------------------------------------------
# workloads (identical description in both arms)
# ------------------------------------------------------------------------------------------------
def shape_of(name, mult=1):
    from bundle_adjustment_b200 import synthetic as S
    n_cams, n_pts, n_obs = S.SHAPES[name]
    return n_cams, n_pts * mult, n_obs * mult


def workload_label(name, mult=1):
    n_cams, n_pts, n_obs = shape_of(name, mult)
    return f"{name}-shape synthetic: {n_cams} cameras / {n_pts} points / {n_obs} observations"


def workload_config(name):
    """The ``config`` both arms print: what is solved, not how."""
    from bundle_adjustment_b200 import synthetic as S
    return {"workload": workload_label(name) + f" (BASELINE.json config {CONFIG_NO[name]})",
            "step": "one solve from the generator's initial guess with the reference's stopping rule: Huber 1 px per coordinate, "
                    "ftol = xtol = 1e-5, gtol = 1e-8, at most 50 iterations (src/bundle_adjuster.py:170-174)",
            "generator": f"bundle_adjustment_b200.synthetic.make_named_problem('{name}'), seed {S.SEEDS[name]}: 0.5 px noise, 5 % "
                         "outliers of +-50 px, initial guess = truth + noise (SURVEY.md 8d)",
            "l2": "inputs larger than L2 (observation streams >= 2 x 24 B x n_obs + 128 B point records), no flush needed"}


def make_problem(name, mult=1):
    from bundle_adjustment_b200 import synthetic as S
    n_cams, n_pts, n_obs = shape_of(name, mult)
    return S.make_flat_problem(n_cams, n_pts, n_obs, S.SEEDS[name], keep_truth=False)


# ------------------------------------------------------------------------------------------------
# CPU legs (the only places this file touches oracle/)
# ------------------------------------------------------------------------------------------------
def cpu_solve(fp):
    """One run of the reference's least_squares call (vectorised residual, tier T1) on problem `fp`."""
    from oracle import ba_oracle as O
    prob = O.Problem.from_flat(fp)
    x0 = fp.x0()
    c0 = O.huber_cost(prob.residuals(x0))
    t0 = time.perf_counter()
    res = O.solve_reference(prob, x0)
    dt = time.perf_counter() - t0
    return {"seconds": dt, "nfev": int(res.nfev), "njev": int(res.njev), "status": int(res.status), "initial_cost": c0,
            "final_cost": float(res.cost), "n_obs": fp.n_obs, "n_cams": fp.n_cams + 1, "n_pts": fp.n_pts,
            "value": fp.n_obs * int(res.nfev) / dt}


def sample_problem(name, n_obs_sample):
    from bundle_adjustment_b200 import synthetic as S
    n_obs = S.SHAPES[name][2]
    return S.make_named_problem(name, scale=min(1.0, n_obs_sample / n_obs), keep_truth=False)


def describe(r, name):
    return (f"scipy.optimize.least_squares(jac_sparsity, loss='huber', xtol=ftol=1e-5, max_nfev=50), vectorised residual (tier T1), "
            f"on {r['n_cams']} cameras / {r['n_pts']} points / {r['n_obs']} observations of the {name} shape: {r['nfev']} evaluations, "
            f"{r['njev']} Jacobians in {r['seconds']:.1f} s, status {r['status']}, Huber cost {r['initial_cost']:.6e} -> {r['final_cost']:.6e}")


def t0_record():
    """Tier T0 (verbatim reference _cost_function): committed figure with provenance - /root/reference does not travel."""
    path = os.path.join(ROOT, "tests", "golden", "t0_baseline.json")
    if not os.path.isfile(path):
        return None
    with open(path) as f:
        t0 = json.load(f)
    return {"us_per_observation_per_call": t0["us_per_observation_per_call"], "slice_observations": t0["slice_observations"],
            "calls_per_outer_iteration": t0["calls_per_outer_iteration"], "where": t0["where"], "threads": t0["threads"],
            "window_run_seconds": t0["window_run"]["seconds"], "source": "tests/golden/t0_baseline.json (oracle/gen_t0_baseline.py)"}


def run_reference(args, rank):
    """--impl reference: the reference's CPU path (oracle port, tier T1) on the host cores."""
    if rank != 0:
        return
    t_start = time.perf_counter()
    name = args.workload
    total = args.warmup + args.steps
    fp_full = make_problem(name)
    full = cpu_solve(fp_full)                               # the workload itself, once
    t_full = full["seconds"]
    left = args.ref_budget_s - (time.perf_counter() - t_start)
    steps_left = total - (1 if args.warmup >= 1 else 0)
    if steps_left * t_full <= left:
        fp, what = fp_full, "every step = the full workload"
    else:
        n_s = int(max(200_000, min(fp_full.n_obs, fp_full.n_obs * left / (steps_left * t_full * 1.1))))
        fp = sample_problem(name, n_s)
        what = (f"the full workload was solved once ({t_full:.1f} s, reported under full_config); {total} such solves do not fit "
                f"the {args.ref_budget_s:.0f} s budget, so the timed steps run on a {fp.n_obs}-observation sample of the same shape")
        del fp_full
    runs = []
    for i in range(total):
        r = full if (i == 0 and args.warmup >= 1 and fp.n_obs == full["n_obs"]) else cpu_solve(fp)
        if i >= args.warmup:
            runs.append(r)
    value = float(np.mean([r["value"] for r in runs]))
    t0 = t0_record()
    if t0:
        # one outer iteration of the verbatim reference = 10 _cost_function calls; extrapolated linearly in the observations
        t0["seconds_per_outer_iteration_full_workload"] = 1e-6 * t0["us_per_observation_per_call"] * full["n_obs"] * t0["calls_per_outer_iteration"]
        t0["value_extrapolated"] = full["n_obs"] / t0["seconds_per_outer_iteration_full_workload"]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean([r["seconds"] for r in runs])), "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(name),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port",
                         "sample": what + "; " + describe(runs[-1], name), "host_cpus": os.cpu_count(), "t0": t0},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "lm_iters_per_step": runs[-1]["nfev"], "final_cost": runs[-1]["final_cost"], "initial_cost": runs[-1]["initial_cost"],
        "status": {1: "gtol", 2: "ftol", 3: "xtol", 4: "ftol and xtol", 0: "iteration limit"}.get(runs[-1]["status"], "?"),
        "full_config": {**{k: full[k] for k in ("seconds", "nfev", "njev", "status", "initial_cost", "final_cost", "n_obs", "value")},
                        "cost_reduction": 1.0 - full["final_cost"] / full["initial_cost"]},
        "solver": "scipy least_squares trf / lsmr, 2-point finite differences over jac_sparsity, loss='huber' (single-threaded, as the reference)",
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# byte / flop accounting
# ------------------------------------------------------------------------------------------------
def algorithmic_bytes(kernel, n_obs, n_obs_adj, n_pts, n_cams, n_seg=0, n_rows=None, tile_lin=True):
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
    if kernel == "lin_points":
        if tile_lin:                         # k_tile_lin: uv 16 + record 8 read, cache 48 written; per point X 32 read, V 48 + g 32
            return 72 * n_obs + 148 * n_pts + 96 * (n_cams + 1)   # written, track tables ~36
        return (24 + 4 + 48) * n_obs + 112 * n_pts + 96 * (n_cams + 1)   # point-major k_lin_points
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


KEYS = ("status_text", "precond_text", "iterations", "accepted", "rejected", "pcg_iterations", "coarse_builds", "coarse_matvecs",
        "initial_cost", "final_cost", "gradient_max", "solve_ms", "launches")


def brief(st):
    return {k: st[k] for k in KEYS}


class Dist:
    def __init__(self, world, local_rank):
        import torch
        self.torch, self.world, self.dist = torch, world, None
        if world > 1:
            import torch.distributed as dist
            self.dist = dist
            torch.cuda.set_device(local_rank)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def new_nccl_id(self):
        """A fresh NCCL unique id (one per distributed solver handle), made on rank 0 and broadcast."""
        if self.dist is None:
            return None
        from bundle_adjustment_b200 import FlatSolver
        torch = self.torch
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if self.dist.get_rank() == 0:
            idt = torch.tensor(list(FlatSolver.nccl_unique_id()), dtype=torch.uint8, device="cuda")
        self.dist.broadcast(idt, 0)
        return bytes(idt.cpu().tolist())

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()

    def max(self, x):
        if self.dist is None:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def close(self):
        if self.dist is not None:
            self.dist.barrier()
            self.dist.destroy_process_group()


def timed_solves(D, solver, n_obs_total, warmup, steps, **opts):
    """Device-resident timed region: W untimed + K timed solves from the same initial guess.  -> (ms per solve, stats, launches)"""
    st = None
    for _ in range(warmup):
        solver.restore_initial()
        st = solver.solve(**opts)
    D.barrier()
    solver.timer_start()
    launches = 0
    for _ in range(steps):
        solver.restore_initial()
        st = solver.solve(**opts)
        launches += st["launches"] + 2
    ms = solver.timer_stop()
    D.barrier()
    ms = D.max(ms)
    return ms / steps, st, launches


def side_workload(D, rank, world, local_rank, name, mult, warmup=1, steps=2):
    """A second problem (weak-scaled workload, or config 5) solved on all ranks: -> dict for the JSON line."""
    from bundle_adjustment_b200 import FlatSolver, sharding
    fp_full = make_problem(name, mult)
    n_obs_total = fp_full.n_obs
    fp = fp_full if world == 1 else sharding.shard_problem(fp_full, rank, world)[0]
    del fp_full
    solver = FlatSolver(local_rank, rank=rank, nranks=world, nccl_id=D.new_nccl_id())
    solver.set_flat(fp)
    ms, st, _ = timed_solves(D, solver, n_obs_total, warmup, steps)
    out = {"workload": workload_label(name, mult), "ms_per_solve": ms, "value": n_obs_total * st["iterations"] / (ms * 1e-3),
           "unit": UNIT, **brief(st)}
    D.barrier()
    solver.close()
    return out


def run_ours(args, rank, world, local_rank):
    # libraries (NCCL) may write to fd 1; keep the JSON line the only thing on stdout
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    from bundle_adjustment_b200 import FlatSolver, sharding

    D = Dist(world, local_rank)
    name = args.workload
    mult = world if args.scaling == "weak" else 1
    fp_full = make_problem(name, mult)
    n_obs_total = fp_full.n_obs
    fp = fp_full if world == 1 else sharding.shard_problem(fp_full, rank, world)[0]

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

    solver = FlatSolver(local_rank, rank=rank, nranks=world, nccl_id=D.new_nccl_id())

    def upload():
        solver.set_problem(fp.fx, fp.fy, fp.cx, fp.cy, host["rvec"], host["tvec"], fp.fixed_rvec, fp.fixed_tvec,
                           host["pts"], host["obs_cam"], host["obs_pt"], host["obs_uv"])

    upload()
    warm = max(args.warmup, 3)

    # ---- timed region 1: device-resident ---------------------------------------------------------
    st = None
    for _ in range(warm):
        solver.restore_initial()
        st = solver.solve()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    ms_step, stats, launches = timed_solves(D, solver, n_obs_total, 0, args.steps)
    clocks = sampler.stop() if sampler else None
    iters = stats["iterations"]
    value = n_obs_total * iters / (ms_step * 1e-3)
    tr = solver.solve_trace()
    trajectory = [[round(float(m), 3), float(c)] for m, c in zip(tr["ms"], tr["cost"])]

    # ---- timed region 2: end to end through the C ABI with host buffers ---------------------------
    upload(); solver.solve(); solver.get_params()
    D.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        upload()
        st_e = solver.solve()
        solver.get_params()
    torch.cuda.synchronize()
    e2e_s = D.max(time.perf_counter() - t0)
    D.barrier()
    e2e_value = n_obs_total * st_e["iterations"] * args.steps / e2e_s

    # ---- per-kernel device times (pooled CUDA events on the solver stream, same solve) --------------
    solver.restore_initial()
    # pcg_check_every=1: the host looks at the convergence flag after every PCG iteration, so no launch of this
    # solve is an early exit and launches == iterations (same arithmetic, same iteration counts)
    st_p = solver.solve(verbose=2, pcg_check_every=1)
    ks = solver.kernel_stats()
    total_kernel_ms = sum(v[1] for v in ks.values())
    layout = solver.layout_info()
    n_obs_adj = int((fp.obs_cam >= 0).sum())
    n_seg, n_rows = layout["segments"], layout["matvec_rows"]
    tile_lin = layout["point_tiles"] > 0

    def ab(k):
        return algorithmic_bytes(k, fp.n_obs, n_obs_adj, fp.n_pts, fp.n_cams, n_seg, n_rows, tile_lin)
    top = max((k for k in ks if ab(k) is not None), key=lambda k: ks[k][1])
    peak, peak_src = peaks()
    avg_ms = ks[top][1] / max(ks[top][0], 1)
    achieved = ab(top) / (avg_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.isfile(tpath) and world == 1:
        with open(tpath) as f:
            traffic = json.load(f).get(name, {}).get(top)
    fp64_peak = solver.fp64_peak()
    fl = algorithmic_flops(top, fp.n_obs, fp.n_pts)
    fp64 = {"peak_tflops": fp64_peak, "kernel": top, "flops_per_launch": fl,
            "achieved_tflops": (fl / (avg_ms * 1e-3) / 1e12) if fl else None,
            "util": (fl / (avg_ms * 1e-3) / 1e12 / fp64_peak) if fl and fp64_peak > 0 else None,
            "how": "ba_fp64_peak: 16 independent DFMA chains per thread, 8 CTAs x 256 threads per SM, best of 3 (CUDA events)"}

    def us(k):
        return 1e3 * ks[k][1] / ks[k][0] if k in ks and ks[k][0] else None
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "frac_of_nominal_8000": achieved / 8000.0, "fp64": fp64,
                "traffic": traffic, "algorithmic_bytes_per_launch": ab(top), "avg_launch_us": 1e3 * avg_ms,
                "launches_per_solve": ks[top][0], "share_of_kernel_time": ks[top][1] / max(total_kernel_ms, 1e-9),
                "peak_source": peak_src, "layout": layout,
                "kernels": {k: {"launches": v[0], "ms": round(v[1], 3), "us_per_launch": round(1e3 * v[1] / v[0], 2),
                                "GBps": (ab(k) or 0) * v[0] / max(v[1], 1e-9) / 1e6,
                                "frac": (ab(k) or 0) * v[0] / max(v[1], 1e-9) / 1e6 / peak}
                            for k, v in ks.items() if v[0] and v[1] > 0.0}}
    pcg_iter = {"tile_pass_us": us("schur_matvec_tiles"), "tail_us": us("pcg_tail"),
                "us_per_pcg_iteration_event": (us("schur_matvec_tiles") or 0) + (us("pcg_tail") or 0),
                "kernels_per_lm_iteration_us": sum(v[1] for k, v in ks.items()
                                                   if k not in ("schur_matvec_tiles", "pcg_tail", "rows_finish", "setup")) * 1e3 / max(st_p["iterations"], 1),
                "how": "pooled CUDA events around every launch of a separate solve with pcg_check_every = 1"}

    # ---- N > 1: the same problem on ONE GPU (rank 0), for parity -------------------------------------
    parity = None
    if world > 1 and args.extras == "auto":
        D.barrier()
        if rank == 0:
            single = FlatSolver(local_rank)
            single.set_flat(fp_full)
            single.solve()
            single.restore_initial()
            s1 = single.solve()
            single.close()
            parity = {"one_gpu": brief(s1), "n_gpu": brief(stats),
                      "final_cost_rel_diff": abs(s1["final_cost"] - stats["final_cost"]) / s1["final_cost"],
                      "initial_cost_rel_diff": abs(s1["initial_cost"] - stats["initial_cost"]) / s1["initial_cost"],
                      "iterations_equal": s1["iterations"] == stats["iterations"],
                      "pcg_iterations": [s1["pcg_iterations"], stats["pcg_iterations"]],
                      "speedup_vs_1gpu_same_run": s1["solve_ms"] / ms_step}
        D.barrier()
    del fp_full
    solver.close()

    # ---- extras: other workloads of BASELINE.json ------------------------------------------------------
    extras = {}
    if args.extras == "auto" and name == "venice":
        if world > 1 and args.scaling == "strong":
            extras["weak"] = side_workload(D, rank, world, local_rank, "venice", world)
        if world in (1, 8):
            extras["config5"] = side_workload(D, rank, world, local_rank, "shard50m", 1)

    if rank != 0:
        D.close()
        return

    # ---- CPU baseline (N = 1): ONE SciPy solve of the workload itself (venice: ~35 s; --cpu-sample-obs bounds it to a sample of the
    #      same shape); the GPU solves the SAME arrays from the same initial guess for a same-problem comparison ----
    cpu, ms_to_ref = None, None
    if world == 1 and not args.no_cpu_baseline:
        fps = sample_problem(name, args.cpu_sample_obs)
        r = cpu_solve(fps)
        with FlatSolver(local_rank) as s2:
            s2.set_flat(fps)
            s2.solve()
            s2.restore_initial()
            g = s2.solve()
            tr2 = s2.solve_trace()
        below = np.flatnonzero(tr2["cost"] <= r["final_cost"])
        whole = fps.n_obs == n_obs_total and mult == 1
        # device time at which the Huber cost is first at or below what the reference's call reached from the same x0 (entry 0 of
        # the trace = the evaluation of the initial point: the reference's xtol rule returns x0 itself on the large shapes)
        ms_to_ref = {"ms": float(tr2["ms"][below[0]]) if below.size else None, "reference_final_cost": r["final_cost"],
                     "reference_initial_cost": r["initial_cost"], "reference_seconds": r["seconds"], "reference_nfev": r["nfev"],
                     "same_problem": "the workload itself" if whole else f"a {fps.n_obs}-observation sample of the workload's shape"}
        cpu = {"value": r["value"], "unit": UNIT, "cores": 1, "kind": "port", "sample": describe(r, name), "host_cpus": os.cpu_count(),
               "t0": t0_record(),
               "same_config": bool(whole),
               "same_problem": {"what": "the GPU arm on the cpu_baseline problem itself (same arrays, same initial guess)",
                                "cpu_seconds": r["seconds"], "cpu_final_cost": r["final_cost"], "cpu_cost_reduction": 1.0 - r["final_cost"] / r["initial_cost"],
                                "gpu_ms": g["solve_ms"], "gpu_final_cost": g["final_cost"], "gpu_status": g["status_text"],
                                "gpu_iterations": g["iterations"],
                                "gpu_ms_to_cpu_final_cost": float(tr2["ms"][below[0]]) if below.size else None,
                                "time_per_solve_ratio": 1e3 * r["seconds"] / g["solve_ms"]}}

    cfg = workload_config(name)
    if mult != 1:
        cfg["workload"] = workload_label(name, mult) + " (weak scaling of BASELINE.json config 4)"
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": cfg,
        "solver": {"parallelism": f"points sharded over {world} GPU(s), cameras replicated", "robust": "huber 1 px per coordinate (IRLS)",
                   "pcg": f"{stats['precond_text']}, forcing term 0.1", "status": stats["status_text"]},
        "lm_iters_per_step": iters, "lm_iters_per_s": iters / (ms_step * 1e-3), "solves_per_s": 1e3 / ms_step,
        "pcg_iters_per_step": stats["pcg_iterations"], "coarse_matvecs_per_step": stats["coarse_matvecs"],
        "final_cost": stats["final_cost"], "initial_cost": stats["initial_cost"], "status": stats["status_text"],
        "ms_to_converged": ms_step if stats["status"] in (1, 2, 3, 4) else None,
        "ms_to_reference_cost": ms_to_ref,
        "trajectory_ms_cost": trajectory,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": 1e3 * e2e_s / args.steps},
        "gpu_launches": int(launches),
        "clocks": clocks, "roofline": roofline, "pcg_iteration": pcg_iter, "cpu_baseline": cpu,
        "parity_vs_1gpu": parity, **extras,
    }
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    print(json.dumps(line), flush=True)
    os.dup2(2, 1)
    D.close()


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
