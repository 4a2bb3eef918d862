/*
 * ba_b200.h - C ABI of libba_b200.so, the B200 (sm_100a) bundle-adjustment back end.
 *
 * Drop-in boundary for the local/global BA step of egirgin/bundle_adjustment
 * (reference: src/bundle_adjuster.py).  The reference is pure Python and has no FFI; a maintainer
 * binds this library with ctypes.  Each entry point below names the reference
 * code it replaces.  Plain pointers and sizes only; all arrays are caller-owned HOST memory, borrowed
 * for the duration of the call; all floating point is IEEE binary64; all indices are int32.
 *
 * Conventions shared with the reference:
 *   pose      x_cam = R(rvec) * X + t          (cv2.projectPoints, src/bundle_adjuster.py:67)
 *   pinhole   u = fx x/z + cx, v = fy y/z + cy (no distortion, no skew)
 *   residual  r[2i], r[2i+1] = observed(u,v)_i - projected(u,v)_i      (src/bundle_adjuster.py:68-70)
 *   params    [rvec_0..rvec_{Nc-1} | t_0..t_{Nc-1} | X_0..X_{Np-1}]      (src/bundle_adjuster.py:157-162)
 *   loss      Huber per scalar residual, threshold huber_delta (=1 px), cost = 0.5 * sum rho(r^2)
 *             (scipy.optimize.least_squares(loss='huber', f_scale=1), src/bundle_adjuster.py:170-174)
 *
 * Return value of every int function: 0 = ok, < 0 = error (text via ba_last_error()).
 * ba_solve additionally reports the termination reason in ba_stats.status (> 0, mirrors SciPy's
 * OptimizeResult.status: 1 gtol, 2 ftol, 3 xtol, 4 ftol and xtol; 0 = iteration limit; 5 = damping
 * limit reached).
 */
#ifndef BA_B200_H
#define BA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ba_handle ba_handle;

#define BA_OK 0
#define BA_ERR_ARG (-1)
#define BA_ERR_CUDA (-2)
#define BA_ERR_NCCL (-3)
#define BA_ERR_STATE (-4)
#define BA_ERR_NONFINITE (-5) /* residuals not finite at the initial point (SciPy raises ValueError there) */

#define BA_ROBUST_NONE 0   /* plain least squares */
#define BA_ROBUST_IRLS 1   /* curvature weight = gradient weight = rho'(r^2)            (default) */
#define BA_ROBUST_TRIGGS 2 /* SciPy's scale_for_robust_loss_function: max(rho'+2 rho'' r^2, eps) */

#define BA_PRECOND_JACOBI 0       /* camera blocks of J^T J                                    */
#define BA_PRECOND_SCHUR_JACOBI 1 /* diagonal blocks of the Schur complement                    */
#define BA_PRECOND_TWO_LEVEL 2    /* Schur-Jacobi plus an additive coarse correction  Z (Z^T S Z)^-1 Z^T : one rigid 6-vector per
                                     aggregate of consecutive cameras; the coarse matrix is built from the Jacobian cache in one
                                     pass over the points (fallback: probed with the tile pass, one mat-vec per colour of the
                                     aggregate graph and component) once the PCG of an LM iteration
                                     has needed more than 24 iterations, inverted on the device and kept (rebuilt when the
                                     iteration count climbs again; skipped for an LM iteration whose damping is more than 4x away
                                     from the one it was built with).  Needs the default kernel path (tiles + cluster tail; on
                                     N > 1 GPUs the peer-memory exchange); where that is unavailable the solve runs with
                                     Schur-Jacobi and says so in ba_stats.precond_used.  Host-driven loops.              */
#define BA_PRECOND_AUTO 3         /* two-level for >= 128 adjustable cameras, else Schur-Jacobi               (default) */

#define BA_PATH_TILES 0    /* fused point-tile passes: one kernel per Schur mat-vec          (default) */
#define BA_PATH_TWO_PASS 1 /* point-major pass + camera-major pass (also used when a track is longer
                              than one tile, 256 observations)                                       */
#define BA_PATH_TILES_ROWS 2 /* fused tiles writing one row per segment (no shared-memory epoch accumulators); also the
                                automatic choice when an epoch of a problem needs more accumulator slots than fit */

typedef struct ba_options {
    int32_t max_iters;        /* LM iterations, accepted + rejected (reference: max_nfev = 50)      */
    double gtol;              /* stop when max |gradient| < gtol               (SciPy default 1e-8) */
    double ftol;              /* stop when cost decrease < ftol * cost          (reference 1e-5)    */
    double xtol;              /* stop when |step| < xtol * (xtol + |x|)         (reference 1e-5)    */
    double huber_delta;       /* Huber threshold in pixels (reference: f_scale = 1.0)               */
    int32_t robust_mode;      /* BA_ROBUST_*                                                        */
    double initial_lambda;    /* LM damping at the first iteration                                  */
    double lambda_min, lambda_max;
    int32_t precond;          /* BA_PRECOND_*                                                       */
    int32_t pcg_max_iters;    /* per LM iteration                                                   */
    double pcg_tol;           /* forcing term: stop PCG when |r|_M^-1 <= pcg_tol * |b|_M^-1          */
    int32_t pcg_check_every;  /* host polls the device "PCG converged" flag every this many iterations */
    int32_t verbose;          /* bit 0: one stderr line per LM iteration; >= 2: per-kernel event timing */
    int32_t schur_path;       /* BA_PATH_*                                                          */
    int32_t pcg_tail;         /* 1 (default): everything between two tile passes of a PCG iteration runs in ONE
                                 thread-block-cluster kernel; 0: row finish / inner products / update kernels */
    int32_t device_loop;      /* single-GPU solves on the default kernel path can run as ONE CUDA graph whose conditional WHILE
                                 nodes loop over the LM and the PCG iterations on the device (one launch and one read-back per
                                 solve).  1 (default): for problems of up to 262144 observations (where it is faster); 2: always;
                                 0: the host drives the loops (always the case for verbose / timed / multi-GPU solves) */
    int32_t coarse_aggregate; /* cameras per aggregate of the two-level preconditioner; 0 (default) = max(32, n_cams / 64)  */
    int32_t coarse_after;     /* build the coarse matrix once an LM iteration's PCG needed more than this many iterations;
                                 0 (default) = 24, negative = from the first LM iteration on                              */
    int32_t point_linesearch; /* 1 (default): after every LM step each point scales its own update by the factor out of
                                 {1, 1/2, 2, 4, 8, 16} with the smallest Huber cost over its observations (cameras at the trial
                                 poses; the total cost can only go down against the plain step).  0: plain LM steps           */
    int32_t dense_schur;      /* 1 (default): single-GPU problems with at most 16 adjustable cameras (the pipeline's windows) assemble the
                                 reduced camera system explicitly and solve it by a dense factorisation (L D L^T by one warp up to 32 unknowns, else Cholesky) - the exact LM step in
                                 two kernels - instead of PCG (csrc/ba_dense.cuh).  0: PCG always                              */
} ba_options;

typedef struct ba_stats {
    int32_t status;           /* termination reason, see top of file                                */
    int32_t iterations;       /* LM iterations run (accepted + rejected)                            */
    int32_t accepted, rejected;
    int32_t pcg_iterations;   /* total over the solve                                               */
    int32_t launches;         /* CUDA kernels launched by the solve                                 */
    double initial_raw_cost;  /* sum r^2 at the initial point  (src/bundle_adjuster.py:165)         */
    double final_raw_cost;    /* sum r^2 at the returned point (src/bundle_adjuster.py:176)         */
    double initial_cost;      /* 0.5 * sum rho at the initial point (SciPy's cost)                  */
    double final_cost;        /* 0.5 * sum rho at the returned point                                */
    double gradient_max;      /* max |g| at the returned point (local camera coordinates)           */
    double final_lambda;
    double solve_ms;          /* CUDA-event time of the device LM loop                              */
    int32_t precond_used;     /* BA_PRECOND_* the PCG actually ran with (TWO_LEVEL only if a coarse matrix was built)   */
    int32_t coarse_builds;    /* coarse matrices built during the solve                                                 */
    int32_t coarse_matvecs;   /* tile-pass mat-vecs spent on probing them (0 with the direct build; not in pcg_iterations) */
    int32_t coarse_unknowns;  /* size of the coarse system (6 x aggregates), 0 if none                                  */
    int32_t dense_solves;     /* LM iterations whose step came from the dense factorisation (0: PCG was used)         */
    int32_t reserved_i;
    double reserved[5];
} ba_stats;

/* Defaults: max_iters 50, gtol 1e-8, ftol = xtol = 1e-5 (the reference's call, src/bundle_adjuster.py:172), huber 1.0, IRLS,
 * lambda 1e-4 (floor 1e-6: keeps the step bounded along the free scale gauge and along low-parallax depth directions),
 * PCG with BA_PRECOND_AUTO, forcing term 0.1. */
void ba_default_options(ba_options* opt);

/* One handle = one GPU + one CUDA stream.  `device` is a CUDA ordinal.  Replaces nothing in the
 * reference (which keeps no state between run() calls); device buffers live inside the handle. */
int ba_create(int device, ba_handle** out);

/* Multi-GPU handle: rank `rank` of `nranks` one-process-per-GPU solvers that each own a contiguous
 * range of the points (and those points' observations) and replicate all cameras.  `nccl_id` is the
 * 128-byte id from ba_nccl_unique_id() on rank 0, distributed by the host (torch.distributed in
 * bench.py).  Camera-side sums and scalar costs are all-reduced over NCCL; nothing else moves. */
int ba_nccl_unique_id(char id_out[128]);
int ba_create_distributed(int device, int rank, int nranks, const char nccl_id[128], ba_handle** out);

/* Distributed handles: COLLECTIVE - every rank unmaps its peers' exchange buffers, the ranks meet in one NCCL all-reduce, and
 * only then is the exported buffer freed (the same order is kept when a later ba_set_problem has to grow it). */
int ba_destroy(ba_handle* h);

/* Upload one problem.  Replaces the flattening at src/bundle_adjuster.py:150-162 on the device side:
 * the host passes the arrays in the reference's own order (observations keyframe-major as
 * _gather_local_data :195-218 emits them; obs_cam = index into the adjustable keyframes or -1 for the
 * fixed keyframe; obs_pt = index into the sorted map-point ids).  The point-major and camera-major
 * permutations the kernels need are built on the device.  n_adj_cams may be 0. */
int ba_set_problem(ba_handle* h, double fx, double fy, double cx, double cy,
                   int32_t n_adj_cams, int32_t n_pts, int32_t n_obs,
                   const double* rvec /* [n_adj_cams*3] */, const double* tvec /* [n_adj_cams*3] */,
                   const double fixed_rvec[3], const double fixed_tvec[3],
                   const double* pts /* [n_pts*3] */,
                   const int32_t* obs_cam /* [n_obs] */, const int32_t* obs_pt /* [n_obs] */,
                   const double* obs_uv /* [n_obs*2] */);

/* Overwrite / read back the parameters of the uploaded problem (same layouts as ba_set_problem).
 * ba_get_params replaces the unpacking half of _update_map (src/bundle_adjuster.py:220-233). */
int ba_set_params(ba_handle* h, const double* rvec, const double* tvec, const double* pts);
int ba_get_params(ba_handle* h, double* rvec, double* tvec, double* pts);

/* Device-side reset: make the parameters uploaded by the last ba_set_problem current again (a
 * device-to-device copy; used to repeat a solve without re-uploading). */
int ba_restore_initial(ba_handle* h);

/* CUDA-event stopwatch on the handle's stream: everything enqueued between start and stop is timed on
 * the device.  ba_timer_stop synchronises and returns milliseconds. */
int ba_timer_start(ba_handle* h);
int ba_timer_stop(ba_handle* h, double* ms_out);

/* Measured FP64 FMA rate of the device (dependent-free DFMA chains on every SM, CUDA events): the denominator of
 * the "FP64 pipe utilisation" figure bench.py reports (not in MEASURED_PEAKS.json).  Needs no problem. */
int ba_fp64_peak(ba_handle* h, double* tflops_out);

/* r_out[2*n_obs] in the reference's observation order.  Replaces _cost_function
 * (src/bundle_adjuster.py:24-72). */
int ba_residuals(ba_handle* h, double* r_out);

/* sum r^2 (what the reference prints and guards on, :165/:176) and 0.5*sum rho(r^2) (SciPy's cost)
 * at the current parameters, with the Huber threshold `huber_delta`. */
int ba_cost(ba_handle* h, double huber_delta, double* raw_out, double* robust_out);

/* Analytic Jacobian blocks of the residuals in the reference's parameterisation:
 * jc_out[n_obs*12]: d r_i / d(rvec_c, t_c) row-major 2x6 (all zero for fixed-keyframe rows),
 * jp_out[n_obs*6] : d r_i / d X_p       row-major 2x3.
 * Replaces the 2-point finite differences SciPy takes over jac_sparsity
 * (src/bundle_adjuster.py:167-171; scipy/optimize/_numdiff.py approx_derivative). */
int ba_jacobian_blocks(ba_handle* h, double* jc_out, double* jp_out);

/* CSR sparsity pattern of the Jacobian, bit-identical to
 * _prepare_sparsity_matrix(...).tocsr() (src/bundle_adjuster.py:74-120): indptr[2*n_obs+1],
 * indices[nnz] with nnz = 18 per adjustable-keyframe observation and 6 per fixed-keyframe one.
 * Pass indices = NULL to query only nnz (returned through nnz_out). */
int ba_sparsity_csr(ba_handle* h, int64_t* indptr, int64_t* indices, int64_t* nnz_out);

/* Gradient and Gauss-Newton Hessian blocks at the current parameters in the solver's local camera
 * coordinates (world-frame twist), for tests: gc[n_adj*6], gp[n_pts*3], U[n_adj*36], V[n_pts*9].
 * Any pointer may be NULL. */
int ba_normal_blocks(ba_handle* h, const ba_options* opt, double* gc, double* gp, double* U, double* V);

/* Test hook: w = (U + lambda D - E (V + lambda D_p)^-1 E^T) u, the damped Schur complement of the
 * Gauss-Newton system at the current parameters applied to u[n_adj*6], through the kernels of
 * opt->schur_path (exactly what one PCG iteration runs).  Single-GPU handles only. */
int ba_schur_matvec(ba_handle* h, const ba_options* opt, double lambda, const double* u, double* w_out);

/* Run the Levenberg-Marquardt loop on the device from the current parameters; the parameters are
 * replaced by the result.  Replaces the least_squares call at src/bundle_adjuster.py:170-174. */
int ba_solve(ba_handle* h, const ba_options* opt, ba_stats* stats_out);

/* Trajectory of the last ba_solve, recorded on the device by the 1-thread decision kernel (time stamps from %globaltimer, so
 * they include everything the solve enqueued and no host time): entry 0 = the initial point, entry i = the state after LM
 * iteration i.  cost = 0.5 * sum rho and raw_cost = sum r^2 of the CURRENT point (unchanged by a rejected step), ms = device time
 * since the start of the solve, pcg_total = PCG iterations so far, accepted = 1 if iteration i was accepted, lambda = the damping
 * iteration i's step was computed with (entry 0: the initial damping).  Up to `cap` entries
 * are written (any array may be NULL); *n_out = entries available (iterations + 1, at most 1024).  What the reference offers
 * instead: verbose=2 prints of least_squares (src/bundle_adjuster.py:173 passes verbose=0). */
int ba_solve_trace(ba_handle* h, int32_t cap, double* cost, double* raw_cost, double* ms, int32_t* pcg_total, int32_t* accepted,
                   double* lambda, int32_t* n_out);

/* Per-kernel accounting of the last ba_solve: for each of the `n` kernel kinds (ba_kernel_name(k)),
 * number of launches and summed CUDA-event milliseconds.  Times are filled when opt->verbose >= 2:
 * every launch is bracketed by a pair of pooled CUDA events on the solver stream (no synchronisation
 * between launches; resolved at the end of the solve).  Otherwise launch counts only. */
int ba_kernel_count(void);
const char* ba_kernel_name(int k);
int ba_kernel_stats(ba_handle* h, int64_t* launches /* [n] */, double* ms /* [n] */);

/* Layout of the uploaded problem, for byte accounting: out[0..11] = n_obs, observations of adjustable
 * cameras, n_pts, n_adj_cams, point tiles, tile segments, point-major windows, camera-major tiles,
 * mat-vec rows (= epoch accumulator slots, or = segments when the accumulators are not in use), epochs,
 * largest number of slots of an epoch, tiles per epoch aimed at (0 = no accumulators).
 * Point tiles = 0 when the fused path is unavailable (a track longer than a tile). */
int ba_layout_info(ba_handle* h, int64_t out[12]);

/* Test hook: copy out the point-tile structure built by ba_set_problem (sizes from ba_layout_info:
 * T = point tiles, S = segments).  ft_pt/ft_obs/ft_seg[T+1], tc_meta[n_obs*2] (camera slot, local point |
 * point-major position << 10), tc_uv[n_obs*2],
 * seg_cam[S], seg_info[S] (first position inside the tile | length << 16), seg_row[S], cam_row_off[n_adj+2], pt_perm[n_pts] (internal point ->
 * caller's index).  Any pointer may be NULL. */
int ba_tile_layout(ba_handle* h, int32_t* ft_pt, int32_t* ft_obs, int32_t* ft_seg, int32_t* tc_meta, double* tc_uv,
                   int32_t* seg_cam, int32_t* seg_info, int32_t* seg_row, int32_t* cam_row_off, int32_t* pt_perm);

/* Test hook: the packed work items of the mat-vec pass with epoch accumulators (csrc/ba_tiles.cuh, k_ft_item_fill; no counterpart in
 * the reference, whose least_squares call has no such structure).  *n_slots = item slots of all tiles (0 if the accumulators are
 * not in use); items[*n_slots] (item = first position | length << 8 | pair << 13 | accumulator slot << 15, 0 = empty lane, eight
 * per quarter-warp); hdr_acc[T * 8] = the pass's tile headers ((p0, points, o0, observations), (first item, item slots, first
 * track row, longest track)).  items / hdr_acc may be NULL (size query). */
int ba_tile_items(ba_handle* h, int64_t* n_slots, int32_t* items, int32_t* hdr_acc);

/* Two-view linear triangulation of n correspondences and the cheirality test, on the device: replaces the
 * reference's  cv2.triangulatePoints(P1, P2, pts1.T, pts2.T)  followed by  X = h[:3] / (h[3] + 1e-6), z > 0 in
 * both views  (src/pipeline.py:315-336).  P1, P2: row-major 3x4 projection matrices; R_rel[9] (row-major), t_rel[3]:
 * pose of the second view (only used for `valid`, may be NULL when valid is NULL); pts1, pts2 [n*2] pixels.
 * Outputs, each may be NULL: hom[n*4] unit homogeneous solutions (sign arbitrary, as OpenCV's), xyz[n*3],
 * valid[n] (1 = in front of both cameras).  Needs no uploaded problem. */
int ba_triangulate(ba_handle* h, const double P1[12], const double P2[12], const double* R_rel, const double* t_rel,
                   int32_t n, const double* pts1, const double* pts2, double* hom, double* xyz, uint8_t* valid);

/* ba_triangulate without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments, outputs and arithmetic as ba_triangulate (the same kernel), but the device staging of pts1 / pts2 / hom / xyz /
 * valid - which ba_triangulate takes from the stream-ordered memory pool - lives in ONE file-scope `static __device__` array of
 * fixed size inside the library: 11.125 doubles per correspondence x BA_TRIANGULATE_STATIC_MAX = 1.46 MB, held by every GPU
 * that loads the library until the process exits.  n > BA_TRIANGULATE_STATIC_MAX is rejected with BA_ERR_ARG.  The call issues
 * no cudaMalloc / cudaFree / cudaMallocAsync and no allocating library routine; the first call on a device resolves the array's
 * device address (cudaGetSymbolAddress) - make one warm-up call before capturing.  All calls on a device share the array: it
 * keeps the previous call's contents (the part a call reads back is cleared first), and calls that OVERLAP on the GPU (two
 * handles / streams of one device) race - serialise them. */
#define BA_TRIANGULATE_STATIC_MAX 16384
int ba_triangulate_static(ba_handle* h, const double P1[12], const double P2[12], const double* R_rel, const double* t_rel,
                          int32_t n, const double* pts1, const double* pts2, double* hom, double* xyz, uint8_t* valid);

/* ---- device-resident map of successive windows ----------------------------------------------------
 * Replaces, for a caller that runs local BA window after window on a growing map, the per-call flattening of the
 * reference: _gather_local_data (src/bundle_adjuster.py:195-218: keyframe-major (kf, mp) rows of the window, sorted unique
 * map point ids, the (kf, mp) -> keypoint dict) and the packing of x0 (:157-162).  The observation rows of every keyframe
 * and the positions of the map points stay in HBM between calls; a call uploads only what is new - the reference pipeline
 * only ever appends (src/pipeline.py:172-173, 215-217, 277-278, 306-308) - and ba_map_window builds on the device the
 * same flat problem ba_set_problem would have received: same row order, same point numbering (rank of the id among the
 * window's ids), the last keypoint of a (kf, mp) pair that is listed twice (:216), rows of unknown map points dropped
 * (:213).  Ids are dense table indices: 0 <= id < 2^24.  Single-GPU handles only.
 *
 *   ba_map_clear                 forget every resident row and point (the counters of ba_map_stats keep counting)
 *   ba_map_append_observations   rows (mp_id[i], uv[2i..2i+1]) of ONE keyframe, appended in the order of its observation list
 *   ba_map_set_points            insert or overwrite positions ([n][3]); the ids become "in the map"
 *   ba_map_forget_points         the ids are no longer in the map (`mp_id in gmap.map_points` false)
 *   ba_map_window                kf_id[0] is the fixed keyframe (:141), kf_id[1..] the adjustable ones in camera order;
 *                                rvec/tvec [n_kfs][3] are their CURRENT poses (240 bytes: uploaded every call).  Uploads the
 *                                problem as ba_set_problem does (so every other entry point works on it) and returns its
 *                                sizes; both 0 (and no problem) when the window holds no observation of a known point
 *   ba_map_window_points         ids of the window's points in problem order (= sorted), [n_pts]
 *   ba_map_commit                after an accepted solve: the solver's current points -> the resident positions
 *   ba_map_stats                 [0] resident rows, [1] rows appended so far, [2] positions set so far, [3] windows built,
 *                                [4] bytes uploaded by the ba_map_* calls so far, [5..7] 0 */
int ba_map_clear(ba_handle* h);
int ba_map_append_observations(ba_handle* h, int32_t kf_id, int32_t n, const int32_t* mp_id, const double* uv);
int ba_map_set_points(ba_handle* h, int32_t n, const int32_t* mp_id, const double* xyz);
int ba_map_forget_points(ba_handle* h, int32_t n, const int32_t* mp_id);
int ba_map_window(ba_handle* h, double fx, double fy, double cx, double cy, int32_t n_kfs, const int32_t* kf_id,
                  const double* rvec, const double* tvec, int32_t* n_pts_out, int32_t* n_obs_out);
int ba_map_window_points(ba_handle* h, int32_t* mp_id_out);
int ba_map_commit(ba_handle* h);
int ba_map_stats(ba_handle* h, int64_t out[8]);

/* Thread-local text of the last error returned on this thread. */
const char* ba_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* BA_B200_H */
