// ba_kernels.cuh - hand-written sm_100a kernels of the Levenberg-Marquardt / Schur / PCG loop.
//
// Dataflow ("recompute"): no per-observation Jacobian or E block is ever stored.  Every pass streams
// the observation record (u, v, camera, point) and re-derives the 2x3 point Jacobian from the camera
// pose and the point; the camera Jacobian follows from it algebraically:
//
//      x_cam = R (X + w x X + v) + t        (camera increment = world-frame twist (w, v))
//      Jp    = d r / d X      = -dpi(x_cam) R                       (2x3)
//      Jc    = d r / d (w, v) = Jp [ -[X]x | I ]                    (2x6)
//
// Two observation orders live in HBM:
//   * point-major  (pm_*): all observations of a point are adjacent.  Per-point sums (V, g_p, the
//     Schur mat-vec's E^T x) are formed inside one CTA: phase A = one thread per observation writes
//     its contribution to shared memory, phase B = one thread per point adds them in a fixed order.
//   * camera-major (cm_*): only observations of adjustable cameras, cut into tiles of <= CM_TILE
//     observations of ONE camera.  Per-camera sums (U, g_c, Schur right-hand side, preconditioner
//     blocks, mat-vec result) are block-reduced per tile (warp shuffles + shared memory, fixed tree)
//     and the few tile partials of a camera are added in tile order by a finishing kernel.
// No floating-point atomics anywhere: results are bit-reproducible run to run.
//
// Gathers (camera records in point-major passes, point records in camera-major passes) are the
// expensive part: a naive per-lane gather costs one L1 wavefront per lane per 16-byte load.  Instead a
// warp gathers COOPERATIVELY: the 32 records it needs are cut into 16-byte chunks, consecutive lanes
// copy consecutive chunks of the same record straight into shared memory with cp.async (LDGSTS), and
// every lane then reads its own record back from shared memory.
//
// All loop-carried scalars (damping, costs, PCG scalars, buffer selector, flags) live in the device
// struct LMState so that the host only enqueues launches and polls a status word.
#pragma once
#include "ba_math.cuh"

#include <cooperative_groups.h>

namespace ba {

constexpr int PM_THREADS = 256;  // point-major CTA; nominal tile = PM_THREADS observations
constexpr int CM_THREADS = 256;  // camera-major CTA
constexpr int CM_ROUNDS = 4;     // observations per thread in the smallest camera-major tile
constexpr int CM_TILE = CM_THREADS * CM_ROUNDS;  // tiles grow up to 8 x this on large problems (fewer 27-value CTA reductions)
constexpr int NSCAL = 16;        // per-iteration scalar slots
constexpr int VEC_THREADS = 128; // camera-vector kernels: one camera per thread
constexpr int MAX_VEC_CTAS = 8192;

// scalar slots (partials are reduced into LMState::scal by k_reduce_scalars)
enum {
    // point-side sums (per-rank partial in multi-GPU mode: slots [0, N_PT_SCAL) are all-reduced)
    S_COST = 0,    // 0.5 * sum rho
    S_RAW = 1,     // sum r^2
    S_GDP = 2,     // g_p . dp
    S_DDP = 3,     // dp^T D_p dp
    S_DP2 = 4,     // |dp|^2
    S_XP2 = 5,     // |X|^2
    S_BAD = 6,     // count of non-finite residuals
    // camera-side sums (replicated on every rank, never all-reduced)
    S_GDC = 7,     // g_c . dc
    S_DDC = 8,     // dc^T D_c dc
    S_RDC = 9,     // r_pcg . dc
    S_DC2 = 10,    // |dc|^2
    S_XC2 = 11,    // |(rvec,t)|^2
    S_GPMAX = 12,  // max |g_p|   (max-reduced, not summed)
};
constexpr int N_PT_SCAL = 7;

struct LMState {
    double lambda, nu;
    double cost[2], raw[2];  // per parameter buffer
    double gmax;             // max |g| at the current point
    double gamma, gamma0, gamma_prev, alpha, beta;  // PCG scalars: gamma = r.M^-1 r
    double model, rho, dcost, step_norm, x_norm;
    double scal[NSCAL];
    int cur;                 // which parameter buffer is current (0/1); trial = cur ^ 1
    int status;              // 0 running, > 0 terminated (ba_b200.h)
    int iter, accepted, rejected;
    int last_accepted;       // 1 if the last decide() accepted
    int pcg_done, pcg_iter, pcg_total;
    int bad;                 // non-finite seen
    unsigned long long t0;   // %globaltimer (ns) when the solve started: origin of the trace's time stamps
};
// One record per LM iteration (entry 0 = the initial point), written by the 1-thread decision kernels: Huber cost and raw sum of
// squares of the CURRENT point after the decision, device time since the start of the solve, PCG iterations so far.
constexpr int TRACE_CAP = 1024;
struct TraceRec { double cost, raw, ms, lambda; int pcg_total, accepted; };
BA_D unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// the last step was rejected: the camera-side linearisation of the current point is still valid
BA_D bool lin_gated(const LMState* st) { return st->iter > 0 && st->last_accepted == 0; }
// Device-side loop control (conditional WHILE nodes of the solve graph, see solve()): 0 = no graph, nothing to set.
BA_D void set_loop(unsigned long long handle, bool go_on) {
    if (handle) cudaGraphSetConditional((cudaGraphConditionalHandle)handle, go_on ? 1u : 0u);
}

struct Params {
    Intr K;
    double delta;  // Huber threshold
    int mode;      // robust mode
    int n_cams;    // adjustable cameras; slot n_cams holds the fixed camera
    int n_pts, n_obs, n_obs_adj;
    int pm_tiles, cm_tiles, pm_cap;  // pm_cap = shared-memory slots per contribution row
    int vec_ctas;                    // grid of the camera-vector kernels
    // parameters: [2][(n_cams+1)*6] (rvec,t), [2][(n_cams+1)*12] (R,t), [2][n_pts*4] points (32-byte slots)
    double* cam_rt;
    double* cam_R;
    double* pt_X;     // [2][pt_stride] point positions per parameter buffer: plane (X0, X1) [n_pad] double2, then
                      //                plane X2 [n_pad] (two planes: conflict-free 16-byte / 8-byte gathers once a
                      //                tile's points sit in shared memory)
    int n_pad;        // n_pts rounded up to even
    long long pt_stride;  // doubles per parameter buffer = 3 * n_pad
    double* pt_z;     // [n_pts*4]     V^-1 E^T u of the two-pass Schur mat-vec
    double* pt_h;     // [n_pts*4]     V^-1 g_p
    // per-point normal-equation data
    double* pt_V;     // [2][n_pts*6]
    double* pt_g;     // [2][n_pts*4]
    double* pt_Vinv;  // [n_pts*6]
    double* pt_D;     // [n_pts*4]  damping diagonal
    // per-camera data (local coordinates)
    double* cam_U;     // [n_cams*21]
    double* cam_g;     // [n_cams*6]
    double* cam_D;     // [n_cams*6]
    double* cam_Minv;  // [n_cams*36]
    double* cam_acc;   // [n_cams*27] tile-summed accumulators (all-reduced in multi-GPU mode)
    // PCG vectors, [(n_cams+1)*6]; the +1 slot (fixed camera) stays zero
    double* v_b;
    double* v_x;
    double* v_r;
    double* v_u;   // M^-1 r  (the vector the Schur mat-vec is applied to)
    double* v_p;
    double* v_s;   // A p
    double* v_w;   // A u
    double* dot_partial;  // [MAX_VEC_CTAS*2]
    // observations
    const double2* o_uv;   // reference order
    const int2* o_idx;     // (camera slot, point)
    const double2* pm_uv;  // point-major
    const int2* pm_idx;
    const int* pt_off;     // [n_pts+1]
    const int* pm_tile_pt; // [pm_tiles+1]
    const double2* cm_uv;  // camera-major (adjustable cameras only)
    const int* cm_pt;
    const int4* cm_tile;      // (camera, first obs, count, 0)
    const int* cam_tile_off;  // [n_cams+1]
    // fused point tiles (ba_tiles.cuh)
    int ft_tiles, n_seg;
    int n_partial_rows;       // rows of pm_partial = max(pm_tiles, ft_tiles)
    const int* ft_pt;         // [ft_tiles+1] first point of a tile
    const int* ft_obs;        // [ft_tiles+1] first observation of a tile (same range in both tile orders)
    const int* ft_seg;        // [ft_tiles+1] first segment of a tile
    const int4* ft_hdr;       // [2*ft_tiles] (p0, points, o0, observations), (s0, segments, first track row, longest track)
    const int* ft_rs;         // per tile: start of track row i in the transposed point-side scratch layout
    const double2* tc_uv;     // [n_obs] tile-camera order
    const int2* tc_meta;      // [n_obs] (camera slot, local point | point-side scratch slot << 10)
    const int* pm_tc;         // [n_obs] point-major index -> tile-camera index
    double2* tc_J;            // [2][3][n_obs] Jacobian cache W^1/2 Jp of the linearisation at parameter buffer b:
                              // planes (jx0, jx1), (jx2, jy0), (jy1, jy2), tile-camera order
    double2* tc_r;            // [2][n_obs] rho = W^-1/2 w_g r of the same linearisation (tile-camera order); only allocated for small
                              // problems (dense reduced system, ba_dense.cuh), nullptr otherwise
    const int* seg_cam;       // [n_seg]
    const int* seg_info;      // [n_seg] first tile-camera position inside the tile | length << 16 (longest first per tile)
    const int* seg_row;       // [n_seg] row slot of a segment (camera-major); accumulator slot inside its epoch when ep_target > 0
    const int* cam_row_off;   // [n_cams+2]
    double* rows;             // [n_seg*6]
    // epoch accumulators (ba_tiles.cuh): a CTA adds the segment sums of ep_len consecutive tiles in shared memory and
    // writes one row per (epoch, camera, split) instead of one per segment
    int ep_target;            // tiles per epoch aimed at (0 = one row per segment)
    const int* cta_ep0;       // [grid] first epoch of a CTA
    const int* ep_g0;         // [epochs+1 (+2 slack)] first accumulator slot (= row) of an epoch
    const int* slot_row;      // [slots] row of an accumulator slot (rows are camera-major)
    // partial sums
    int cam_in_backsub;  // dense path: the back-substitution pass carries one extra CTA that moves the cameras (k_tile_pass<1>)
    int lin_gate;        // device-side LM loop: the camera-side linearisation kernels return at once after a rejected step
                         // (the host loop does not launch them then)
    double* cm_partial;  // [cm_tiles*27]
    double* pm_partial;  // [n_partial_rows*NSCAL]
    double* cam_partial; // [n_cams*NSCAL]
    LMState* st;
    TraceRec* trace;     // [TRACE_CAP]
};

// position of point p in parameter buffer b
BA_D const double* pt_base(const Params& P, int b) { return P.pt_X + (size_t)b * P.pt_stride; }
BA_D void load_pt(const Params& P, int b, int p, double X[3]) {
    const double* base = pt_base(P, b);
    const double2 a = __ldg(reinterpret_cast<const double2*>(base) + p);
    X[0] = a.x; X[1] = a.y; X[2] = __ldg(base + 2 * (size_t)P.n_pad + p);
}
BA_D void store_pt(const Params& P, int b, int p, double x0, double x1, double x2) {
    double* base = P.pt_X + (size_t)b * P.pt_stride;
    reinterpret_cast<double2*>(base)[p] = make_double2(x0, x1);
    base[2 * (size_t)P.n_pad + p] = x2;
}

// ---------------------------------------------------------------------------------------------
// reductions (fixed trees -> deterministic)
// ---------------------------------------------------------------------------------------------
BA_D double warp_sum(double v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
    return v;
}
BA_D double warp_max(double v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, d));
    return v;
}

// Sum N per-thread values over the CTA; thread k < N gets total k in out[k] (shared, N doubles).
template <int N, int THREADS>
BA_D void block_sum(double (&v)[N], double* sm /* [THREADS/32 * N] */, double* out /* [N] */) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < N; ++k) {
        const double s = warp_sum(v[k]);
        if (lane == 0) sm[w * N + k] = s;
    }
    __syncthreads();
    if (threadIdx.x < N) {
        double s = 0.0;
#pragma unroll
        for (int i = 0; i < THREADS / 32; ++i) s += sm[i * N + threadIdx.x];
        out[threadIdx.x] = s;
    }
}

// ---------------------------------------------------------------------------------------------
// loads
// ---------------------------------------------------------------------------------------------
BA_D void load_cam(const double* __restrict__ base, int slot, double Rt[12]) {
    const double2* p = reinterpret_cast<const double2*>(base + (size_t)slot * 12);
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const double2 a = __ldg(p + k);
        Rt[2 * k] = a.x;
        Rt[2 * k + 1] = a.y;
    }
}
BA_D void load_vec6(const double* __restrict__ base, int slot, double x[6]) {
    const double2* p = reinterpret_cast<const double2*>(base + (size_t)slot * 6);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double2 a = __ldg(p + k);
        x[2 * k] = a.x;
        x[2 * k + 1] = a.y;
    }
}
// 3-vector stored in a 4-double slot at `ptr`
BA_D void load_v3(const double* __restrict__ ptr, double X[3]) {
    const double2* q = reinterpret_cast<const double2*>(ptr);
    const double2 a = __ldg(q), b = __ldg(q + 1);
    X[0] = a.x; X[1] = a.y; X[2] = b.x;
}
BA_D void store_v3(double* ptr, double a, double b, double c) {
    double2* q = reinterpret_cast<double2*>(ptr);
    q[0] = make_double2(a, b);
    q[1] = make_double2(c, 0.0);
}
BA_D void load_sym3(const double* __restrict__ base, int p, double S[6]) {
    const double2* q = reinterpret_cast<const double2*>(base + (size_t)p * 6);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double2 a = __ldg(q + k);
        S[2 * k] = a.x;
        S[2 * k + 1] = a.y;
    }
}

// cp.async (LDGSTS): 16 bytes global -> shared, no register staging.  .ca keeps the line in L1 (camera
// records are re-used by many warps), .cg bypasses it (point records are used a handful of times).
BA_D void cp_async16_ca(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}
BA_D void cp_async16_cg(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}
BA_D void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// Cooperative gather of 32 camera records into `stage` (per-warp shared memory, 32 * NCH * 16 bytes).
// Record of lane L = NCH 16-byte chunks: chunks [0,6) = R|t of camera `cam` (from camR), chunks [6,9) =
// its 6-vector (from vec) when NCH == 9.  Chunk j of the warp's 32*NCH is copied by lane j % 32.
template <int NCH>
BA_D void warp_gather_cameras(char* stage, int cam, const double* __restrict__ camR, const double* __restrict__ vec) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
        const int j = 32 * k + lane;
        const int src = j / NCH, part = j - src * NCH;
        const int c = __shfl_sync(0xffffffffu, cam, src);
        const char* g = (NCH == 6 || part < 6) ? reinterpret_cast<const char*>(camR) + (size_t)c * 96 + part * 16
                                               : reinterpret_cast<const char*>(vec) + (size_t)c * 48 + (part - 6) * 16;
        cp_async16_ca(stage + j * 16, g);
    }
}
template <int NCH>
BA_D void read_staged(const char* stage, double* out /* [2*NCH] */) {
    const double2* p = reinterpret_cast<const double2*>(stage + (threadIdx.x & 31) * (NCH * 16));
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
        const double2 a = p[k];
        out[2 * k] = a.x;
        out[2 * k + 1] = a.y;
    }
}

// ---------------------------------------------------------------------------------------------
// camera preparation: R(rvec) for buffer `which` (0/1 absolute, 2 = current, 3 = trial)
// ---------------------------------------------------------------------------------------------
BA_D int resolve_buf(const LMState* st, int which) { return which < 2 ? which : (which == 2 ? st->cur : st->cur ^ 1); }

__global__ void k_cam_prepare(Params P, int which) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c > P.n_cams) return;
    const int b = resolve_buf(P.st, which);
    const double* rt = P.cam_rt + (size_t)b * (P.n_cams + 1) * 6 + (size_t)c * 6;
    double* out = P.cam_R + (size_t)b * (P.n_cams + 1) * 12 + (size_t)c * 12;
    double R[9];
    rodrigues(rt, R);
#pragma unroll
    for (int k = 0; k < 9; ++k) out[k] = R[k];
    out[9] = rt[3]; out[10] = rt[4]; out[11] = rt[5];
}

// ---------------------------------------------------------------------------------------------
// API kernels, reference observation order
// ---------------------------------------------------------------------------------------------
__global__ void k_residuals(Params P, double2* __restrict__ r_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n_obs) return;
    const int b = P.st->cur;
    const double2 uv = __ldg(P.o_uv + i);
    const int2 id = __ldg(P.o_idx + i);
    double Rt[12], X[3];
    load_cam(P.cam_R + (size_t)b * (P.n_cams + 1) * 12, id.x, Rt);
    load_pt(P, b, id.y, X);
    double rx, ry;
    project_residual(Rt, X, uv.x, uv.y, P.K, rx, ry);
    r_out[i] = make_double2(rx, ry);
}

// Jacobian blocks in the reference's (rvec, t, X) parameterisation:
//   d r / d rvec = Jp (-[X]x) Jr(rvec),   d r / d t = Jp R^T = -dpi,   d r / d X = Jp.
__global__ void k_jacobian_blocks(Params P, double* __restrict__ jc_out, double* __restrict__ jp_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n_obs) return;
    const int b = P.st->cur;
    const double2 uv = __ldg(P.o_uv + i);
    const int2 id = __ldg(P.o_idx + i);
    double Rt[12], X[3];
    load_cam(P.cam_R + (size_t)b * (P.n_cams + 1) * 12, id.x, Rt);
    load_pt(P, b, id.y, X);
    ObsLin o;
    linearize_obs(Rt, X, uv.x, uv.y, P.K, o);
    double* jp = jp_out + (size_t)i * 6;
    double* jc = jc_out + (size_t)i * 12;
#pragma unroll
    for (int k = 0; k < 3; ++k) { jp[k] = o.jx[k]; jp[3 + k] = o.jy[k]; }
    if (id.x >= P.n_cams) {
#pragma unroll
        for (int k = 0; k < 12; ++k) jc[k] = 0.0;
        return;
    }
    const double* rvec = P.cam_rt + (size_t)b * (P.n_cams + 1) * 6 + (size_t)id.x * 6;
    double Jr[9], ax[3], ay[3];
    right_jacobian(rvec, Jr);
    cross3(X, o.jx, ax);  // row of Jp (-[X]x) = (X x j)^T
    cross3(X, o.jy, ay);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        jc[k] = ax[0] * Jr[k] + ax[1] * Jr[3 + k] + ax[2] * Jr[6 + k];
        jc[6 + k] = ay[0] * Jr[k] + ay[1] * Jr[3 + k] + ay[2] * Jr[6 + k];
        jc[3 + k] = o.jx[0] * Rt[3 * k] + o.jx[1] * Rt[3 * k + 1] + o.jx[2] * Rt[3 * k + 2];  // Jp R^T
        jc[9 + k] = o.jy[0] * Rt[3 * k] + o.jy[1] * Rt[3 * k + 1] + o.jy[2] * Rt[3 * k + 2];
    }
}

// ---------------------------------------------------------------------------------------------
// point-major tile kernels
// ---------------------------------------------------------------------------------------------
struct PmTile { int p0, p1, o0, n; };
BA_D PmTile pm_tile(const Params& P) {
    PmTile t;
    t.p0 = __ldg(P.pm_tile_pt + blockIdx.x);
    t.p1 = __ldg(P.pm_tile_pt + blockIdx.x + 1);
    t.o0 = __ldg(P.pt_off + t.p0);
    t.n = __ldg(P.pt_off + t.p1) - t.o0;
    return t;
}
// dynamic shared memory of a point-major kernel: NROW contribution rows of pm_cap doubles, then one
// cooperative-gather stage per warp
template <int NROW, int NCH>
__host__ __device__ constexpr size_t pm_smem_bytes(int cap) {
    return (size_t)NROW * cap * sizeof(double);
}

// K-LINP: linearise at buffer `which` in point-major order.
//   per point : V_p = sum Jp^T W Jp (6), g_p = sum Jp^T wg r (3)           -> pt_V[b], pt_g[b]
//   per tile  : 0.5*sum rho, sum r^2, max|g_p|, #bad                       -> pm_partial
__global__ void __launch_bounds__(PM_THREADS) k_lin_points(Params P, int which) {
    extern __shared__ __align__(16) double sm[];
    __shared__ double red[PM_THREADS / 32 * 3];
    __shared__ double tot[3];
    const int b = resolve_buf(P.st, which);
    const PmTile T = pm_tile(P);
    const int cap = P.pm_cap;
    const double* camR = P.cam_R + (size_t)b * (P.n_cams + 1) * 12;
    double acc[3] = {0.0, 0.0, 0.0};  // cost, raw, bad
    for (int base = 0; base < T.n; base += PM_THREADS) {
        const int i = base + threadIdx.x;
        const bool act = i < T.n;
        double2 uv = make_double2(0.0, 0.0);
        int2 id = make_int2(P.n_cams, T.p0);
        if (act) { uv = __ldg(P.pm_uv + T.o0 + i); id = __ldg(P.pm_idx + T.o0 + i); }
        if (!act) continue;
        double X[3], Rt[12];
        load_cam(camR, id.x, Rt);
        load_pt(P, b, id.y, X);
        ObsLin o;
        linearize_obs(Rt, X, uv.x, uv.y, P.K, o);
        double wHx, wgx, r0x, wHy, wgy, r0y;
        robust_weights(o.rx, P.delta, P.mode, wHx, wgx, r0x);
        robust_weights(o.ry, P.delta, P.mode, wHy, wgy, r0y);
        acc[0] += 0.5 * (r0x + r0y);
        acc[1] += o.rx * o.rx + o.ry * o.ry;
        if (!(isfinite(o.rx) && isfinite(o.ry))) acc[2] += 1.0;
        if (P.tc_J) {
            const double sx = sqrt(wHx), sy = sqrt(wHy);
            double2* J = P.tc_J + (size_t)b * 3 * P.n_obs + __ldg(P.pm_tc + T.o0 + i);
            J[0] = make_double2(sx * o.jx[0], sx * o.jx[1]);
            J[P.n_obs] = make_double2(sx * o.jx[2], sy * o.jy[0]);
            J[2 * (size_t)P.n_obs] = make_double2(sy * o.jy[1], sy * o.jy[2]);
        }
        const double ax = wHx * o.jx[0], bx = wHx * o.jx[1], cx = wHx * o.jx[2];
        const double ay = wHy * o.jy[0], by = wHy * o.jy[1], cy = wHy * o.jy[2];
        sm[0 * cap + i] = ax * o.jx[0] + ay * o.jy[0];
        sm[1 * cap + i] = ax * o.jx[1] + ay * o.jy[1];
        sm[2 * cap + i] = ax * o.jx[2] + ay * o.jy[2];
        sm[3 * cap + i] = bx * o.jx[1] + by * o.jy[1];
        sm[4 * cap + i] = bx * o.jx[2] + by * o.jy[2];
        sm[5 * cap + i] = cx * o.jx[2] + cy * o.jy[2];
        const double gx = wgx * o.rx, gy = wgy * o.ry;
        sm[6 * cap + i] = o.jx[0] * gx + o.jy[0] * gy;
        sm[7 * cap + i] = o.jx[1] * gx + o.jy[1] * gy;
        sm[8 * cap + i] = o.jx[2] * gx + o.jy[2] * gy;
    }
    __syncthreads();
    double gmax = 0.0;
    double* V = P.pt_V + (size_t)b * P.n_pts * 6;
    double* G = P.pt_g + (size_t)b * P.n_pts * 4;
    // one thread per (point, value): nine short sums side by side instead of one thread walking nine rows
    for (int w = threadIdx.x; w < (T.p1 - T.p0) * 9; w += PM_THREADS) {
        const int j = w / 9, k = w - 9 * j;
        const int p = T.p0 + j;
        const int a = __ldg(P.pt_off + p) - T.o0, e = __ldg(P.pt_off + p + 1) - T.o0;
        const double* row = sm + (size_t)k * cap;
        double sk = 0.0;
        for (int i = a; i < e; ++i) sk += row[i];
        if (k < 6) {
            V[(size_t)p * 6 + k] = sk;
        } else {
            G[(size_t)p * 4 + (k - 6)] = sk;
            gmax = fmax(gmax, fabs(sk));
        }
    }
    __syncthreads();
    block_sum<3, PM_THREADS>(acc, red, tot);
    __syncthreads();
    const double wm = warp_max(gmax);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = wm;
    __syncthreads();
    if (threadIdx.x == 0) {
        double m = 0.0;
        for (int i = 0; i < PM_THREADS / 32; ++i) m = fmax(m, red[i]);
        double* out = P.pm_partial + (size_t)blockIdx.x * NSCAL;
        out[S_COST] = tot[0]; out[S_RAW] = tot[1]; out[S_BAD] = tot[2]; out[S_GPMAX] = m;
    }
}

// K-PTPREP: V_d = V + lambda D_p, D_p = clamp(diag V); Vinv = V_d^-1; h = Vinv g_p.
BA_D void point_prepare_one(const Params& P, int p, double dmin, double dmax) {
    const int b = P.st->cur;
    const double lam = P.st->lambda;
    double V[6], g[3];
    load_sym3(P.pt_V + (size_t)b * P.n_pts * 6, p, V);
    load_v3(P.pt_g + (size_t)b * P.n_pts * 4 + (size_t)p * 4, g);
    const double d0 = fmin(fmax(V[0], dmin), dmax), d1 = fmin(fmax(V[3], dmin), dmax), d2 = fmin(fmax(V[5], dmin), dmax);
    V[0] += lam * d0; V[3] += lam * d1; V[5] += lam * d2;
    double Vi[6], h[3];
    if (!sym3_inverse(V, Vi)) {  // not SPD (can only come from non-finite input): damped diagonal instead
        Vi[0] = 1.0 / (lam * d0 + dmin); Vi[3] = 1.0 / (lam * d1 + dmin); Vi[5] = 1.0 / (lam * d2 + dmin);
        Vi[1] = Vi[2] = Vi[4] = 0.0;
    }
    sym3_mulv(Vi, g, h);
    double2* vo = reinterpret_cast<double2*>(P.pt_Vinv + (size_t)p * 6);
    vo[0] = make_double2(Vi[0], Vi[1]); vo[1] = make_double2(Vi[2], Vi[3]); vo[2] = make_double2(Vi[4], Vi[5]);
    store_v3(P.pt_h + (size_t)p * 4, h[0], h[1], h[2]);
    store_v3(P.pt_D + (size_t)p * 4, d0, d1, d2);
}
__global__ void k_point_prepare(Params P, double dmin, double dmax) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < P.n_pts) point_prepare_one(P, p, dmin, dmax);
}

// K-P1 (MODE 0): z_p = Vinv_p * sum_i Jp_i^T W_i Jc_i vec_c         (first half of the Schur mat-vec)
// K-BACKSUB (MODE 1): dp = -h_p - z_p(x), X_trial = X + dp, plus the step's scalar products.
template <int MODE>
__global__ void __launch_bounds__(PM_THREADS) k_point_pass(Params P, const double* __restrict__ vec) {
    extern __shared__ __align__(16) double sm[];
    __shared__ double red[PM_THREADS / 32 * 4];
    __shared__ double tot[4];
    if (MODE == 0 && P.st->pcg_done) return;
    const int b = P.st->cur;
    const PmTile T = pm_tile(P);
    const int cap = P.pm_cap;
    const double* camR = P.cam_R + (size_t)b * (P.n_cams + 1) * 12;
    for (int base = 0; base < T.n; base += PM_THREADS) {
        const int i = base + threadIdx.x;
        const bool act = i < T.n;
        double2 uv = make_double2(0.0, 0.0);
        int2 id = make_int2(P.n_cams, T.p0);
        if (act) { uv = __ldg(P.pm_uv + T.o0 + i); id = __ldg(P.pm_idx + T.o0 + i); }
        if (!act) continue;
        double X[3], rec[18];
        load_cam(camR, id.x, rec);
        load_vec6(vec, id.x, rec + 12);
        load_pt(P, b, id.y, X);
        const double* x = rec + 12;
        ObsLin o;
        linearize_obs(rec, X, uv.x, uv.y, P.K, o);
        double wHx, wgx, r0x, wHy, wgy, r0y;
        robust_weights(o.rx, P.delta, P.mode, wHx, wgx, r0x);
        robust_weights(o.ry, P.delta, P.mode, wHy, wgy, r0y);
        // d = w x X + v
        const double d0 = fma(x[1], X[2], fma(-x[2], X[1], x[3]));
        const double d1 = fma(x[2], X[0], fma(-x[0], X[2], x[4]));
        const double d2 = fma(x[0], X[1], fma(-x[1], X[0], x[5]));
        const double ax = wHx * (o.jx[0] * d0 + o.jx[1] * d1 + o.jx[2] * d2);
        const double ay = wHy * (o.jy[0] * d0 + o.jy[1] * d1 + o.jy[2] * d2);
        sm[0 * cap + i] = o.jx[0] * ax + o.jy[0] * ay;
        sm[1 * cap + i] = o.jx[1] * ax + o.jy[1] * ay;
        sm[2 * cap + i] = o.jx[2] * ax + o.jy[2] * ay;
    }
    __syncthreads();
    double acc[4] = {0.0, 0.0, 0.0, 0.0};  // g.dp, dp D dp, |dp|^2, |X|^2
    for (int j = threadIdx.x; j < T.p1 - T.p0; j += PM_THREADS) {
        const int p = T.p0 + j;
        const int a = __ldg(P.pt_off + p) - T.o0, e = __ldg(P.pt_off + p + 1) - T.o0;
        double s[3] = {0.0, 0.0, 0.0};
        for (int i = a; i < e; ++i) {
            s[0] += sm[i]; s[1] += sm[cap + i]; s[2] += sm[2 * cap + i];
        }
        double Vi[6], z[3];
        load_sym3(P.pt_Vinv, p, Vi);
        sym3_mulv(Vi, s, z);
        if (MODE == 0) {
            store_v3(P.pt_z + (size_t)p * 4, z[0], z[1], z[2]);
        } else {
            double h[3], X[3], g[3], D[3];
            load_v3(P.pt_h + (size_t)p * 4, h);
            load_pt(P, b, p, X);
            load_v3(P.pt_g + (size_t)b * P.n_pts * 4 + (size_t)p * 4, g);
            load_v3(P.pt_D + (size_t)p * 4, D);
            const double dp0 = -h[0] - z[0], dp1 = -h[1] - z[1], dp2 = -h[2] - z[2];
            store_pt(P, b ^ 1, p, X[0] + dp0, X[1] + dp1, X[2] + dp2);
            acc[0] += g[0] * dp0 + g[1] * dp1 + g[2] * dp2;
            acc[1] += D[0] * dp0 * dp0 + D[1] * dp1 * dp1 + D[2] * dp2 * dp2;
            acc[2] += dp0 * dp0 + dp1 * dp1 + dp2 * dp2;
            acc[3] += X[0] * X[0] + X[1] * X[1] + X[2] * X[2];
        }
    }
    if (MODE == 1) {
        __syncthreads();
        block_sum<4, PM_THREADS>(acc, red, tot);
        __syncthreads();
        if (threadIdx.x == 0) {
            double* out = P.pm_partial + (size_t)blockIdx.x * NSCAL;
            out[S_GDP] = tot[0]; out[S_DDP] = tot[1]; out[S_DP2] = tot[2]; out[S_XP2] = tot[3];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// camera-major tile kernels.  KIND 0: U (21) + g_c (6).  KIND 1: Schur right-hand side (6) +
// preconditioner correction (21).  KIND 2: second half of the Schur mat-vec (6).
// Accumulator layout: [0..20] symmetric 6x6 upper triangle (sym6), [21..26] 6-vector.  KIND 2 uses [0..5].
// ---------------------------------------------------------------------------------------------
template <int KIND> struct CmN { static constexpr int value = (KIND == 2) ? 6 : 27; };
// 16-byte chunks gathered per observation: X (2) [+ z (2)] [+ h (2) + Vinv (3)]
template <int KIND> struct CmCh { static constexpr int value = (KIND == 0) ? 2 : (KIND == 2 ? 4 : 7); };
template <int KIND>
__host__ __device__ constexpr size_t cm_smem_bytes() {
    return 0;
}

// WARP = false: one CTA per tile, long tiles (thousands of observations of one camera), CTA reduction of the N sums.
// WARP = true : one warp per tile (<= CM_TILE observations): the N sums of a tile are a shuffle tree, no shared memory,
//               no barrier - faster when a camera has only a few hundred observations (many cameras, or a rank of a
//               multi-GPU run that holds 1/N of every camera's observations).
template <int KIND, bool WARP, int PRE = 1, int THREADS = CM_THREADS>
__global__ void __launch_bounds__(THREADS, 2) k_camera_pass(Params P, const double* __restrict__ vec, int with_schur_diag) {
    constexpr int N = CmN<KIND>::value;
    constexpr int NPRE = (KIND == 0) ? 3 : ((KIND == 1 && PRE == 2) ? 12 : 6);  // doubles gathered one trip ahead: X [+ z | h] [+ Vinv]
    __shared__ double red[WARP ? 1 : THREADS / 32 * N];
    __shared__ double tot[WARP ? 1 : N];
    // the camera's pose in shared memory (broadcast reads), not in 24 registers of every thread: with the 27 accumulators and
    // the gathers of two trips in flight the kernel otherwise spills at its 128 registers (round 1: STACK 80, LDL 22 / STL 24)
    __shared__ double sRt[(WARP ? THREADS / 32 : 1) * 12];
    if (KIND == 2 && P.st->pcg_done) return;
    if (KIND == 0 && P.lin_gate && lin_gated(P.st)) return;
    const int lane = WARP ? (threadIdx.x & 31) : threadIdx.x;
    constexpr int STRIDE = WARP ? 32 : THREADS;
    const int tile_id = WARP ? blockIdx.x * (THREADS / 32) + (threadIdx.x >> 5) : blockIdx.x;
    if (WARP && tile_id >= P.cm_tiles) return;
    const int b = P.st->cur;
    const int4 tile = __ldg(P.cm_tile + tile_id);
    const int cam = tile.x, o0 = tile.y, n = tile.z;
    // the records are fetched ahead: the point index two trips ahead, because it is the address of the gathers, and the
    // gathers themselves (point, and z or h) one trip ahead - a trip is about 200 FP64 instructions, a gather a round trip to L2
    int p_cur = 0, p_next = 0;
    if (lane < n) p_cur = __ldg(P.cm_pt + o0 + lane);
    if (lane + STRIDE < n) p_next = __ldg(P.cm_pt + o0 + lane + STRIDE);
    double* Rt = sRt + (WARP ? (threadIdx.x >> 5) * 12 : 0);
    if (lane < 12) Rt[lane] = __ldg(P.cam_R + (size_t)b * (P.n_cams + 1) * 12 + (size_t)cam * 12 + lane);
    double x[6];
    if (KIND == 2) load_vec6(vec, cam, x);
    double acc[N];
#pragma unroll
    for (int k = 0; k < N; ++k) acc[k] = 0.0;
    auto gather = [&](int p, double* rec) {
        load_pt(P, b, p, rec);
        if constexpr (KIND == 2) load_v3(P.pt_z + (size_t)p * 4, rec + 3);
        if constexpr (KIND == 1) load_v3(P.pt_h + (size_t)p * 4, rec + 3);
        if constexpr (KIND == 1 && PRE == 2) load_sym3(P.pt_Vinv, p, rec + 6);
    };
    double2 uv_next = make_double2(0.0, 0.0);
    double rec_next[NPRE];
#pragma unroll
    for (int k = 0; k < NPRE; ++k) rec_next[k] = 0.0;
    if (lane < n) { uv_next = __ldg(P.cm_uv + o0 + lane); gather(p_cur, rec_next); }
    if (WARP) __syncwarp(); else __syncthreads();
    for (int i = lane; i < n; i += STRIDE) {
        const double2 uvr = uv_next;
        double rec[NPRE];
#pragma unroll
        for (int k = 0; k < NPRE; ++k) rec[k] = rec_next[k];
        double Vi[6];
        if constexpr (KIND == 1 && PRE != 2) load_sym3(P.pt_Vinv, p_cur, Vi);  // needed late in the trip
        if constexpr (KIND == 1 && PRE == 2) {
#pragma unroll
            for (int k = 0; k < 6; ++k) Vi[k] = rec[6 + k];
        }
        p_cur = p_next;
        if (i + 2 * STRIDE < n) p_next = __ldg(P.cm_pt + o0 + i + 2 * STRIDE);
        if (i + STRIDE < n) { uv_next = __ldg(P.cm_uv + o0 + i + STRIDE); gather(p_cur, rec_next); }
        double2 uv[1] = {uvr};
        constexpr int r = 0;
        const double* X = rec;
        ObsLin o;
        linearize_obs(Rt, X, uv[r].x, uv[r].y, P.K, o);
        double wHx, wgx, r0x, wHy, wgy, r0y;
        robust_weights(o.rx, P.delta, P.mode, wHx, wgx, r0x);
        robust_weights(o.ry, P.delta, P.mode, wHy, wgy, r0y);
        if constexpr (KIND == 0) {
            double jcx[6], jcy[6];  // rows of Jc = (X x j, j)
            cross3(X, o.jx, jcx); cross3(X, o.jy, jcy);
#pragma unroll
            for (int k = 0; k < 3; ++k) { jcx[3 + k] = o.jx[k]; jcy[3 + k] = o.jy[k]; }
            const double gx = wgx * o.rx, gy = wgy * o.ry;
            int q = 0;
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const double sx = wHx * jcx[a], sy = wHy * jcy[a];
#pragma unroll
                for (int c = a; c < 6; ++c) { acc[q] = fma(sx, jcx[c], fma(sy, jcy[c], acc[q])); ++q; }  // two FMAs, not mul + fma + add
                acc[21 + a] = fma(jcx[a], gx, fma(jcy[a], gy, acc[21 + a]));
            }
        } else if constexpr (KIND == 1) {
            double W[6];  // W = Jp^T diag(wH) Jp (sym 3x3)
            {
                const double ax = wHx * o.jx[0], bx = wHx * o.jx[1], cx = wHx * o.jx[2];
                const double ay = wHy * o.jy[0], by = wHy * o.jy[1], cy = wHy * o.jy[2];
                W[0] = ax * o.jx[0] + ay * o.jy[0]; W[1] = ax * o.jx[1] + ay * o.jy[1]; W[2] = ax * o.jx[2] + ay * o.jy[2];
                W[3] = bx * o.jx[1] + by * o.jy[1]; W[4] = bx * o.jx[2] + by * o.jy[2]; W[5] = cx * o.jx[2] + cy * o.jy[2];
            }
            const double* h = rec + 3;
            double m[3], xm[3];
            sym3_mulv(W, h, m);  // E h = G^T (W h)
            cross3(X, m, xm);
#pragma unroll
            for (int k = 0; k < 3; ++k) { acc[21 + k] += xm[k]; acc[24 + k] += m[k]; }
            if (with_schur_diag) {
                // T = W Vinv W (sym), correction = G^T T G with G = [-[X]x | I]
                const double Wf[9] = {W[0], W[1], W[2], W[1], W[3], W[4], W[2], W[4], W[5]};
                double A[9];  // A = Vinv W
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const double col[3] = {Wf[c], Wf[3 + c], Wf[6 + c]};
                    double r3[3];
                    sym3_mulv(Vi, col, r3);
                    A[c] = r3[0]; A[3 + c] = r3[1]; A[6 + c] = r3[2];
                }
                double Tm[9];  // T = W A (symmetric)
#pragma unroll
                for (int a = 0; a < 3; ++a)
#pragma unroll
                    for (int c = 0; c < 3; ++c) Tm[3 * a + c] = Wf[3 * a] * A[c] + Wf[3 * a + 1] * A[3 + c] + Wf[3 * a + 2] * A[6 + c];
                double C[9];  // C = [X]x T : column c = X x T[:,c]
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const double tc[3] = {Tm[c], Tm[3 + c], Tm[6 + c]};
                    double cc[3];
                    cross3(X, tc, cc);
                    C[c] = cc[0]; C[3 + c] = cc[1]; C[6 + c] = cc[2];
                }
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    // top-left block [X]x T [X]x^T: row a = X x C[a,:]
                    const double cr[3] = {C[3 * a], C[3 * a + 1], C[3 * a + 2]};
                    double tl[3];
                    cross3(X, cr, tl);
#pragma unroll
                    for (int c = a; c < 3; ++c) acc[sym6(a, c)] += tl[c];
#pragma unroll
                    for (int c = 0; c < 3; ++c) acc[sym6(a, 3 + c)] += C[3 * a + c];  // top-right = [X]x T
                }
#pragma unroll
                for (int a = 0; a < 3; ++a)
#pragma unroll
                    for (int c = a; c < 3; ++c) acc[sym6(3 + a, 3 + c)] += Tm[3 * a + c];
            }
        } else {
            const double* z = rec + 3;
            const double d0 = fma(x[1], X[2], fma(-x[2], X[1], x[3])) - z[0];
            const double d1 = fma(x[2], X[0], fma(-x[0], X[2], x[4])) - z[1];
            const double d2 = fma(x[0], X[1], fma(-x[1], X[0], x[5])) - z[2];
            const double bx = wHx * (o.jx[0] * d0 + o.jx[1] * d1 + o.jx[2] * d2);
            const double by = wHy * (o.jy[0] * d0 + o.jy[1] * d1 + o.jy[2] * d2);
            double m[3], xm[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) m[k] = o.jx[k] * bx + o.jy[k] * by;
            cross3(X, m, xm);
#pragma unroll
            for (int k = 0; k < 3; ++k) { acc[k] += xm[k]; acc[3 + k] += m[k]; }
        }
    }
    double* out = P.cm_partial + (size_t)tile_id * 27;
    if constexpr (WARP) {
#pragma unroll
        for (int k = 0; k < N; ++k) {
            const double tsum = warp_sum(acc[k]);  // fixed shuffle order; total in lane 0
            if (lane == 0) out[k] = tsum;
        }
    } else {
        block_sum<N, THREADS>(acc, red, tot);
        __syncthreads();
        if (threadIdx.x < N) out[threadIdx.x] = tot[threadIdx.x];
    }
}

// FP64 FMA rate: 16 independent chains per thread, nothing else in the loop (ba_fp64_peak).
__global__ void __launch_bounds__(256) k_fp64_peak(double* __restrict__ sink, int iters, double m) {
    double a[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) a[k] = 1e-3 * (threadIdx.x + k);
    const double c = 1e-7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k) a[k] = fma(a[k], m, c);
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 16; ++k) s += a[k];
    sink[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Sum the tile partials of each camera in tile order: cam_acc[c*n_val + k] (k < n_val).
// gated: 1 = part of a PCG iteration, 2 = part of the camera-side linearisation (device-side LM loop).
__global__ void k_camera_sum(Params P, int n_val, int gated) {
    if (gated == 1 && P.st->pcg_done) return;
    if (gated == 2 && P.lin_gate && lin_gated(P.st)) return;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.n_cams * n_val) return;
    const int c = idx / n_val, k = idx - c * n_val;
    const int t0 = __ldg(P.cam_tile_off + c), t1 = __ldg(P.cam_tile_off + c + 1);
    double s = 0.0;
    for (int t = t0; t < t1; ++t) s += P.cm_partial[(size_t)t * 27 + k];
    P.cam_acc[(size_t)c * n_val + k] = s;
}

// After K-LINC (+ all-reduce): store U, g_c, D_c = clamp(diag U).
__global__ void k_camera_finish_lin(Params P, double dmin, double dmax) {
    if (P.lin_gate && lin_gated(P.st)) return;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P.n_cams) return;
    const double* a = P.cam_acc + (size_t)c * 27;
    for (int k = 0; k < 21; ++k) P.cam_U[(size_t)c * 21 + k] = a[k];
    for (int k = 0; k < 6; ++k) {
        P.cam_g[(size_t)c * 6 + k] = a[21 + k];
        P.cam_D[(size_t)c * 6 + k] = fmin(fmax(a[sym6(k, k)], dmin), dmax);
    }
}

// ---------------------------------------------------------------------------------------------
// camera-vector kernels (one camera per thread, grid = vec_ctas CTAs of VEC_THREADS)
//
// Preconditioned CG in the single-reduction form of Chronopoulos & Gear, so that the two inner
// products of an iteration are available right after the mat-vec and the vector updates need no
// grid-wide barrier:
//      u = M^-1 r,  w = A u,  gamma = r.u,  delta = w.u
//      beta = gamma / gamma_prev (0 at first),  alpha = gamma / (delta - beta * gamma / alpha_prev)
//      p = u + beta p,  s = w + beta s,  x += alpha p,  r -= alpha s
// Inner products: every CTA writes its partial, all CTAs (k_pcg_update) or the last CTA to arrive
// (an integer ticket; the sum itself is done by one warp in a fixed order) add them up.
// ---------------------------------------------------------------------------------------------
BA_D double cta_sum_vec(double v, double* sm /* >= VEC_THREADS/32 */) {  // result valid in thread 0
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    double s = 0.0;
    if (threadIdx.x == 0)
        for (int i = 0; i < VEC_THREADS / 32; ++i) s += sm[i];
    return s;
}
// sum of n per-CTA partials (stride 2 doubles) by warp 0, fixed order; result in lane 0 of warp 0
BA_D double warp0_sum_partials(const double* part, int n) {
    double v = 0.0;
    for (int i = threadIdx.x; i < n; i += 32) v += part[2 * i];
    return warp_sum(v);
}

BA_D void minv_apply(const double* __restrict__ Minv, const double* __restrict__ r, double* z) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < 6; ++j) s += Minv[6 * i + j] * r[j];
        z[i] = s;
    }
}

// After K-SCHURC (+ all-reduce): M_c = U_c + lambda D_c - corr_c, Minv_c; b_c = -g_c + (E Vinv g_p)_c;
// PCG start: x = 0, r = b, u = Minv r, p = s = 0, gamma = r.u.
__global__ void __launch_bounds__(VEC_THREADS) k_camera_finish_schur(Params P, int with_schur_diag, int* ticket,
                                                                     unsigned long long pcg_loop) {
    __shared__ double sm[VEC_THREADS / 32];
    __shared__ int last;
    const int c = blockIdx.x * VEC_THREADS + threadIdx.x;
    double part = 0.0;
    if (c < P.n_cams) {
        const double lam = P.st->lambda;
        const double* a = P.cam_acc + (size_t)c * 27;
        const double* U = P.cam_U + (size_t)c * 21;
        double M[36];
        for (int i = 0; i < 6; ++i)
            for (int j = i; j < 6; ++j) {
                double v = U[sym6(i, j)] - (with_schur_diag ? a[sym6(i, j)] : 0.0);
                if (i == j) v += lam * P.cam_D[(size_t)c * 6 + i];
                M[6 * i + j] = v; M[6 * j + i] = v;
            }
        if (!spd6_inverse(M)) {  // numerically indefinite: damped block of J^T J, then its diagonal
            for (int i = 0; i < 6; ++i)
                for (int j = i; j < 6; ++j) {
                    double v = U[sym6(i, j)];
                    if (i == j) v += lam * P.cam_D[(size_t)c * 6 + i];
                    M[6 * i + j] = v; M[6 * j + i] = v;
                }
            if (!spd6_inverse(M)) {
                for (int i = 0; i < 36; ++i) M[i] = 0.0;
                for (int i = 0; i < 6; ++i) M[7 * i] = 1.0 / ((1.0 + lam) * P.cam_D[(size_t)c * 6 + i]);
            }
        }
        for (int k = 0; k < 36; ++k) P.cam_Minv[(size_t)c * 36 + k] = M[k];
        double r[6], u[6];
        for (int k = 0; k < 6; ++k) r[k] = a[21 + k] - P.cam_g[(size_t)c * 6 + k];
        minv_apply(M, r, u);
        for (int k = 0; k < 6; ++k) {
            const size_t i = (size_t)c * 6 + k;
            P.v_b[i] = r[k]; P.v_r[i] = r[k]; P.v_u[i] = u[k];
            P.v_x[i] = 0.0; P.v_p[i] = 0.0; P.v_s[i] = 0.0;
            part += r[k] * u[k];
        }
    }
    const double cs = cta_sum_vec(part, sm);
    if (threadIdx.x == 0) {
        P.dot_partial[2 * blockIdx.x + 1] = cs;
        __threadfence();
        last = (atomicAdd(ticket, 1) == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (last && threadIdx.x < 32) {
        __threadfence();
        const double g = warp0_sum_partials(P.dot_partial + 1, gridDim.x);
        if (threadIdx.x == 0) {
            LMState* st = P.st;
            st->gamma = g; st->gamma0 = g; st->gamma_prev = g; st->alpha = 0.0; st->beta = 0.0; st->pcg_iter = 0;
            st->pcg_done = !(g > 0.0);  // zero (or non-finite) right-hand side: nothing to solve
            *ticket = 0;
            set_loop(pcg_loop, g > 0.0);
        }
    }
}

// w = (summed mat-vec partials) + lambda D u;  per-CTA partial of delta = w.u.
// FROM_TILES: add the camera's tile partials here (single GPU); otherwise read cam_acc (all-reduced).
template <bool FROM_TILES>
__global__ void __launch_bounds__(VEC_THREADS) k_pcg_dots(Params P) {
    __shared__ double sm[VEC_THREADS / 32];
    if (P.st->pcg_done) return;
    const int c = blockIdx.x * VEC_THREADS + threadIdx.x;
    double part = 0.0;
    if (c < P.n_cams) {
        const double lam = P.st->lambda;
        double w[6];
        if (FROM_TILES) {
            const int t0 = __ldg(P.cam_tile_off + c), t1 = __ldg(P.cam_tile_off + c + 1);
#pragma unroll
            for (int k = 0; k < 6; ++k) w[k] = 0.0;
            for (int t = t0; t < t1; ++t)
#pragma unroll
                for (int k = 0; k < 6; ++k) w[k] += P.cm_partial[(size_t)t * 27 + k];
        } else {
#pragma unroll
            for (int k = 0; k < 6; ++k) w[k] = P.cam_acc[(size_t)c * 6 + k];
        }
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const size_t i = (size_t)c * 6 + k;
            const double u = P.v_u[i];
            const double wk = w[k] + lam * P.cam_D[i] * u;
            P.v_w[i] = wk;
            part += wk * u;
        }
    }
    const double cs = cta_sum_vec(part, sm);
    if (threadIdx.x == 0) P.dot_partial[2 * blockIdx.x] = cs;
}

__global__ void __launch_bounds__(VEC_THREADS) k_pcg_update(Params P, double tol, int max_iters, int n_dot_partials, int* ticket) {
    __shared__ double sm[VEC_THREADS / 32];
    __shared__ double sc[2];
    __shared__ int last;
    LMState* st = P.st;
    if (st->pcg_done) return;
    const double gamma = st->gamma, alpha_prev = st->alpha, gamma_prev = st->gamma_prev;
    const int first = (st->pcg_iter == 0);
    if (threadIdx.x < 32) {
        const double d = warp0_sum_partials(P.dot_partial, n_dot_partials);
        if (threadIdx.x == 0) {
            const double beta = first ? 0.0 : gamma / gamma_prev;
            const double den = first ? d : d - beta * gamma / alpha_prev;
            sc[0] = (den > 0.0) ? gamma / den : 0.0;
            sc[1] = beta;
        }
    }
    __syncthreads();
    const double alpha = sc[0], beta = sc[1];
    const int c = blockIdx.x * VEC_THREADS + threadIdx.x;
    double part = 0.0;
    if (c < P.n_cams) {
        double r[6], u[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const size_t i = (size_t)c * 6 + k;
            const double p = P.v_u[i] + beta * P.v_p[i];
            const double s = P.v_w[i] + beta * P.v_s[i];
            P.v_p[i] = p; P.v_s[i] = s;
            P.v_x[i] += alpha * p;
            r[k] = P.v_r[i] - alpha * s;
            P.v_r[i] = r[k];
        }
        minv_apply(P.cam_Minv + (size_t)c * 36, r, u);
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            P.v_u[(size_t)c * 6 + k] = u[k];
            part += r[k] * u[k];
        }
    }
    __syncthreads();
    const double cs = cta_sum_vec(part, sm);
    if (threadIdx.x == 0) {
        P.dot_partial[2 * blockIdx.x + 1] = cs;
        __threadfence();
        last = (atomicAdd(ticket, 1) == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (last && threadIdx.x < 32) {
        __threadfence();
        const double g = warp0_sum_partials(P.dot_partial + 1, gridDim.x);
        if (threadIdx.x == 0) {
            st->gamma_prev = gamma;
            st->gamma = g; st->alpha = alpha; st->beta = beta;
            const int it = st->pcg_iter + 1;
            st->pcg_iter = it; st->pcg_total += 1;
            if (!(alpha > 0.0) || !(g > tol * tol * st->gamma0) || it >= max_iters) st->pcg_done = 1;
            *ticket = 0;
        }
    }
}

// Trial camera poses from the PCG solution x: R <- R Exp(w), t <- t + R v; rvec = Log(R).
BA_D void camera_update_one(const Params& P, int c) {
    const int b = P.st->cur;
    const size_t nc1 = (size_t)P.n_cams + 1;
    const double* rt = P.cam_rt + (size_t)b * nc1 * 6 + (size_t)c * 6;
    double* rt_new = P.cam_rt + (size_t)(b ^ 1) * nc1 * 6 + (size_t)c * 6;
    double* R_new = P.cam_R + (size_t)(b ^ 1) * nc1 * 12 + (size_t)c * 12;
    const double* Rt = P.cam_R + (size_t)b * nc1 * 12 + (size_t)c * 12;
    if (c == P.n_cams) {  // fixed camera: copy
        for (int k = 0; k < 6; ++k) rt_new[k] = rt[k];
        for (int k = 0; k < 12; ++k) R_new[k] = Rt[k];
        return;
    }
    double x[6], E[9], Rn[9], rv[3];
    for (int k = 0; k < 6; ++k) x[k] = P.v_x[(size_t)c * 6 + k];
    rodrigues(x, E);
    mat3_mul(Rt, E, Rn);
    log_so3(Rn, rv);
    double tn[3];
    for (int k = 0; k < 3; ++k) tn[k] = Rt[9 + k] + Rt[3 * k] * x[3] + Rt[3 * k + 1] * x[4] + Rt[3 * k + 2] * x[5];
    rodrigues(rv, Rn);  // keep R consistent with the stored rvec (what ba_get_params returns)
    for (int k = 0; k < 3; ++k) { rt_new[k] = rv[k]; rt_new[3 + k] = tn[k]; }
    for (int k = 0; k < 9; ++k) R_new[k] = Rn[k];
    for (int k = 0; k < 3; ++k) R_new[9 + k] = tn[k];
    double* out = P.cam_partial + (size_t)c * NSCAL;
    double gd = 0, dd = 0, rd = 0, d2 = 0, x2 = 0;
    for (int k = 0; k < 6; ++k) {
        gd += P.cam_g[(size_t)c * 6 + k] * x[k];
        dd += P.cam_D[(size_t)c * 6 + k] * x[k] * x[k];
        rd += P.v_r[(size_t)c * 6 + k] * x[k];
        d2 += x[k] * x[k];
        x2 += rt[k] * rt[k];
    }
    out[S_GDC] = gd; out[S_DDC] = dd; out[S_RDC] = rd; out[S_DC2] = d2; out[S_XC2] = x2;
}
__global__ void k_camera_update(Params P) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c <= P.n_cams) camera_update_one(P, c);
}

// ---------------------------------------------------------------------------------------------
// single-CTA scalar kernels (once per LM iteration)
// ---------------------------------------------------------------------------------------------
constexpr int RED_THREADS = 1024;

BA_D double cta_sum_all(double v, double* sm /* >= 33 */) {  // all threads get the total
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sm[w] = v;
    __syncthreads();
    if (w == 0) {
        double s = (lane < (int)(blockDim.x >> 5)) ? sm[lane] : 0.0;
        s = warp_sum(s);
        if (lane == 0) sm[32] = s;
    }
    __syncthreads();
    return sm[32];
}

// Accept / reject the trial point, update the damping (Nielsen), test ftol / xtol.
struct DecideArgs { double ftol, xtol, lam_min, lam_max; int max_iters; unsigned long long lm_loop; };
BA_D void lm_decide_body(const Params& P, double ftol, double xtol, double lam_min, double lam_max, int max_iters,
                         unsigned long long lm_loop) {
    LMState* st = P.st;
    if (st->status != 0) {  // already terminated (gtol at the current point)
        set_loop(lm_loop, false);
        return;
    }
    const double* s = st->scal;
    const double lambda_used = st->lambda;  // the damping this iteration's step was computed with (trace)
    const int cur = st->cur, tri = cur ^ 1;
    const double cost_new = s[S_COST], raw_new = s[S_RAW];
    const bool finite_new = isfinite(cost_new) && !(s[S_BAD] > 0.0);
    const double gd = s[S_GDP] + s[S_GDC];
    const double dDd = s[S_DDP] + s[S_DDC];
    // exact decrease of the quadratic model for the inexact (PCG) step: the camera rows of
    // (H + lambda D) d + g equal -r_pcg, the point rows vanish (exact back-substitution)
    const double model = -0.5 * gd + 0.5 * s[S_RDC] + 0.5 * st->lambda * dDd;
    const double cost = st->cost[cur];
    const double dcost = cost - cost_new;
    const double rho = (model > 0.0 && finite_new) ? dcost / model : -1.0;
    const double step = sqrt(s[S_DP2] + s[S_DC2]);
    const double xn = sqrt(s[S_XP2] + s[S_XC2]);
    st->model = model; st->rho = rho; st->dcost = dcost; st->step_norm = step; st->x_norm = xn;
    st->iter += 1;
    const bool accept = finite_new && rho > 0.0 && dcost >= 0.0;
    if (accept) {
        st->cost[tri] = cost_new; st->raw[tri] = raw_new;
        st->cur = tri;
        st->accepted += 1; st->last_accepted = 1;
        const double t = 2.0 * rho - 1.0;
        st->lambda = fmax(st->lambda * fmax(1.0 / 3.0, 1.0 - t * t * t), lam_min);
        st->nu = 2.0;
        const bool f_ok = dcost <= ftol * cost;
        const bool x_ok = step < xtol * (xtol + xn);
        if (st->status == 0) {
            if (f_ok && x_ok) st->status = 4;
            else if (f_ok) st->status = 2;
            else if (x_ok) st->status = 3;
        }
    } else {
        st->rejected += 1; st->last_accepted = 0;
        st->lambda *= st->nu;
        st->nu *= 2.0;
        if (st->status == 0 && st->lambda > lam_max) st->status = 5;
    }
    if (st->status == 0 && st->iter >= max_iters) st->status = 6;  // iteration limit (reported as 0)
    if (P.trace && st->iter < TRACE_CAP)
        P.trace[st->iter] = TraceRec{st->cost[st->cur], st->raw[st->cur], 1e-6 * (double)(global_ns() - st->t0), lambda_used, st->pcg_total, accept ? 1 : 0};
    set_loop(lm_loop, st->status == 0);
}
__global__ void k_lm_decide(Params P, double ftol, double xtol, double lam_min, double lam_max, int max_iters,
                            unsigned long long lm_loop) {
    lm_decide_body(P, ftol, xtol, lam_min, lam_max, max_iters, lm_loop);
}

// Reduce per-tile / per-camera scalar partials into st->scal (fixed order).  The masks select slots.
// decide != 0 (single GPU, end of an LM iteration): thread 0 goes straight on to the accept / reject decision - one launch less.
// prep != 0 (small problems): once the decision is made the whole cluster prepares the points (V^-1, h, D at the new damping and
// the new current point) for the next iteration - another launch less.
__global__ void __launch_bounds__(RED_THREADS) k_reduce_scalars(Params P, unsigned pm_mask, unsigned cam_mask, int decide, DecideArgs da,
                                                                int prep, double dmin, double dmax) {
    // One thread-block cluster (1..16 CTAs): CTA r takes the row groups r, r + CL, ...; 16 lanes per partial row (one
    // 128-byte line): thread t owns slot t & 15.  CTA totals meet in CTA 0's shared memory and are added in rank order.
    static_assert(NSCAL == 16, "row layout");
    __shared__ double sm[RED_THREADS];
    __shared__ double part[16 * 16];  // [rank][slot], CTA 0 only
    cooperative_groups::cluster_group cl = cooperative_groups::this_cluster();
    const int R = (int)cl.block_rank(), CL = (int)cl.num_blocks();
    const int s = threadIdx.x & 15, g = threadIdx.x >> 4;
    constexpr int G = RED_THREADS / 16;
    const bool pm = (pm_mask >> s) & 1u, cm = (cam_mask >> s) & 1u, is_max = (s == S_GPMAX);
    double v = 0.0;
    if (pm) {
        for (int t = R * G + g; t < P.n_partial_rows; t += CL * G) {
            const double x = P.pm_partial[(size_t)t * NSCAL + s];
            v = is_max ? fmax(v, x) : v + x;
        }
    }
    if (cm) {
        for (int c = R * G + g; c < P.n_cams; c += CL * G) v += P.cam_partial[(size_t)c * NSCAL + s];
    }
    sm[g * 16 + s] = v;
    __syncthreads();
    if (threadIdx.x < 16) {  // fixed order over the 64 row groups
        double tot = 0.0;
        for (int i = 0; i < G; ++i) {
            const double x = sm[i * 16 + s];
            tot = is_max ? fmax(tot, x) : tot + x;
        }
        cl.map_shared_rank(part, 0)[R * 16 + s] = tot;
    }
    cl.sync();
    if (R == 0 && threadIdx.x < 16 && (pm || cm)) {
        double tot = 0.0;
        for (int r = 0; r < CL; ++r) {
            const double x = part[r * 16 + s];
            tot = is_max ? fmax(tot, x) : tot + x;
        }
        P.st->scal[s] = tot;
    }
    if (decide && R == 0 && threadIdx.x < 32) {
        __syncwarp();  // the sixteen totals of warp 0 are in place
        if (threadIdx.x == 0) lm_decide_body(P, da.ftol, da.xtol, da.lam_min, da.lam_max, da.max_iters, da.lm_loop);
    }
    if (prep) {
        __threadfence();
        cl.sync();  // the decision (current buffer, damping) is visible to the whole cluster
        for (int p = R * RED_THREADS + threadIdx.x; p < P.n_pts; p += CL * RED_THREADS) point_prepare_one(P, p, dmin, dmax);
    }
}

// gmax = max(|g_c|, |g_p|) at the current point; gtol test.
__global__ void __launch_bounds__(RED_THREADS) k_gradient_check(Params P, double gtol) {
    __shared__ double sm[32];
    if (P.lin_gate && lin_gated(P.st)) return;
    double v = 0.0;
    for (int i = threadIdx.x; i < P.n_cams * 6; i += RED_THREADS) v = fmax(v, fabs(P.cam_g[i]));
    v = warp_max(v);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double m = P.st->scal[S_GPMAX];
        for (int i = 0; i < RED_THREADS / 32; ++i) m = fmax(m, sm[i]);
        P.st->gmax = m;
        if (P.st->status == 0 && m < gtol) P.st->status = 1;
    }
}

// Record the cost of buffer `which` from st->scal (after k_lin_points + k_reduce_scalars).
__global__ void k_store_cost(Params P, int which) {
    LMState* st = P.st;
    const int b = resolve_buf(st, which);
    st->cost[b] = st->scal[S_COST];
    st->raw[b] = st->scal[S_RAW];
    if (st->scal[S_BAD] > 0.0 || !isfinite(st->scal[S_COST])) st->bad = 1;
    if (P.trace && st->iter == 0) P.trace[0] = TraceRec{st->cost[b], st->raw[b], 1e-6 * (double)(global_ns() - st->t0), st->lambda, 0, 0};
}

}  // namespace ba
