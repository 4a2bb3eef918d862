// ba_dense.cuh - the reduced camera system of a SMALL problem (at most DN_MAX_CAMS adjustable cameras: the sliding windows of
// the reference's pipeline, src/pipeline.py:99 with BA_WINDOW_SIZE = 5, and its final global BA over a handful of keyframes)
// assembled explicitly and solved by a dense Cholesky factorisation, instead of preconditioned CG:
//
//      S = blockdiag(U_c + lambda D_c) - sum_p  E_p (V_p + lambda D_p)^-1 E_p^T          (6 Nc x 6 Nc, Nc <= 16)
//
// With 24 .. 96 unknowns a PCG solve is 7 .. 20 iterations of two launch-bound kernels each and leaves an inexact step;
// ONE kernel here does the whole camera side of an LM iteration and gives the exact step.
//
// PAIRS.  Two cameras are coupled through every point both observe.  ba_set_problem lists all ordered pairs (i, j) of
// observations of adjustable cameras that share a point - i = j included, a camera that lists a point twice included - sorted
// by bin = camera(i) * Nc + camera(j).  One CTA per bin adds the 6x6 blocks of its pairs,
//      E_i Vinv E_j^T = C_i^T (Jt_i Vinv Jt_j^T) C_j,     C = Jt [-[X]x | I]  (2x6),   Jt = W^1/2 Jp  (the Jacobian cache),
// thread-strided with register accumulators, a shuffle tree and a fixed-order sum over the warps: no floating-point atomics,
// bit-reproducible.
#pragma once
#include "ba_tiles.cuh"

namespace ba {

constexpr int DN_MAX_CAMS = 16;           // 96 unknowns: the factorisation runs in one CTA's shared memory
constexpr int DN_MAX_OBS = 1 << 17;
constexpr int DN_MAX_PAIRS = 1 << 22;
constexpr int DN_THREADS = 256;

// number of pairs observation i (point-major index) opens: the adjustable observations of its point
__global__ void k_dn_pair_count(int n_obs, int n_cams, const int2* __restrict__ pm_idx, const int* __restrict__ pt_off,
                                int* __restrict__ cnt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_obs) return;
    int c = 0;
    if (i < n_obs && pm_idx[i].x < n_cams) {
        const int p = pm_idx[i].y;
        for (int j = pt_off[p]; j < pt_off[p + 1]; ++j) c += pm_idx[j].x < n_cams ? 1 : 0;
    }
    cnt[i] = c;  // cnt[n_obs] = 0: the exclusive scan's last entry is the total
}
__global__ void k_dn_pair_fill(int n_obs, int n_cams, const int2* __restrict__ pm_idx, const int* __restrict__ pt_off,
                               const int* __restrict__ pm_tc, const int* __restrict__ off, unsigned* __restrict__ key,
                               int* __restrict__ idx, int4* __restrict__ pairs) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    const int2 me = pm_idx[i];
    if (me.x >= n_cams) return;
    int k = off[i];
    const int ti = pm_tc[i];
    for (int j = pt_off[me.y]; j < pt_off[me.y + 1]; ++j) {
        const int cj = pm_idx[j].x;
        if (cj >= n_cams) continue;
        key[k] = (unsigned)(me.x * n_cams + cj);
        idx[k] = k;
        pairs[k] = make_int4(ti, pm_tc[j], me.y, 0);
        ++k;
    }
}
__global__ void k_dn_pair_gather(int n, const int* __restrict__ perm, const int4* __restrict__ in, int4* __restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) out[k] = in[perm[k]];
}

// ONE kernel per LM iteration does the whole camera side of a small problem: CTA (c, c') adds its bin
//      S_cc' = delta_cc' U_c  -  sum over the pairs of C_i^T (Jt_i Vinv Jt_j^T) C_j                (row-major 6x6)
// and the diagonal CTAs, from their i = j pairs (= the observations of camera c), also diag(U_c), the gradient
// g_c = sum C_i^T rho_i  (rho = W^-1/2 w_g r, stored beside the Jacobian cache by the tile linearisation) and E h = sum C_i^T Jt_i h_p.
// That replaces the two camera-major passes, their tile sums and finish kernels.  The CTA that finishes LAST (integer ticket;
// what it computes does not depend on which one it is) assembles S + lambda D, factorises it (Cholesky in shared memory),
// solves S x = E h - g_c, and writes x, g_c, D_c, the gradient test.  Fixed summation order everywhere: bit-reproducible.
constexpr int DN_MAX_SPLIT = 8;  // CTAs that share a bin
constexpr int DN_EXTRA = 18;  // per camera: diag(U) 6, g_c 6, E h 6

__host__ __device__ inline size_t dn_solve_smem(int n_cams) {
    const size_t n = 6 * (size_t)n_cams;
    return (n * (n + 1) + 4 * n) * sizeof(double);
}

template <bool WARP>
BA_D void dn_sync() {
    if (WARP) __syncwarp(); else __syncthreads();
}
// A (n x n, leading dimension ld, lower triangle) = L L^T in place (the diagonal ends up holding 1 / L_kk, so that the two
// triangular solves multiply instead of divide), then y <- A^-1 y, by a group of nt threads (t = index in the
// group): one warp (WARP) or the whole CTA.  Returns true - in every thread of the group - when a pivot is not positive.
template <bool WARP>
BA_D bool dn_cholesky_solve(double* A, double* y, int n, int ld, int t, int nt) {
    for (int k = 0; k < n; ++k) {  // right-looking Cholesky
        const double d = A[k * ld + k];
        if (!(d > 0.0) || !isfinite(d)) return true;  // same value in every thread: uniform exit
        const double ljj = sqrt(d), ilj = 1.0 / ljj;
        dn_sync<WARP>();
        for (int i = k + t; i < n; i += nt) A[i * ld + k] = (i == k) ? ilj : A[i * ld + k] * ilj;  // the diagonal holds 1 / L_kk
        dn_sync<WARP>();
        const int m = n - k - 1;  // trailing block: entries (i, j), k < j <= i < n
        for (int l = t; l < m * m; l += nt) {
            const int q = l / m;
            const int i = k + 1 + q, j = k + 1 + (l - q * m);
            if (j <= i) A[i * ld + j] -= A[i * ld + k] * A[j * ld + k];
        }
        dn_sync<WARP>();
    }
    for (int k = 0; k < n; ++k) {  // L y = b
        const double yk = y[k] * A[k * ld + k];  // every thread computes the same value
        dn_sync<WARP>();
        if (t == 0) y[k] = yk;
        for (int i = k + 1 + t; i < n; i += nt) y[i] -= A[i * ld + k] * yk;
        dn_sync<WARP>();
    }
    for (int k = n - 1; k >= 0; --k) {  // L^T x = y
        const double xk = y[k] * A[k * ld + k];
        dn_sync<WARP>();
        if (t == 0) y[k] = xk;
        for (int i = t; i < k; i += nt) y[i] -= A[k * ld + i] * xk;
        dn_sync<WARP>();
    }
    return false;
}

#ifdef BA_DN_STAMPS
// experiment builds only (scripts/window_stamps.py): %globaltimer at the stage boundaries of the last launch
__device__ unsigned long long g_dn_stamps[16];
#define DN_STAMP(i) do { if (threadIdx.x == 0) g_dn_stamps[i] = global_ns(); } while (0)
#else
#define DN_STAMP(i) do { } while (0)
#endif

// n <= 32 unknowns (the pipeline's windows: 4 adjustable keyframes = 24): ONE warp, lane i owning row i, square-root-free
// (A = L D L^T, L unit lower), left-looking.  What bounds a factorisation this small is the chain of dependent operations
// from one pivot to the next, so that chain stays in registers and shuffles - pivot by shuffle, reciprocal, L_ik = s_i / d_k,
// one FMA for the last term of the next column - while the rest of the next column's inner products (terms of the columns
// before k, which do not depend on this pivot) issue in the shadow of the reciprocal: loads only, uniform trip counts.
// The lower triangle receives L; the upper triangle (free) keeps the undivided entries t_km = L_km d_m that the inner
// products of row k's column need, so every operand is a conflict-free row read or a broadcast.
// Measured at n = 24 (scripts/window_stamps.py): 8.4 us for factorisation + both solves (0.28 us per column: two shuffles, the
// reciprocal and three dependent FP64 operations at about 40 cycles each; 33 ns per row of a solve), against 11.4-14.4 us for
// right-looking variants with the pivot in shared memory and a library rsqrt, and 11.9 us for rows held in registers (unrolled).
// y <- A^-1 y; returns true (in every lane) on a non-positive pivot.
BA_D bool dn_ldlt_solve_warp(double* A, int ld, double* y, int n, int lane) {
    constexpr unsigned FULL = 0xffffffffu;
    const bool mine = lane < n;
    double* row = A + (size_t)min(lane, n - 1) * ld;
    double s = mine ? row[0] : 0.0;  // entry (i, k) of the current column with every earlier column's term subtracted
    double inv = 0.0;                // 1 / d_i in lane i
    bool bad = false;                // same in every lane: the pivots are broadcast
    for (int k = 0; k < n; ++k) {
        const double d = __shfl_sync(FULL, s, k);
        const double tk1 = __shfl_sync(FULL, s, min(k + 1, 31));  // t_{k+1,k}
        // branch-free: the library division is a call with a slow path, which the scheduler does not overlap with the inner
        // products below.  Hardware seed (rcp.approx.ftz.f64: about 20 bits) and two Newton steps give 1 / d to an ulp or
        // two for the pivots that pass the test (checked after the loop: a bad pivot only poisons numbers nobody uses).
        bad |= !(d > 1e-280) || !(d < 1e280);
        double r;
        asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
        r = fma(r, fma(-d, r, 1.0), r);
        r = fma(r, fma(-d, r, 1.0), r);
        // entry (i, k + 1) minus the terms of the columns m < k (independent of r)
        const int k1 = min(k + 1, n - 1);
        const double* col = A + k1;
        double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
        int m = 0;
        for (; m + 3 < k; m += 4) {
            p0 += row[m] * col[m * ld];
            p1 += row[m + 1] * col[(m + 1) * ld];
            p2 += row[m + 2] * col[(m + 2) * ld];
            p3 += row[m + 3] * col[(m + 3) * ld];
        }
        for (; m < k; ++m) p0 += row[m] * col[m * ld];
        const double part = row[k1] - ((p0 + p1) + (p2 + p3));
        if (lane == k) inv = r;
        const bool below = mine && lane > k;
        const double l = below ? s * r : 0.0;  // L_ik
        if (below) {
            row[k] = l;
            A[k * ld + lane] = s;              // t_ik, read as the multiplier of row i's columns later
        }
        s = part - l * tk1;
        __syncwarp();
    }
    if (bad) return true;
    DN_STAMP(9);
    double rhs = mine ? y[lane] : 0.0;
    for (int k = 0; k < n; ++k) {  // L z = b (unit diagonal)
        const double zk = __shfl_sync(FULL, rhs, k);
        if (mine && lane > k) rhs -= row[k] * zk;
    }
    rhs *= inv;                    // D w = z
    DN_STAMP(10);
    for (int k = n - 1; k >= 0; --k) {  // L^T x = w
        const double xk = __shfl_sync(FULL, rhs, k);
        if (lane < k) rhs -= A[k * ld + lane] * xk;
    }
    DN_STAMP(11);
    if (mine) y[lane] = rhs;
    return false;
}

__global__ void __launch_bounds__(DN_THREADS) k_dense_schur(Params P, const int4* __restrict__ pairs, const int* __restrict__ bin_off,
                                                            double* __restrict__ S_bins, double* __restrict__ extra,
                                                            int* __restrict__ ticket, int* __restrict__ n_solves, double dmin,
                                                            double dmax, double gtol, int split) {
    extern __shared__ __align__(16) double dsm[];
    __shared__ double red[DN_THREADS / 32 * (36 + DN_EXTRA)];
    __shared__ double tot[36 + DN_EXTRA];
    __shared__ int last, bad;
    const int tid = threadIdx.x;
    // `split` CTAs share a bin (contiguous parts of its pair list): a window has 16 bins of a few thousand pairs, and one CTA
    // per bin would walk them in a dozen dependent-load round trips; the last CTA adds the parts in order
    const int nc = P.n_cams, bin = blockIdx.x / split, part = blockIdx.x - bin * split;
    const int ci = bin / nc, cj = bin - ci * nc;
    const bool diag = ci == cj;
    const int b = P.st->cur;
    if (blockIdx.x == 0) DN_STAMP(0);
    const double2* Jb = P.tc_J + (size_t)b * 3 * P.n_obs;
    const double2* Rb = P.tc_r + (size_t)b * P.n_obs;
    double acc[36 + DN_EXTRA];
#pragma unroll
    for (int k = 0; k < 36 + DN_EXTRA; ++k) acc[k] = 0.0;
    const int b0_ = __ldg(bin_off + bin), b1_ = __ldg(bin_off + bin + 1);
    const int k0 = b0_ + (int)((long long)(b1_ - b0_) * part / split), k1 = b0_ + (int)((long long)(b1_ - b0_) * (part + 1) / split);
    for (int k = k0 + tid; k < k1; k += DN_THREADS) {
        const int4 pr = __ldg(pairs + k);
        double X[3], Vi[6];
        load_pt(P, b, pr.z, X);
        {
            const double2* v = reinterpret_cast<const double2*>(P.pt_Vinv + (size_t)pr.z * 6);
            const double2 a0 = v[0], a1 = v[1], a2 = v[2];
            Vi[0] = a0.x; Vi[1] = a0.y; Vi[2] = a1.x; Vi[3] = a1.y; Vi[4] = a2.x; Vi[5] = a2.y;  // 00 01 02 11 12 22
        }
        const double2 i0 = Jb[pr.x], i1 = Jb[(size_t)P.n_obs + pr.x], i2 = Jb[2 * (size_t)P.n_obs + pr.x];
        const double2 j0 = Jb[pr.y], j1 = Jb[(size_t)P.n_obs + pr.y], j2 = Jb[2 * (size_t)P.n_obs + pr.y];
        const double jxi[3] = {i0.x, i0.y, i1.x}, jyi[3] = {i1.y, i2.x, i2.y};
        const double jxj[3] = {j0.x, j0.y, j1.x}, jyj[3] = {j1.y, j2.x, j2.y};
        double tx[3], ty[3];  // Vinv jx_j, Vinv jy_j
        sym3_mulv(Vi, jxj, tx);
        sym3_mulv(Vi, jyj, ty);
        const double K00 = jxi[0] * tx[0] + jxi[1] * tx[1] + jxi[2] * tx[2], K01 = jxi[0] * ty[0] + jxi[1] * ty[1] + jxi[2] * ty[2];
        const double K10 = jyi[0] * tx[0] + jyi[1] * tx[1] + jyi[2] * tx[2], K11 = jyi[0] * ty[0] + jyi[1] * ty[1] + jyi[2] * ty[2];
        double cxi[6], cyi[6], cxj[6], cyj[6];  // rows of C = (X x j, j)
        cross3(X, jxi, cxi); cross3(X, jyi, cyi); cross3(X, jxj, cxj); cross3(X, jyj, cyj);
#pragma unroll
        for (int q = 0; q < 3; ++q) { cxi[3 + q] = jxi[q]; cyi[3 + q] = jyi[q]; cxj[3 + q] = jxj[q]; cyj[3 + q] = jyj[q]; }
        const bool self = diag && pr.x == pr.y;  // the observation itself: its U, g_c and E h terms
        const double u = self ? 1.0 : 0.0;
        double Tx[6], Ty[6];  // (delta_ij I - K) C_j
#pragma unroll
        for (int q = 0; q < 6; ++q) { Tx[q] = (u - K00) * cxj[q] - K01 * cyj[q]; Ty[q] = -K10 * cxj[q] + (u - K11) * cyj[q]; }
#pragma unroll
        for (int a = 0; a < 6; ++a)
#pragma unroll
            for (int q = 0; q < 6; ++q) acc[6 * a + q] = fma(cxi[a], Tx[q], fma(cyi[a], Ty[q], acc[6 * a + q]));
        if (self) {
            const double2 rho = Rb[pr.x];
            double h[3];
            load_v3(P.pt_h + (size_t)pr.z * 4, h);
            const double ex = jxi[0] * h[0] + jxi[1] * h[1] + jxi[2] * h[2], ey = jyi[0] * h[0] + jyi[1] * h[1] + jyi[2] * h[2];
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                acc[36 + a] = fma(cxi[a], cxi[a], fma(cyi[a], cyi[a], acc[36 + a]));
                acc[42 + a] = fma(cxi[a], rho.x, fma(cyi[a], rho.y, acc[42 + a]));
                acc[48 + a] = fma(cxi[a], ex, fma(cyi[a], ey, acc[48 + a]));
            }
        }
    }
    if (blockIdx.x == 0) DN_STAMP(1);
    block_sum<36 + DN_EXTRA, DN_THREADS>(acc, red, tot);
    __syncthreads();
    if (blockIdx.x == 0) DN_STAMP(2);
    if (tid < 36) S_bins[((size_t)bin * split + part) * 36 + tid] = tot[tid];
    if (diag && tid < DN_EXTRA) extra[((size_t)ci * split + part) * DN_EXTRA + tid] = tot[36 + tid];
    __threadfence();
    __syncthreads();
    if (tid == 0) last = (atomicAdd(ticket, 1) == (int)gridDim.x - 1);
    __syncthreads();
    if (!last) return;
    __threadfence();
    DN_STAMP(3);

    // ---- the last CTA: S + lambda D, Cholesky, solve -------------------------------------------------------------------
    const int n = 6 * nc, ld = n + 1;
    double* A = dsm;                       // [n][ld], lower triangle
    double* y = dsm + (size_t)n * ld;      // [n] right-hand side -> solution
    double* D = y + n;                     // [n]
    double* G = D + n;                     // [n] gradient
    LMState* st = P.st;
    const double lam = st->lambda;
    if (tid == 0) { bad = 0; *ticket = 0; }
    for (int l = tid; l < n; l += DN_THREADS) {
        const int c = l / 6, a = l - 6 * c;
        double du = 0.0, g = 0.0, eh = 0.0;
        for (int sp = 0; sp < split; ++sp) {  // the parts in order
            const double* e = extra + ((size_t)c * split + sp) * DN_EXTRA;
            du += __ldcg(e + a); g += __ldcg(e + 6 + a); eh += __ldcg(e + 12 + a);
        }
        const double d = fmin(fmax(du, dmin), dmax);
        D[l] = d;
        G[l] = g;
        y[l] = eh - g;
        P.cam_D[l] = d; P.cam_g[l] = g; P.v_b[l] = eh - g;
    }
    __syncthreads();
    DN_STAMP(4);
    for (int l = tid; l < n * n; l += DN_THREADS) {
        const int r = l / n, c = l - r * n;
        if (c > r) continue;
        const int cr = r / 6, a = r - 6 * cr, cc = c / 6, q = c - 6 * cc;
        const double* src = S_bins + (size_t)(cr * nc + cc) * split * 36 + 6 * a + q;
        double part[DN_MAX_SPLIT];  // all loads in flight, then the adds in order
#pragma unroll
        for (int sp = 0; sp < DN_MAX_SPLIT; ++sp) part[sp] = sp < split ? __ldcg(src + sp * 36) : 0.0;
        double v = 0.0;
#pragma unroll
        for (int sp = 0; sp < DN_MAX_SPLIT; ++sp) v += part[sp];
        if (r == c) v += lam * D[r];
        A[r * ld + c] = v;
    }
    // gradient test at the current point (skipped after a rejected step: the point and its gradient are unchanged)
    if (tid == 0 && !lin_gated(st)) {
        double m = st->scal[S_GPMAX];
        for (int l = 0; l < n; ++l) m = fmax(m, fabs(G[l]));
        st->gmax = m;
        if (st->status == 0 && m < gtol) st->status = 1;
    }
    __syncthreads();
    DN_STAMP(5);
    if (n <= 32) {
        // up to five adjustable cameras: one warp, a row per lane
        if (tid < 32) {
            const bool b0 = dn_ldlt_solve_warp(A, ld, y, n, tid);
            if (tid == 0) bad = b0 ? 1 : 0;
        }
    } else {
        // (one warp with the CTA routine's element-strided loops was tried for n <= 48: slower than the CTA)
        const bool b0 = dn_cholesky_solve<false>(A, y, n, ld, tid, DN_THREADS);
        if (tid == 0) bad = b0 ? 1 : 0;
    }
    __syncthreads();
    DN_STAMP(6);
    // a pivot that is not positive (cannot happen for a damped Schur complement in exact arithmetic) leaves x = 0: the points
    // alone move, the model decrease stays exact, LM carries on
    for (int l = tid; l < n; l += DN_THREADS) {
        const double x = bad ? 0.0 : y[l];
        P.v_x[l] = isfinite(x) ? x : 0.0;
        P.v_r[l] = 0.0;  // PCG residual of the step: exact
    }
    if (tid == 0) {
        st->pcg_iter = 0;
        st->pcg_done = 1;
        if (n_solves) *n_solves += bad ? 0 : 1;
    }
    // the trial poses and the step's camera-side scalars follow in the back-substitution launch (its extra CTA, ba_tiles.cuh)
    DN_STAMP(7);
    DN_STAMP(8);
}

}  // namespace ba
