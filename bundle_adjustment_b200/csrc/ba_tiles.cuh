// ba_tiles.cuh - the fused point-tile passes of the Schur solver.
//
// A TILE is a contiguous range of (locality-ordered) points together with ALL their observations,
// at most FT_THREADS of them, so that one CTA holds one thread per observation.  Because every
// observation of a point lives in the tile, both halves of the implicit Schur mat-vec
//
//      z_p = V_p^-1 sum_i Jp_i^T W_i Jc_i u_c(i)            (per point,  gather over cameras)
//      w_c = sum_i Jc_i^T W_i (Jc_i u_c - Jp_i z_p(i))      (per camera, gather over points)
//
// run inside one kernel with z_p in shared memory: the observation record is streamed ONCE per PCG
// iteration, the 2x3 Jacobian block is derived once and kept in registers between the two halves,
// and nothing per-observation or per-point is written to HBM.
//
// Inside a tile the observations are stored sorted by camera ("tile-camera order", built once per
// problem on the device).  Consequences:
//   * neighbouring lanes read the SAME camera record -> the camera gather is a broadcast load (a few
//     L1 wavefronts per warp instead of one per lane);
//   * a camera's observations form a contiguous run of threads; runs are cut into SEGMENTS of at most
//     FT_SEG_MAX observations, each segment is summed in index order by one thread per component and
//     leaves one 6-double (mat-vec) or 27-double (linearisation) ROW in HBM;
//   * rows are laid out camera-major, so the per-camera finish is a contiguous, fixed-order sum.
// No floating-point atomics: every sum has a fixed order, results are bit-reproducible.
//
// Per-observation word (tc_word): bits 0-9 local point index, 10-19 position of the observation in
// the tile's point-major order (where its per-point contribution is parked), 20-31 local segment.
#pragma once
#include "ba_kernels.cuh"

namespace ba {

constexpr int FT_THREADS = 512;        // threads per CTA = observation capacity of a tile
constexpr int FT_SEG_MAX = 32;         // observations per segment (one output row each)
constexpr int FT_LD = FT_THREADS + 2;  // row stride of the per-observation scratch (doubles); 2*FT_LD mod 32 = 4 banks
constexpr int FT_ROWS_MV = 6;          // scratch rows of the mat-vec pass
constexpr size_t ft_smem_bytes(int rows) { return (size_t)(FT_THREADS * 3 + rows * FT_LD) * sizeof(double); }

struct TileHdr { int p0, m, o0, n, s0, ns; };
BA_D TileHdr tile_hdr(const Params& P, int t) {
    TileHdr T;
    T.p0 = __ldg(P.ft_pt + t);
    T.m = __ldg(P.ft_pt + t + 1) - T.p0;
    T.o0 = __ldg(P.ft_obs + t);
    T.n = __ldg(P.ft_obs + t + 1) - T.o0;
    T.s0 = __ldg(P.ft_seg + t);
    T.ns = __ldg(P.ft_seg + t + 1) - T.s0;
    return T;
}

// MODE 0: one application of the Schur complement's point part to `vec`:
//         rows[seg] = sum over the segment of Jc^T W (Jc vec_c - Jp z_p),  z_p = Vinv_p sum Jp^T W Jc vec_c.
// MODE 1: back-substitution with the PCG solution `vec`: dp = -h_p - z_p(vec), X_trial = X + dp, and the
//         step's scalar products (per tile partials).
template <int MODE>
__global__ void __launch_bounds__(FT_THREADS, 2) k_tile_pass(Params P, const double* __restrict__ vec) {
    extern __shared__ __align__(16) double sm[];
    double* sP = sm;                   // [m][3]  point positions (phase A), then z_p (phase C)
    double* sO = sm + FT_THREADS * 3;  // scratch rows of FT_LD doubles
    __shared__ double red[FT_THREADS / 32 * 4];
    __shared__ double tot[4];
    if (MODE == 0 && P.st->pcg_done) return;
    const int b = P.st->cur;
    const int tid = threadIdx.x;
    const TileHdr T = tile_hdr(P, blockIdx.x);
    const bool act = tid < T.n;

    // ---- phase 0: stage the tile's points, fetch this thread's observation and camera ----------
    double2 uv = make_double2(0.0, 0.0);
    unsigned word = 0;
    double rec[18];
    if (act) {
        uv = __ldg(P.tc_uv + T.o0 + tid);
        word = __ldg(P.tc_word + T.o0 + tid);
    }
    if (tid < T.m) {
        const double2* q = reinterpret_cast<const double2*>(pt_x(P, b, T.p0 + tid));
        const double2 a = __ldg(q), c = __ldg(q + 1);
        sP[3 * tid] = a.x; sP[3 * tid + 1] = a.y; sP[3 * tid + 2] = c.x;
    }
    if (act) {
        const int cam = __ldg(P.seg_cam + T.s0 + (word >> 20));
        load_cam(P.cam_R + (size_t)b * (P.n_cams + 1) * 12, cam, rec);
        load_vec6(vec, cam, rec + 12);
    }
    __syncthreads();

    // ---- phase A: linearise; park Jp^T W Jc vec_c at the observation's point-major position ------
    double X[3] = {0.0, 0.0, 0.0}, jx[3] = {0.0, 0.0, 0.0}, jy[3] = {0.0, 0.0, 0.0};
    double ax = 0.0, ay = 0.0, wHx = 0.0, wHy = 0.0;
    const int pl = word & 1023u;
    if (act) {
        const int pos = (word >> 10) & 1023u;
        X[0] = sP[3 * pl]; X[1] = sP[3 * pl + 1]; X[2] = sP[3 * pl + 2];
        ObsLin o;
        linearize_obs(rec, X, uv.x, uv.y, P.K, o);
        double wgx, r0x, wgy, r0y;
        robust_weights(o.rx, P.delta, P.mode, wHx, wgx, r0x);
        robust_weights(o.ry, P.delta, P.mode, wHy, wgy, r0y);
        const double* x = rec + 12;
        const double d0 = fma(x[1], X[2], fma(-x[2], X[1], x[3]));  // d = w x X + v
        const double d1 = fma(x[2], X[0], fma(-x[0], X[2], x[4]));
        const double d2 = fma(x[0], X[1], fma(-x[1], X[0], x[5]));
#pragma unroll
        for (int k = 0; k < 3; ++k) { jx[k] = o.jx[k]; jy[k] = o.jy[k]; }
        ax = wHx * (jx[0] * d0 + jx[1] * d1 + jx[2] * d2);
        ay = wHy * (jy[0] * d0 + jy[1] * d1 + jy[2] * d2);
        sO[pos] = jx[0] * ax + jy[0] * ay;
        sO[FT_LD + pos] = jx[1] * ax + jy[1] * ay;
        sO[2 * FT_LD + pos] = jx[2] * ax + jy[2] * ay;
    }
    __syncthreads();

    // ---- phase B: one thread per point: sum its track in observation order, apply Vinv -----------
    double acc[4] = {0.0, 0.0, 0.0, 0.0};  // MODE 1: g.dp, dp D dp, |dp|^2, |X|^2
    if (tid < T.m) {
        const int p = T.p0 + tid;
        const int a = __ldg(P.pt_off + p) - T.o0, e = __ldg(P.pt_off + p + 1) - T.o0;
        double s[3] = {0.0, 0.0, 0.0};
        for (int i = a; i < e; ++i) {
            s[0] += sO[i]; s[1] += sO[FT_LD + i]; s[2] += sO[2 * FT_LD + i];
        }
        double Vi[6], z[3];
        load_sym3(P.pt_Vinv, p, Vi);
        sym3_mulv(Vi, s, z);
        if (MODE == 0) {
            sP[3 * tid] = z[0]; sP[3 * tid + 1] = z[1]; sP[3 * tid + 2] = z[2];
        } else {
            double h[3], g[3], D[3];
            const double Xp[3] = {sP[3 * tid], sP[3 * tid + 1], sP[3 * tid + 2]};
            load_v3(P.pt_h + (size_t)p * 4, h);
            load_v3(P.pt_g + (size_t)b * P.n_pts * 4 + (size_t)p * 4, g);
            load_v3(P.pt_D + (size_t)p * 4, D);
            const double dp0 = -h[0] - z[0], dp1 = -h[1] - z[1], dp2 = -h[2] - z[2];
            store_v3(pt_x(P, b ^ 1, p), Xp[0] + dp0, Xp[1] + dp1, Xp[2] + dp2);
            acc[0] = g[0] * dp0 + g[1] * dp1 + g[2] * dp2;
            acc[1] = D[0] * dp0 * dp0 + D[1] * dp1 * dp1 + D[2] * dp2 * dp2;
            acc[2] = dp0 * dp0 + dp1 * dp1 + dp2 * dp2;
            acc[3] = Xp[0] * Xp[0] + Xp[1] * Xp[1] + Xp[2] * Xp[2];
        }
    }
    if (MODE == 1) {
        block_sum<4, FT_THREADS>(acc, red, tot);
        __syncthreads();
        if (tid == 0) {
            double* out = P.pm_partial + (size_t)blockIdx.x * NSCAL;
            out[S_GDP] = tot[0]; out[S_DDP] = tot[1]; out[S_DP2] = tot[2]; out[S_XP2] = tot[3];
        }
        return;
    }
    __syncthreads();

    // ---- phase C: camera-side product of every observation, parked at its tile-camera position ----
    if (act) {
        const double z0 = sP[3 * pl], z1 = sP[3 * pl + 1], z2 = sP[3 * pl + 2];
        const double bx = ax - wHx * (jx[0] * z0 + jx[1] * z1 + jx[2] * z2);
        const double by = ay - wHy * (jy[0] * z0 + jy[1] * z1 + jy[2] * z2);
        double mm[3], xm[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) mm[k] = jx[k] * bx + jy[k] * by;
        cross3(X, mm, xm);
#pragma unroll
        for (int k = 0; k < 3; ++k) { sO[k * FT_LD + tid] = xm[k]; sO[(3 + k) * FT_LD + tid] = mm[k]; }
    }
    __syncthreads();

    // ---- phase D: one thread per (segment, component): fixed-order sum -> one row per segment -----
    for (int w = tid; w < T.ns * 6; w += FT_THREADS) {
        const int seg = w / 6, k = w - 6 * seg;
        const int a = __ldg(P.seg_start + T.s0 + seg) - T.o0, e = __ldg(P.seg_start + T.s0 + seg + 1) - T.o0;
        const double* src = sO + k * FT_LD;
        double s = 0.0;
        for (int i = a; i < e; ++i) s += src[i];
        P.rows[(size_t)__ldg(P.seg_row + T.s0 + seg) * 6 + k] = s;
    }
}

// ---------------------------------------------------------------------------------------------
// per-camera finish of the mat-vec rows: one warp per camera adds the camera's rows (contiguous,
// fixed lane/shuffle order).  TO_ACC: write the sums to cam_acc (all-reduced by the caller, multi-GPU);
// otherwise finish the PCG inner product here: w = sum + lambda D u, per-CTA partial of w.u.
// ---------------------------------------------------------------------------------------------
constexpr int RS_WARPS = 8;  // cameras per CTA

template <bool TO_ACC>
__global__ void __launch_bounds__(RS_WARPS * 32) k_rows_finish(Params P) {
    __shared__ double sm[RS_WARPS];
    if (P.st->pcg_done) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int c = blockIdx.x * RS_WARPS + wid;
    double part = 0.0;
    if (c < P.n_cams) {
        const int r0 = __ldg(P.cam_row_off + c), r1 = __ldg(P.cam_row_off + c + 1);
        double a[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        for (int r = r0 + lane; r < r1; r += 32) {
            const double2* q = reinterpret_cast<const double2*>(P.rows + (size_t)r * 6);
            const double2 v0 = q[0], v1 = q[1], v2 = q[2];
            a[0] += v0.x; a[1] += v0.y; a[2] += v1.x; a[3] += v1.y; a[4] += v2.x; a[5] += v2.y;
        }
#pragma unroll
        for (int k = 0; k < 6; ++k) a[k] = warp_sum(a[k]);
        if (lane == 0) {
            if (TO_ACC) {
#pragma unroll
                for (int k = 0; k < 6; ++k) P.cam_acc[(size_t)c * 6 + k] = a[k];
            } else {
                const double lam = P.st->lambda;
#pragma unroll
                for (int k = 0; k < 6; ++k) {
                    const size_t i = (size_t)c * 6 + k;
                    const double u = P.v_u[i];
                    const double wk = a[k] + lam * P.cam_D[i] * u;
                    P.v_w[i] = wk;
                    part += wk * u;
                }
            }
        }
    }
    if (TO_ACC) return;
    if (lane == 0) sm[wid] = part;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
#pragma unroll
        for (int i = 0; i < RS_WARPS; ++i) s += sm[i];
        P.dot_partial[2 * blockIdx.x] = s;
    }
}

// ---------------------------------------------------------------------------------------------
// tile construction (once per problem)
// ---------------------------------------------------------------------------------------------
// key = (tile << 32) | camera slot for the observation at point-major index j
__global__ void k_ft_keys(int n_obs, int n_tiles, const int* __restrict__ ft_obs, const int2* __restrict__ pm_idx,
                          unsigned long long* __restrict__ key, int* __restrict__ iota) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_obs) return;
    int lo = 0, hi = n_tiles;  // last tile with ft_obs[t] <= j
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (ft_obs[mid] <= j) lo = mid; else hi = mid;
    }
    key[j] = ((unsigned long long)lo << 32) | (unsigned)pm_idx[j].x;
    iota[j] = j;
}
// start of the (tile, camera) run that tile-camera position i belongs to (as a max-scan input)
__global__ void k_ft_run_heads(int n_obs, const unsigned long long* __restrict__ key_s, int* __restrict__ head) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    head[i] = (i == 0 || key_s[i] != key_s[i - 1]) ? i : 0;
}
__global__ void k_ft_seg_flags(int n_obs, const int* __restrict__ run_start, int* __restrict__ flag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    flag[i] = ((i - run_start[i]) % FT_SEG_MAX) == 0;
}
// seg_inc = inclusive sum of the segment-head flags, i.e. segment id + 1
__global__ void k_ft_tile_seg(int n_tiles, int n_obs, int n_seg, const int* __restrict__ ft_obs, const int* __restrict__ seg_inc,
                              int* __restrict__ ft_seg, int* __restrict__ seg_start) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > n_tiles) return;
    if (t == n_tiles) { ft_seg[t] = n_seg; seg_start[n_seg] = n_obs; return; }
    const int o = ft_obs[t];
    ft_seg[t] = o < n_obs ? seg_inc[o] - 1 : n_seg;  // trailing tiles of unobserved points own no segment
}
__global__ void k_ft_fill(int n_obs, const unsigned long long* __restrict__ key_s, const int* __restrict__ src,
                          const int* __restrict__ flag, const int* __restrict__ seg_inc, const int* __restrict__ ft_obs,
                          const int* __restrict__ ft_pt, const int* __restrict__ ft_seg, const double2* __restrict__ pm_uv,
                          const int2* __restrict__ pm_idx, double2* __restrict__ tc_uv, unsigned* __restrict__ tc_word,
                          int* __restrict__ seg_cam, int* __restrict__ seg_start, unsigned* __restrict__ seg_key) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    const int j = src[i];
    const int t = (int)(key_s[i] >> 32), cam = (int)(unsigned)key_s[i];
    const int s = seg_inc[i] - 1;
    const unsigned pl = (unsigned)(pm_idx[j].y - ft_pt[t]), pos = (unsigned)(j - ft_obs[t]), sl = (unsigned)(s - ft_seg[t]);
    tc_uv[i] = pm_uv[j];
    tc_word[i] = pl | (pos << 10) | (sl << 20);
    if (flag[i]) { seg_cam[s] = cam; seg_start[s] = i; seg_key[s] = (unsigned)cam; }
}
__global__ void k_ft_seg_rows(int n_seg, const int* __restrict__ sorted_seg, int* __restrict__ seg_row) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n_seg) seg_row[sorted_seg[r]] = r;
}
__global__ void k_iota(int n, int* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = i;
}

}  // namespace ba
