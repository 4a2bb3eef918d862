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
//     FT_SEG_MAX observations, each segment is summed in index order by one thread per PAIR of components
//     (the per-observation products are parked as three double2) and leaves one 6-double ROW in HBM - or adds
//     to a shared-memory accumulator slot that becomes a row when its epoch ends;
//   * rows are laid out camera-major, so the per-camera finish is a contiguous, fixed-order sum.
// No floating-point atomics: every sum has a fixed order, results are bit-reproducible.
//
// Per-observation record (tc_meta): .x camera slot, .y = local point index | position of the observation
// in the tile's point-major order << 10 (where its per-point contribution is parked).
#pragma once
#include "ba_kernels.cuh"

#include <cub/block/block_scan.cuh>

namespace ba {

constexpr int FT_THREADS = 256;        // threads per CTA = observation capacity of a tile
constexpr int FT_PTS = 128;            // point capacity of a tile
constexpr int FT_CTAS_PER_SM = 1024 / FT_THREADS;  // register file: 64 registers per thread
constexpr int FT_SEG_MAX = 16;         // observations per segment (one output row each)
constexpr int FT_LD = FT_THREADS + 3;  // row stride of the camera-side scratch (phases C/D), in double2: three rows of value PAIRS
                                       // (xm0,xm1) (xm2,mm0) (mm1,mm2); 3 mod 8 sixteen-byte bank groups between the pairs of a segment
constexpr int FT_LDB = FT_THREADS + 4; // row stride of the point-side scratch (phases A/B): 2*FT_LDB mod 32 = 8 banks, so the
                                       // three rows read by the four points of a half-warp are disjoint
constexpr int FT_ROWS_MV = 6;          // doubles per observation in the camera-side scratch of the mat-vec pass (three double2 rows)
constexpr int FT_ACC_CAP = 96;         // accumulator slots (6 doubles + row number each) of an epoch; what is left of 228 KB / 4 CTAs

// Tiles of persistent CTA b of a grid of G over T tiles, and the way a CTA cuts its run into EPOCHS of about
// `target` tiles (equal lengths, the last one may be shorter).  Host and device use the same two functions.
__host__ __device__ inline int ft_cta_begin(int b, int T, int G) { return (int)((long long)b * T / G); }
__host__ __device__ inline int ft_cta_of_tile(int t, int T, int G) { return (int)((((long long)t + 1) * G - 1) / T); }
__host__ __device__ inline int ft_epoch_len(int len, int target) {
    const int ne = (len + target - 1) / target;
    return ne > 0 ? (len + ne - 1) / ne : 1;
}

// dynamic shared memory of a tile pass (16-byte aligned regions): per-observation scratch rows | Jacobian
// cache planes [3][FT_THREADS] double2 | observation records | X as planes (X0,X1)[FT_PTS], X2[FT_PTS+2] | z likewise | Vinv [FT_PTS][6] |
// Vinv x2 ... | track offsets x2 | track-row starts x2 | segment table x2 | segment rows x2 | two mbarriers.  The x2 regions are double-buffered
// by tile parity; everything else is refilled by the TMA as soon as its last reader has passed a barrier.
constexpr int FT_META_N = FT_THREADS + 2, FT_OFF_N = FT_PTS + 8, FT_SEG_N = FT_THREADS + 8;
// With epoch accumulators the segment table and the row table (2 x 2 x FT_SEG_N ints) hold the tile's packed WORK ITEMS of
// phase D instead (k_ft_item_fill): two buffers of FT_ITEM_N ints, i.e. at most FT_ITEM_QMAX quarter-warps of eight items.
constexpr int FT_ITEM_N = 2 * FT_SEG_N, FT_ITEM_QMAX = FT_ITEM_N / 8;
struct FtSmem {
    double* sA;  // [3][FT_LDB] point-side scratch (phases A/B); separate from sO so that phase D of a tile and phase A of the next overlap
    double* sO;  // [3][FT_LD] double2: (X x m, m) of every observation as three value pairs (phases C/D)
    double2* sJ; int2* sMeta; double2* sX01; double* sX2; double2* sZ01; double* sZ2; double* sVi;
    int* sOff; int* sRs; int* sSeg; int* sRow;
    unsigned long long* bar;
    double* sAcc;  // [FT_ACC_CAP][6] epoch accumulators (only when the pass is instantiated with them)
    int* sSlot;    // [FT_ACC_CAP] row of every slot of the current epoch
};
__host__ __device__ constexpr size_t ft_smem_bytes(int rows, bool acc = false) {
    return (size_t)(rows * FT_LD + 3 * FT_LDB) * 8 + 3 * FT_THREADS * 16 + FT_META_N * 8 + (size_t)FT_PTS * (2 + 2 + 6) * 8 +
           2 * (size_t)(FT_PTS + 2) * 8 + (size_t)(FT_OFF_N + 5 * FT_SEG_N) * 4 + 16 + (acc ? FT_ACC_CAP * 52 : 0);
}
BA_D FtSmem ft_carve(double* sm, int rows) {
    FtSmem S;
    S.sO = sm;
    S.sA = S.sO + rows * FT_LD;
    S.sJ = reinterpret_cast<double2*>(S.sA + 3 * FT_LDB);
    S.sMeta = reinterpret_cast<int2*>(S.sJ + 3 * FT_THREADS);
    S.sX01 = reinterpret_cast<double2*>(S.sMeta + FT_META_N);
    S.sX2 = reinterpret_cast<double*>(S.sX01 + FT_PTS);
    S.sZ01 = reinterpret_cast<double2*>(S.sX2 + FT_PTS + 2);
    S.sZ2 = reinterpret_cast<double*>(S.sZ01 + FT_PTS);
    S.sVi = S.sZ2 + FT_PTS + 2;
    S.sOff = reinterpret_cast<int*>(S.sVi + FT_PTS * 6);
    S.sRs = S.sOff + FT_OFF_N;
    S.sSeg = S.sRs + FT_SEG_N;
    S.sRow = S.sSeg + 2 * FT_SEG_N;
    S.bar = reinterpret_cast<unsigned long long*>(S.sRow + 2 * FT_SEG_N);
    S.sAcc = reinterpret_cast<double*>(S.bar + 2);
    S.sSlot = reinterpret_cast<int*>(S.sAcc + FT_ACC_CAP * 6);
    return S;
}

// ---- TMA (bulk asynchronous copy, global -> shared) and its mbarrier --------------------------------
BA_D unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
BA_D void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
BA_D void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("{\n.reg .b64 st;\nmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}\n" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
BA_D void mbar_wait(unsigned long long* bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(ok)
                     : "r"(smem_u32(bar)), "r"(parity)
                     : "memory");
    } while (!ok);
}
// bytes must be a positive multiple of 16, both addresses 16-byte aligned
BA_D void tma_load(void* smem, const void* gmem, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(smem_u32(smem)),
                 "l"(gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

struct TileHdr { int p0, m, o0, n, s0, ns, r0, maxlen; };
BA_D TileHdr tile_hdr(const Params& P, int t) {
    const int4 a = __ldg(P.ft_hdr + 2 * t), b = __ldg(P.ft_hdr + 2 * t + 1);
    TileHdr T;
    T.p0 = a.x; T.m = a.y; T.o0 = a.z; T.n = a.w; T.s0 = b.x; T.ns = b.y; T.r0 = b.z; T.maxlen = b.w;
    return T;
}

// MODE 0: one application of the Schur complement's point part to `vec`:
//         rows[seg] = sum over the segment of Jc^T W (Jc vec_c - Jp z_p),  z_p = Vinv_p sum Jp^T W Jc vec_c.
// MODE 1: back-substitution with the PCG solution `vec`: dp = -h_p - z_p(vec), X_trial = X + dp, and the
//         step's scalar products (per tile partials).
// The pass streams the JACOBIAN CACHE of the current linearisation: Jt = W^1/2 Jp (2x3, 48 bytes per
// observation, tile-camera order, written by k_lin_points).  With the weights folded in,
//         a = Jt (w x X + v),  s_p = sum Jt^T a,  z = Vinv s,  b = a - Jt z,  m = Jt^T b,  row += [X x m ; m],
// so the pass needs no camera pose, no measurement, no division: per observation 56 bytes from HBM, the
// camera's 6-vector from L1 (broadcast inside a camera run) and the point from shared memory.
//
// Persistent CTAs (grid = 2 per SM) each walk a contiguous run of tiles.  All HBM traffic of a tile is moved by
// the TMA (cp.async.bulk) while the PREVIOUS tile is being computed: group 1 (observation records,
// Jacobian cache, points, segment table) is re-armed right after phase A has pulled its operands into
// registers, together with group 2 (Vinv, track offsets, track rows; double-buffered) - by the first lane of
// the last warp, which has no work in phase B; each group completes on its own mbarrier.
struct TileCopy {  // 16-byte aligned source windows of one tile
    int meta_lo, off_lo, seg_lo;  // first element copied (<= the tile's first) of tc_meta / pt_off / seg arrays
};

template <int MODE, bool ACC = false>
BA_D void ft_issue_group1(const Params& P, const FtSmem& S, const TileHdr& T, int b, int buf, unsigned long long* bar) {
    const int meta_lo = T.o0 & ~1, seg_lo = T.s0 & ~3;
    const unsigned meta_b = (unsigned)(((T.o0 + T.n - meta_lo) * 8 + 15) & ~15);
    const unsigned j_b = (unsigned)T.n * 16u, x01_b = (unsigned)T.m * 16u;
    const int x2_lo = T.p0 & ~1;
    const unsigned x2_b = (unsigned)(((T.p0 + T.m - x2_lo) * 8 + 15) & ~15);
    // ACC: (s0, ns) of the header = (first work item, work item slots) of the tile, both multiples of 8
    const unsigned seg_b = (unsigned)(((T.s0 + T.ns - seg_lo) * 4 + 15) & ~15);
    unsigned total = x01_b + x2_b;
    if (T.n > 0) total += meta_b + 3 * j_b;
    if (MODE == 0 && seg_b) total += ACC ? seg_b : 2 * seg_b;
    mbar_expect_tx(bar, total);
    if (T.n > 0) {
        tma_load(S.sMeta, P.tc_meta + meta_lo, meta_b, bar);
        const double2* J = P.tc_J + (size_t)b * 3 * P.n_obs + T.o0;
#pragma unroll
        for (int r = 0; r < 3; ++r) tma_load(S.sJ + r * FT_THREADS, J + (size_t)r * P.n_obs, j_b, bar);
    }
    tma_load(S.sX01, reinterpret_cast<const double2*>(pt_base(P, b)) + T.p0, x01_b, bar);
    tma_load(S.sX2, pt_base(P, b) + 2 * (size_t)P.n_pad + x2_lo, x2_b, bar);
    if (MODE == 0 && seg_b) {
        if (ACC) {
            tma_load(S.sSeg + buf * FT_ITEM_N, P.seg_info + seg_lo, seg_b, bar);
        } else {
            tma_load(S.sSeg + buf * FT_SEG_N, P.seg_info + seg_lo, seg_b, bar);
            tma_load(S.sRow + buf * FT_SEG_N, P.seg_row + seg_lo, seg_b, bar);
        }
    }
}
BA_D void ft_issue_group2(const Params& P, const FtSmem& S, const TileHdr& T, unsigned long long* bar) {
    const int off_lo = T.p0 & ~3;
    const unsigned vi_b = (unsigned)T.m * 48u;
    const unsigned off_b = (unsigned)(((T.p0 + T.m + 1 - off_lo) * 4 + 15) & ~15);
    const unsigned rs_b = (unsigned)((T.maxlen * 4 + 15) & ~15);  // r0 is a multiple of 4
    mbar_expect_tx(bar, vi_b + off_b + rs_b);
    tma_load(S.sVi, P.pt_Vinv + (size_t)T.p0 * 6, vi_b, bar);
    tma_load(S.sOff, P.pt_off + off_lo, off_b, bar);
    if (rs_b) tma_load(S.sRs, P.ft_rs + T.r0, rs_b, bar);
}

// ACC (mat-vec only): the segment sums of an EPOCH (a few consecutive tiles of the CTA) are added up in shared
// memory, slot = (camera, split index) of the epoch, and leave ONE row per slot when the epoch ends - several
// times fewer rows for k_rows_finish to read than one per segment.  Fixed order (tile by tile), no atomics.
template <int MODE, bool ACC = false>
__global__ void __launch_bounds__(FT_THREADS, FT_CTAS_PER_SM) k_tile_pass(Params P, const double* __restrict__ vec) {
    extern __shared__ __align__(16) double sm[];
    __shared__ double red[FT_THREADS / 32 * 4];
    __shared__ double tot[4];
    __shared__ int hs[8];
    __shared__ int es[4];  // [0] first tile of the CTA, [1] epoch length, [2] first epoch, [3] 6 x slots of the current epoch
    if (MODE == 0 && P.st->pcg_done) return;
    int G = gridDim.x;
    if (MODE == 1 && P.cam_in_backsub) {
        // small problems (dense reduced system): the trial poses R <- R Exp(w), t <- t + R v are a 4 us chain of FP64 trigonometry
        // in one thread per camera; the last CTA of this launch runs it beside the point update instead of after the solve
        G -= 1;
        if ((int)blockIdx.x == G) {
            for (int c = threadIdx.x; c <= P.n_cams; c += FT_THREADS) camera_update_one(P, c);
            return;
        }
    }
    // contiguous run of tiles per CTA: neighbouring tiles share cameras, so the camera vectors stay in L1
    int t = ft_cta_begin(blockIdx.x, P.ft_tiles, G);
    const int t_end = ft_cta_begin(blockIdx.x + 1, P.ft_tiles, G);
    if (t >= t_end) return;
    const FtSmem S = ft_carve(sm, FT_ROWS_MV);
    const int b = P.st->cur;
    const int tid = threadIdx.x;
    TileHdr T = tile_hdr(P, t);
    if (ACC) {
        for (int l = tid; l < FT_ACC_CAP * 6; l += FT_THREADS) S.sAcc[l] = 0.0;
        if (tid == 0) { es[0] = t; es[1] = ft_epoch_len(t_end - t, P.ep_target); es[2] = P.cta_ep0[blockIdx.x]; }
    }
    if (tid == 0) {
        mbar_init(S.bar, 1);
        mbar_init(S.bar + 1, 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
        ft_issue_group1<MODE, ACC>(P, S, T, b, 0, S.bar);
        ft_issue_group2(P, S, T, S.bar + 1);
    }
    __syncthreads();

    int ep_pos = 0;  // tiles of the current epoch that are behind this one (a counter instead of (t - es[0]) % es[1]: a
                     // division by a run-time value in every thread and tile was 7 % of the pass's instructions)
    for (unsigned it = 0;; ++it) {
        const int buf = it & 1;
        const int tn = t + 1;
        const bool have_next = tn < t_end;
        TileHdr Tn = T;
        const bool issuer = tid == FT_THREADS - 32;  // first lane of the last warp, which is idle in phase B
        if (have_next && issuer) Tn = tile_hdr(P, tn);
        const bool act = tid < T.n;

        // ---- phase A: operands to registers; a = Jt (w x X + v); park Jt^T a at the point-major position --
        mbar_wait(S.bar, it & 1);
        int2 meta = make_int2(P.n_cams, 0);
        double2 j0 = make_double2(0.0, 0.0), j1 = j0, j2 = j0;  // (jx0, jx1) (jx2, jy0) (jy1, jy2)
        if (act) {
            meta = S.sMeta[(T.o0 & 1) + tid];
            j0 = S.sJ[tid]; j1 = S.sJ[FT_THREADS + tid]; j2 = S.sJ[2 * FT_THREADS + tid];
        }
        double x[6];
        load_vec6(vec, meta.x, x);
        double X[3] = {0.0, 0.0, 0.0};
        double ax = 0.0, ay = 0.0;
        const int pl = meta.y & 1023;
        if (act) {
            const int pos = (meta.y >> 10) & 1023;
            const double2 x01 = S.sX01[pl];
            X[0] = x01.x; X[1] = x01.y; X[2] = S.sX2[(T.p0 & 1) + pl];
            const double d0 = fma(x[1], X[2], fma(-x[2], X[1], x[3]));  // d = w x X + v
            const double d1 = fma(x[2], X[0], fma(-x[0], X[2], x[4]));
            const double d2 = fma(x[0], X[1], fma(-x[1], X[0], x[5]));
            ax = j0.x * d0 + j0.y * d1 + j1.x * d2;
            ay = j1.y * d0 + j2.x * d1 + j2.y * d2;
            S.sA[pos] = j0.x * ax + j1.y * ay;
            S.sA[FT_LDB + pos] = j0.y * ax + j2.x * ay;
            S.sA[2 * FT_LDB + pos] = j1.x * ax + j2.y * ay;
        }
        mbar_wait(S.bar + 1, it & 1);  // every thread sees group 2 complete BEFORE the barrier, so re-arming is safe after it
        __syncthreads();
        if (have_next && issuer) {
            // phase A has its operands in registers: observation records, Jacobian cache, points and the segment table of
            // the next tile may land (MODE 1 reads the points of phase B from HBM for that reason).  The next header goes to
            // shared memory here; everybody reads it after later barriers.
            ft_issue_group1<MODE, ACC>(P, S, Tn, b, buf ^ 1, S.bar);
            hs[0] = Tn.p0; hs[1] = Tn.m; hs[2] = Tn.o0; hs[3] = Tn.n; hs[4] = Tn.s0; hs[5] = Tn.ns; hs[6] = Tn.r0; hs[7] = Tn.maxlen;
        }
        if (ACC && tid >= FT_THREADS - 32 && ep_pos == 0) {
            // first tile of an epoch: the last warp (idle in phase B) fetches the rows of the epoch's slots; the
            // previous epoch was flushed before the barrier above, this one is flushed after later barriers
            const int ep = es[2] + (t - es[0]) / es[1];
            const int g0 = __ldg(P.ep_g0 + ep), n = __ldg(P.ep_g0 + ep + 1) - g0;
            for (int l = tid - (FT_THREADS - 32); l < n; l += 32) S.sSlot[l] = __ldg(P.slot_row + g0 + l);
            if (issuer) es[3] = n * 6;
        }

        // ---- phase B: four lanes per point: lane k < 3 adds component k of the track in observation
        //      order, the quad exchanges the three sums and lane k forms row k of Vinv * s --------------
        double acc[4] = {0.0, 0.0, 0.0, 0.0};  // MODE 1: g.dp, dp D dp, |dp|^2, |X|^2
        {
            const int k = tid & 3, lane = tid & 31;
            const int* off = S.sOff + (T.p0 & 3);
            const int* rsb = S.sRs;
            const double* sVi = S.sVi;
            for (int q0 = (tid & ~31) >> 2; q0 < T.m; q0 += FT_THREADS / 4) {  // warp-uniform bounds
                const int q = q0 + (lane >> 2);
                const bool valid = q < T.m;
                // transposed track layout: observation i of point q sits at rs[i] + q, so the lanes of
                // neighbouring points read neighbouring words; tracks of a warp are (nearly) equally long
                // idle lanes (fourth lane of a quad, points past the tile's last) shadow a working lane - same address, a
                // broadcast - so that the head of the loop, up to the shortest track of the warp (the tracks of a warp are
                // sorted and nearly equally long), needs no predicates; their sums are never used
                const int qs = min(q, T.m - 1);
                const int len_s = off[qs + 1] - off[qs];
                const int len = (valid && k < 3) ? len_s : 0;
                const int wmax = __reduce_max_sync(0xffffffffu, len);
                const int whead = __reduce_min_sync(0xffffffffu, len_s) & ~3;
                const double* src = S.sA + min(k, 2) * FT_LDB + qs;
                double sk = 0.0;
                for (int i0 = 0; i0 < whead; i0 += 4) {  // loads first, then the (ordered) adds
                    const int4 r4 = *reinterpret_cast<const int4*>(rsb + i0);  // broadcast
                    const double v0 = src[r4.x], v1 = src[r4.y], v2 = src[r4.z], v3 = src[r4.w];
                    sk += v0; sk += v1; sk += v2; sk += v3;
                }
                for (int i0 = whead; i0 < wmax; i0 += 4) {  // the tail: x + 0.0 is exact
                    const int4 r4 = *reinterpret_cast<const int4*>(rsb + i0);  // broadcast
                    const int rs[4] = {r4.x, r4.y, r4.z, r4.w};
                    double v[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) v[j] = (i0 + j < len) ? src[rs[j]] : 0.0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) sk += v[j];
                }
                const double s0 = __shfl_sync(0xffffffffu, sk, (lane & ~3));
                const double s1 = __shfl_sync(0xffffffffu, sk, (lane & ~3) + 1);
                const double s2 = __shfl_sync(0xffffffffu, sk, (lane & ~3) + 2);
                double zk = 0.0;
                if (valid && k < 3) {
                    const double* Vi = sVi + 6 * q;  // 00 01 02 11 12 22
                    const double v0 = Vi[k], v1 = Vi[k == 0 ? 1 : (k == 1 ? 3 : 4)], v2 = Vi[k == 2 ? 5 : (k == 1 ? 4 : 2)];
                    zk = v0 * s0 + v1 * s1 + v2 * s2;
                }
                if (MODE == 0) {
                    if (valid && k < 2) reinterpret_cast<double*>(S.sZ01)[2 * q + k] = zk;
                    if (valid && k == 2) S.sZ2[q] = zk;
                } else if (valid && k < 3) {
                    const size_t p = (size_t)(T.p0 + q), p4 = p * 4 + k;
                    const double Xk = __ldg(pt_base(P, b) + (k < 2 ? 2 * p + k : 2 * (size_t)P.n_pad + p));
                    const double dp = -__ldg(P.pt_h + p4) - zk;
                    const double gk = __ldg(P.pt_g + (size_t)b * P.n_pts * 4 + p4), Dk = __ldg(P.pt_D + p4);
                    double* out = P.pt_X + (size_t)(b ^ 1) * P.pt_stride;
                    out[k < 2 ? 2 * p + k : 2 * (size_t)P.n_pad + p] = Xk + dp;
                    acc[0] += gk * dp; acc[1] += Dk * dp * dp; acc[2] += dp * dp; acc[3] += Xk * Xk;
                }
            }
        }
        if (MODE == 1) {
            block_sum<4, FT_THREADS>(acc, red, tot);
            __syncthreads();
            if (tid == 0) {
                double* out = P.pm_partial + (size_t)t * NSCAL;
                out[S_GDP] = tot[0]; out[S_DDP] = tot[1]; out[S_DP2] = tot[2]; out[S_XP2] = tot[3];
            }
            if (have_next && issuer) ft_issue_group2(P, S, Tn, S.bar + 1);  // phase B is over: V^-1 and the track tables may be refilled
            if (!have_next) break;
            __syncthreads();
        } else {
            __syncthreads();
            // phase B is over: V^-1, track offsets and track rows of the next tile may land (single-buffered)
            if (have_next && issuer) ft_issue_group2(P, S, Tn, S.bar + 1);

            // ---- phase C: camera-side product of every observation, parked at its tile-camera position --
            if (act) {
                const double2 z01 = S.sZ01[pl];
                const double z2 = S.sZ2[pl];
                const double bx = ax - (j0.x * z01.x + j0.y * z01.y + j1.x * z2);
                const double by = ay - (j1.y * z01.x + j2.x * z01.y + j2.y * z2);
                const double mm[3] = {j0.x * bx + j1.y * by, j0.y * bx + j2.x * by, j1.x * bx + j2.y * by};
                double xm[3];
                cross3(X, mm, xm);
                double2* o2 = reinterpret_cast<double2*>(S.sO) + tid;  // three 16-byte stores instead of six 8-byte ones
                o2[0] = make_double2(xm[0], xm[1]); o2[FT_LD] = make_double2(xm[2], mm[0]); o2[2 * FT_LD] = make_double2(mm[1], mm[2]);
            }
            __syncthreads();

            // ---- phase D: one thread per (segment, pair of components): fixed-order sum -> one row per segment ---
            if constexpr (ACC) {
                // Work items (segment, pair) packed by k_ft_item_fill: the eight lanes of a quarter-warp start in eight
                // different 16-byte bank groups, so every step of the loop below is one conflict-free wavefront per
                // quarter-warp (lanes side by side in the segment-major numbering collide 2.3-fold on average).
                // item = first position | length << 8 | pair << 13 | accumulator slot << 15; 0 = empty lane.
                const int* items = S.sSeg + buf * FT_ITEM_N;
                for (int w0 = tid & ~31; w0 < T.ns; w0 += FT_THREADS) {  // warp-uniform bounds
                    const int w = w0 + (tid & 31);
                    const int item = w < T.ns ? items[w] : 0;
                    const int a = item & 255, len = (item >> 8) & 31, pr = (item >> 13) & 3;
                    const int wmax = __reduce_max_sync(0xffffffffu, len);
                    const double2* src = reinterpret_cast<const double2*>(S.sO) + pr * FT_LD + a;
                    // a slot occurs at most once per tile; its running sum is fetched ahead of the segment's loop
                    double2* slot = reinterpret_cast<double2*>(S.sAcc + (item >> 15) * 6 + 2 * pr);
                    const double2 before = len ? *slot : make_double2(0.0, 0.0);
                    double sx = 0.0, sy = 0.0;
                    for (int j0 = 0; j0 < wmax; j0 += 2) {
                        const double2 z2 = make_double2(0.0, 0.0);
                        const double2 v0 = (j0 < len) ? src[j0] : z2, v1 = (j0 + 1 < len) ? src[j0 + 1] : z2;
                        sx += v0.x; sy += v0.y;
                        sx += v1.x; sy += v1.y;
                    }
                    if (len) *slot = make_double2(before.x + sx, before.y + sy);
                }
            } else {
                const int* seg_i = S.sSeg + buf * FT_SEG_N + (T.s0 & 3);
                const int* seg_r = S.sRow + buf * FT_SEG_N + (T.s0 & 3);
                for (int w0 = tid & ~31; w0 < T.ns * 3; w0 += FT_THREADS) {  // warp-uniform bounds
                    const int w = w0 + (tid & 31);
                    const bool valid = w < T.ns * 3;
                    const int seg = w / 3, pr = w - 3 * seg;
                    const int info = valid ? seg_i[seg] : 0;  // first tile-camera position | length << 16; longest first
                    const int a = info & 0xffff, len = info >> 16;
                    const int wmax = __reduce_max_sync(0xffffffffu, len);
                    const double2* src = reinterpret_cast<const double2*>(S.sO) + pr * FT_LD + a;
                    double sx = 0.0, sy = 0.0;
                    for (int j0 = 0; j0 < wmax; j0 += 2) {
                        const double2 z2 = make_double2(0.0, 0.0);
                        const double2 v0 = (j0 < len) ? src[j0] : z2, v1 = (j0 + 1 < len) ? src[j0 + 1] : z2;
                        sx += v0.x; sy += v0.y;
                        sx += v1.x; sy += v1.y;
                    }
                    if (valid) *reinterpret_cast<double2*>(P.rows + (size_t)seg_r[seg] * 6 + 2 * pr) = make_double2(sx, sy);
                }
            }
        }
        // end of an epoch: its slots become rows (camera-major, like the per-segment rows).  Every thread flushes and
        // clears the same words; the next tile adds to them only after further barriers.
        auto flush = [&]() {
            const int n6 = es[3];
            for (int l = tid; l < n6; l += FT_THREADS) {
                const int slot = l / 6;
                P.rows[(size_t)S.sSlot[slot] * 6 + (l - 6 * slot)] = S.sAcc[l];
                S.sAcc[l] = 0.0;
            }
        };
        if (MODE == 0) {
            // No barrier between phase D of this tile and phase A of the next: they use different scratch (sO / sA), and
            // everything the next phase A reads arrived through the TMA.  Only the end of an epoch needs all of phase D.
            const bool epoch_end = ACC && (!have_next || ep_pos + 1 == es[1]);
            ep_pos = epoch_end ? 0 : ep_pos + 1;
            if (epoch_end) { __syncthreads(); flush(); }
            if (!have_next) break;
        }
        // the next tile's header was put into shared memory by the issuing thread before the last barrier(s)
        T.p0 = hs[0]; T.m = hs[1]; T.o0 = hs[2]; T.n = hs[3]; T.s0 = hs[4]; T.ns = hs[5]; T.r0 = hs[6]; T.maxlen = hs[7];
        t = tn;
    }
}

// ---------------------------------------------------------------------------------------------
// Point-side linearisation on the tiles (replaces k_lin_points when the tile layout exists): one CTA per
// tile, one thread per observation in tile-camera order.  Against the point-major kernel the camera pose
// is a broadcast load inside a camera run (not one gather per lane), the Jacobian cache is written at the
// thread's own position (coalesced, not scattered through pm_tc) and the measurement comes from tc_uv.
// The nine per-point sums (V 6, g_p 3) use the transposed track layout of the mat-vec: observation i of
// local point q parks at rs[i] + q, thread (q, value) adds the track in observation order.
// ---------------------------------------------------------------------------------------------
constexpr int FT_LDL = FT_THREADS + 2;  // 2 mod 16 words: the nine values of two neighbouring points fall on (nearly) disjoint banks

BA_D void cp_async8(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
BA_D void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}

// Per-point line search on the trial step (LS = true: the linearisation of the TRIAL point, after back-substitution and
// camera update).  With the trial cameras held fixed the cost is a sum over the points, so every point may scale ITS step
// dp by its own factor: X = X_cur + alpha dp, alpha in {1, 1/2, 2, 4, 8, 16}, whichever has the smallest Huber cost over the
// point's observations (alpha = 1 unless another is better by more than 1e-9 relative).  The total cost can only go down
// against the plain trial point, so the accept / reject logic is unchanged.  Why: under IRLS a point whose observations sit
// in the linear regime of the Huber loss (an outlier pulling along a poorly constrained depth, a point running to infinity)
// moves by a small fraction of the distance to its piecewise-linear optimum per LM iteration - gain ratios of 2 for dozens
// of iterations; the search lets those few points jump (CPU statement: oracle/lm_schur.py::point_linesearch).  The six
// candidate costs of an observation go through the same transposed track layout as the nine sums of the linearisation
// (before them, in the same scratch), the chosen position replaces the trial point in HBM and is linearised right away.
#ifdef BA_DN_STAMPS
__device__ unsigned long long g_ls_moved[8];  // experiment builds: points whose line search chose candidate k
#endif
constexpr int LS_K = 6;
__device__ __constant__ double kLsAlpha[LS_K] = {1.0, 0.5, 2.0, 4.0, 8.0, 16.0};

// Persistent CTAs over contiguous runs of tiles.  Everything a tile reads from HBM except the camera poses - the records and
// measurements of its observations, its points (LS: the current points too), its track offsets and track rows - is staged
// for the NEXT tile with cp.async while the current tile is computed; a tile ends with cp.async.wait_group 0 and the CTA
// barrier it needs anyway, which publishes the next tile's data to all threads.  The tile headers are fetched two tiles
// ahead.  (Before the points were staged, the first use of a point gathered from HBM / L2 and the track-table loads were
// 14-18 % of the stall samples of this kernel, profiles/r02z.)
BA_D void cp_async4(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
template <bool LS>
__global__ void __launch_bounds__(FT_THREADS, FT_CTAS_PER_SM) k_tile_lin(Params P, int which) {
    __shared__ double sA[9 * FT_LDL];  // LS: planes 0-5 hold the candidate costs, planes 6-8 their per-point sums (sC)
    __shared__ __align__(16) double2 sUv[2][FT_THREADS];
    __shared__ __align__(16) int2 sMeta[2][FT_THREADS];
    __shared__ __align__(16) double2 sP01[2][FT_PTS];  // the tile's points in parameter buffer b (LS: replaced by the chosen positions)
    __shared__ double sP2[2][FT_PTS];
    __shared__ __align__(16) double2 sQ01[LS ? 2 : 1][LS ? FT_PTS : 1];  // LS: the current points (buffer b ^ 1)
    __shared__ double sQ2[LS ? 2 : 1][LS ? FT_PTS : 1];
    __shared__ int sRs[2][FT_THREADS];
    __shared__ int sOff[2][FT_PTS + 1];
    __shared__ double red[FT_THREADS / 32 * 3];
    __shared__ double tot[3];
    static_assert(LS_K * FT_PTS <= 3 * FT_LDL, "candidate sums alias three planes of the scratch");
    double* const sC = sA + 6 * FT_LDL;
    const int b = resolve_buf(P.st, which);
    const int tid = threadIdx.x;
    int t = ft_cta_begin(blockIdx.x, P.ft_tiles, gridDim.x);
    const int t_end = ft_cta_begin(blockIdx.x + 1, P.ft_tiles, gridDim.x);
    const double* camR = P.cam_R + (size_t)b * (P.n_cams + 1) * 12;
    double2* const Jb = P.tc_J + (size_t)b * 3 * P.n_obs;
    double* const Vb = P.pt_V + (size_t)b * P.n_pts * 6;
    double* const Gb = P.pt_g + (size_t)b * P.n_pts * 4;
    double acc[3] = {0.0, 0.0, 0.0};  // cost, raw, bad
    double gmax = 0.0;
    auto hdr_rows = [&](int tile) { return __ldg(reinterpret_cast<const int2*>(P.ft_hdr + 2 * tile + 1) + 1); };
    auto stage = [&](int buf, const int4& h, const int2& g) {  // h = (p0, m, o0, n), g = (first track row, longest track)
        if (tid < h.w) {
            cp_async16(&sUv[buf][tid], P.tc_uv + h.z + tid);
            cp_async8(&sMeta[buf][tid], P.tc_meta + h.z + tid);
        }
        if (tid < h.y) {
            const double* xb = pt_base(P, b);
            cp_async16(&sP01[buf][tid], reinterpret_cast<const double2*>(xb) + h.x + tid);
            cp_async8(&sP2[buf][tid], xb + 2 * (size_t)P.n_pad + h.x + tid);
            if constexpr (LS) {
                const double* xc = pt_base(P, b ^ 1);
                cp_async16(&sQ01[buf][tid], reinterpret_cast<const double2*>(xc) + h.x + tid);
                cp_async8(&sQ2[buf][tid], xc + 2 * (size_t)P.n_pad + h.x + tid);
            }
        }
        if (tid <= h.y) cp_async4(&sOff[buf][tid], P.pt_off + h.x + tid);
        if (tid < g.y) cp_async4(&sRs[buf][tid], P.ft_rs + g.x + tid);
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };
    int4 ha = make_int4(0, 0, 0, 0), han = ha;
    int2 hbn = make_int2(0, 0);
    if (t < t_end) {
        ha = __ldg(P.ft_hdr + 2 * t);
        han = ha;
        if (t + 1 < t_end) { han = __ldg(P.ft_hdr + 2 * (t + 1)); hbn = hdr_rows(t + 1); }
        stage(0, ha, hdr_rows(t));
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        __syncthreads();
    }
    for (int it = 0; t < t_end; ++t, ++it) {
        const int buf = it & 1;
        int4 hann = han;
        int2 hbnn = hbn;
        if (t + 2 < t_end) { hann = __ldg(P.ft_hdr + 2 * (t + 2)); hbnn = hdr_rows(t + 2); }
        const int p0 = ha.x, m = ha.y, o0 = ha.z, n = ha.w;
        if (t + 1 < t_end) stage(buf ^ 1, han, hbn);  // into the buffers tile t - 1 used: everybody has passed its last barrier
        const int* const rs = sRs[buf];
        const int* const off = sOff[buf];
        if constexpr (LS) {
            if (tid < n) {
                const int2 meta = sMeta[buf][tid];
                const double2 uv = sUv[buf][tid];
                const int pos = (meta.y >> 10) & 1023, pl = meta.y & 1023;
                double Rt[12];
                load_cam(camR, meta.x, Rt);
                const double2 x01 = sP01[buf][pl], q01 = sQ01[buf][pl];
                const double X0[3] = {q01.x, q01.y, sQ2[buf][pl]};
                const double d[3] = {x01.x - X0[0], x01.y - X0[1], sP2[buf][pl] - X0[2]};
                // x_cam(alpha) = (R X0 + t) + alpha R d
                const double ax = fma(Rt[0], X0[0], fma(Rt[1], X0[1], fma(Rt[2], X0[2], Rt[9])));
                const double ay = fma(Rt[3], X0[0], fma(Rt[4], X0[1], fma(Rt[5], X0[2], Rt[10])));
                const double az = fma(Rt[6], X0[0], fma(Rt[7], X0[1], fma(Rt[8], X0[2], Rt[11])));
                const double dx = fma(Rt[0], d[0], fma(Rt[1], d[1], Rt[2] * d[2]));
                const double dy = fma(Rt[3], d[0], fma(Rt[4], d[1], Rt[5] * d[2]));
                const double dz = fma(Rt[6], d[0], fma(Rt[7], d[1], Rt[8] * d[2]));
#pragma unroll
                for (int k = 0; k < LS_K; ++k) {
                    const double al = kLsAlpha[k];
                    // candidates only have to be RANKED (a factor other than 1 must win by 1e-9 relative): the reciprocal of
                    // the depth by two Newton steps from the single-precision one (~ 2^-45) instead of a full FP64 division,
                    // which was most of this loop's instructions
                    const double z = fma(al, dz, az);
                    double iz = (double)__frcp_rn((float)z);
                    iz = iz * fma(-z, iz, 2.0);
                    iz = iz * fma(-z, iz, 2.0);
                    const double rx = uv.x - fma(P.K.fx, fma(al, dx, ax) * iz, P.K.cx);
                    const double ry = uv.y - fma(P.K.fy, fma(al, dy, ay) * iz, P.K.cy);
                    const double axr = fabs(rx), ayr = fabs(ry);
                    const double cx_ = (P.mode == ROBUST_NONE || axr <= P.delta) ? rx * rx : P.delta * (2.0 * axr - P.delta);
                    const double cy_ = (P.mode == ROBUST_NONE || ayr <= P.delta) ? ry * ry : P.delta * (2.0 * ayr - P.delta);
                    const double c = 0.5 * (cx_ + cy_);
                    sA[k * FT_LDL + pos] = isfinite(c) ? c : 1e300;
                }
            }
            __syncthreads();
            for (int w = tid; w < m * LS_K; w += FT_THREADS) {
                const int q = w / LS_K, k = w - LS_K * q;
                const int len = off[q + 1] - off[q];
                const double* src = sA + k * FT_LDL + q;
                double sk = 0.0;
                for (int i = 0; i < len; ++i) sk += src[rs[i]];
                sC[k * FT_PTS + q] = sk;
            }
            __syncthreads();
            if (tid < m) {
                const double c1 = sC[tid];
                const double thr = c1 - 1e-9 * fabs(c1);
                double best = c1;
                int kb = 0;
#pragma unroll
                for (int k = 1; k < LS_K; ++k) {
                    const double c = sC[k * FT_PTS + tid];
                    if (c < thr && c < best) { best = c; kb = k; }
                }
#ifdef BA_DN_STAMPS
                atomicAdd(&g_ls_moved[kb], 1ull);
#endif
                if (kb != 0) {
                    const double al = kLsAlpha[kb];
                    const double2 x01 = sP01[buf][tid], q01 = sQ01[buf][tid];
                    const double X0[3] = {q01.x, q01.y, sQ2[buf][tid]};
                    double X1[3] = {x01.x, x01.y, sP2[buf][tid]};
#pragma unroll
                    for (int k = 0; k < 3; ++k) X1[k] = fma(al, X1[k] - X0[k], X0[k]);
                    store_pt(P, b, p0 + tid, X1[0], X1[1], X1[2]);
                    sP01[buf][tid] = make_double2(X1[0], X1[1]);
                    sP2[buf][tid] = X1[2];
                }
            }
            __syncthreads();  // the chosen positions; and the candidate costs and their sums in sA have all been read
        }
        if (tid < n) {
            const int2 meta = sMeta[buf][tid];
            const double2 uv = sUv[buf][tid];
            const int pos = (meta.y >> 10) & 1023, pl = meta.y & 1023;
            double Rt[12];
            load_cam(camR, meta.x, Rt);  // (LS: again - from L1 - rather than twelve registers held across three barriers)
            const double2 x01 = sP01[buf][pl];
            const double X[3] = {x01.x, x01.y, sP2[buf][pl]};
            ObsLin o;
            linearize_obs(Rt, X, uv.x, uv.y, P.K, o);
            double wHx, wgx, r0x, wHy, wgy, r0y;
            robust_weights(o.rx, P.delta, P.mode, wHx, wgx, r0x);
            robust_weights(o.ry, P.delta, P.mode, wHy, wgy, r0y);
            acc[0] += 0.5 * (r0x + r0y);
            acc[1] += o.rx * o.rx + o.ry * o.ry;
            if (!(isfinite(o.rx) && isfinite(o.ry))) acc[2] += 1.0;
            const double sx = sqrt(wHx), sy = sqrt(wHy);
            double2* J = Jb + o0 + tid;
            J[0] = make_double2(sx * o.jx[0], sx * o.jx[1]);
            J[P.n_obs] = make_double2(sx * o.jx[2], sy * o.jy[0]);
            J[2 * (size_t)P.n_obs] = make_double2(sy * o.jy[1], sy * o.jy[2]);
            const double ax = wHx * o.jx[0], bx = wHx * o.jx[1], cx = wHx * o.jx[2];
            const double ay = wHy * o.jy[0], by = wHy * o.jy[1], cy = wHy * o.jy[2];
            double* dst = sA + pos;
            dst[0 * FT_LDL] = ax * o.jx[0] + ay * o.jy[0];
            dst[1 * FT_LDL] = ax * o.jx[1] + ay * o.jy[1];
            dst[2 * FT_LDL] = ax * o.jx[2] + ay * o.jy[2];
            dst[3 * FT_LDL] = bx * o.jx[1] + by * o.jy[1];
            dst[4 * FT_LDL] = bx * o.jx[2] + by * o.jy[2];
            dst[5 * FT_LDL] = cx * o.jx[2] + cy * o.jy[2];
            const double gx = wgx * o.rx, gy = wgy * o.ry;
            if (P.tc_r) P.tc_r[(size_t)b * P.n_obs + o0 + tid] = make_double2(gx / sx, gy / sy);
            dst[6 * FT_LDL] = o.jx[0] * gx + o.jy[0] * gy;
            dst[7 * FT_LDL] = o.jx[1] * gx + o.jy[1] * gy;
            dst[8 * FT_LDL] = o.jx[2] * gx + o.jy[2] * gy;
        }
        __syncthreads();
        double* V = Vb + (size_t)p0 * 6;
        double* G = Gb + (size_t)p0 * 4;
        for (int w = tid; w < m * 9; w += FT_THREADS) {
            const int q = w / 9, k = w - 9 * q;
            const int len = off[q + 1] - off[q];
            const double* src = sA + k * FT_LDL + q;
            double sk = 0.0;
            for (int i = 0; i < len; ++i) sk += src[rs[i]];
            if (k < 6) {
                V[q * 6 + k] = sk;
            } else {
                G[q * 4 + (k - 6)] = sk;
                gmax = fmax(gmax, fabs(sk));
            }
        }
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");  // the next tile's data (staged at the top of this one)
        __syncthreads();  // ... becomes visible to everybody; and the scratch is rewritten by the next tile
        ha = han; han = hann; hbn = hbnn;
    }
    block_sum<3, FT_THREADS>(acc, red, tot);
    __syncthreads();
    const double wm = warp_max(gmax);
    if ((tid & 31) == 0) red[tid >> 5] = wm;
    __syncthreads();
    if (tid == 0) {
        double mx = 0.0;
        for (int i = 0; i < FT_THREADS / 32; ++i) mx = fmax(mx, red[i]);
        double* out = P.pm_partial + (size_t)blockIdx.x * NSCAL;
        out[S_COST] = tot[0]; out[S_RAW] = tot[1]; out[S_BAD] = tot[2]; out[S_GPMAX] = mx;
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
// Transposed track layout of a tile (its points are ordered by track length, longest first):
// observation i of local point q is parked at rs[i] + q, rs[i] = sum over i' < i of #{points longer than i'}.
__global__ void __launch_bounds__(FT_THREADS) k_ft_rowstart(const int* __restrict__ ft_pt, const int* __restrict__ pt_off,
                                                            const int* __restrict__ rs0, const int* __restrict__ maxlen,
                                                            int* __restrict__ rs) {
    typedef cub::BlockScan<int, FT_THREADS> Scan;
    __shared__ typename Scan::TempStorage tmp;
    const int t = blockIdx.x, i = threadIdx.x;
    const int p0 = ft_pt[t], m = ft_pt[t + 1] - p0, ml = maxlen[t];
    int mi = 0;
    if (i < ml) {  // first q whose track is not longer than i
        int lo = 0, hi = m;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (pt_off[p0 + mid + 1] - pt_off[p0 + mid] > i) lo = mid + 1; else hi = mid;
        }
        mi = lo;
    }
    int ex;
    Scan(tmp).ExclusiveSum(mi, ex);
    if (i < ml) rs[rs0[t] + i] = ex;
}
__global__ void k_ft_fill(int n_obs, const unsigned long long* __restrict__ key_s, const int* __restrict__ src,
                          const int* __restrict__ flag, const int* __restrict__ seg_inc, const int* __restrict__ ft_pt,
                          const int* __restrict__ pt_off, const int* __restrict__ rs0, const int* __restrict__ rs,
                          const double2* __restrict__ pm_uv, const int2* __restrict__ pm_idx, double2* __restrict__ tc_uv,
                          int2* __restrict__ tc_meta, int* __restrict__ pm_tc, int* __restrict__ seg_cam,
                          int* __restrict__ seg_start, const int* __restrict__ run_start, int* __restrict__ seg_split) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    const int j = src[i];
    const int t = (int)(key_s[i] >> 32), cam = (int)(unsigned)key_s[i];
    const int s = seg_inc[i] - 1;
    const int p = pm_idx[j].y;
    const unsigned pl = (unsigned)(p - ft_pt[t]);
    const unsigned pos = (unsigned)rs[rs0[t] + (j - pt_off[p])] + pl;
    tc_uv[i] = pm_uv[j];
    tc_meta[i] = make_int2(cam, (int)(pl | (pos << 10)));
    pm_tc[j] = i;
    if (flag[i]) { seg_cam[s] = cam; seg_start[s] = i; seg_split[s] = (i - run_start[i]) / FT_SEG_MAX; }
}
// Segments of a tile are renumbered longest first (so that the lanes that sum segments side by side run
// equally long loops): sort key = (tile, FT_SEG_MAX - length).
__global__ void k_ft_seg_keys(int n_seg, int n_tiles, const int* __restrict__ seg_start, const int* __restrict__ ft_obs,
                              unsigned* __restrict__ key, int* __restrict__ iota) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_seg) return;
    const int a = seg_start[s], len = seg_start[s + 1] - a;
    int lo = 0, hi = n_tiles;  // last tile with ft_obs[t] <= a
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (ft_obs[mid] <= a) lo = mid; else hi = mid;
    }
    key[s] = ((unsigned)lo << 5) | (unsigned)(FT_SEG_MAX - len);
    iota[s] = s;
}
__global__ void k_ft_seg_apply(int n_seg, const unsigned* __restrict__ key_s, const int* __restrict__ order,
                               const int* __restrict__ seg_start, const int* __restrict__ seg_cam_in,
                               const int* __restrict__ ft_obs, int* __restrict__ seg_info, int* __restrict__ seg_cam,
                               unsigned* __restrict__ cam_key, int* __restrict__ iota, const int* __restrict__ seg_split_in,
                               int* __restrict__ seg_tile, int* __restrict__ seg_split) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_seg) return;
    const int s = order[r], t = (int)(key_s[r] >> 5);
    const int a = seg_start[s], len = seg_start[s + 1] - a;
    seg_info[r] = (a - ft_obs[t]) | (len << 16);
    seg_cam[r] = seg_cam_in[s];
    cam_key[r] = (unsigned)seg_cam_in[s];
    iota[r] = r;
    seg_tile[r] = t;
    seg_split[r] = seg_split_in[s];
}
__global__ void k_ft_headers(int n_tiles, const int* __restrict__ ft_pt, const int* __restrict__ ft_obs,
                             const int* __restrict__ ft_seg, const int* __restrict__ rs0, const int* __restrict__ maxlen,
                             int4* __restrict__ hdr) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tiles) return;
    hdr[2 * t] = make_int4(ft_pt[t], ft_pt[t + 1] - ft_pt[t], ft_obs[t], ft_obs[t + 1] - ft_obs[t]);
    hdr[2 * t + 1] = make_int4(ft_seg[t], ft_seg[t + 1] - ft_seg[t], rs0[t], maxlen[t]);
}
__global__ void k_ft_seg_rows(int n_seg, const int* __restrict__ sorted_seg, int* __restrict__ seg_row) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n_seg) seg_row[sorted_seg[r]] = r;
}
// ---- epoch accumulator slots ------------------------------------------------------------------------
// key of segment r = (first tile of its epoch << 40) | (camera << 8) | split index of the segment in its run
__global__ void k_ft_slot_keys(int n_seg, int n_tiles, int grid, int target, const int* __restrict__ seg_tile,
                               const int* __restrict__ seg_cam, const int* __restrict__ seg_split,
                               unsigned long long* __restrict__ key, int* __restrict__ iota) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_seg) return;
    const int t = seg_tile[r];
    const int b = ft_cta_of_tile(t, n_tiles, grid);
    const int t0 = ft_cta_begin(b, n_tiles, grid), t1 = ft_cta_begin(b + 1, n_tiles, grid);
    const int el = ft_epoch_len(t1 - t0, target);
    const int ts = t0 + ((t - t0) / el) * el;
    key[r] = ((unsigned long long)ts << 40) | ((unsigned long long)(unsigned)seg_cam[r] << 8) | (unsigned)seg_split[r];
    iota[r] = r;
}
__global__ void k_ft_slot_heads(int n_seg, const unsigned long long* __restrict__ key_s, int* __restrict__ head) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_seg) head[i] = (i == 0 || key_s[i] != key_s[i - 1]) ? 1 : 0;
}
// first slot of every epoch (ep_ts = first tile of the epoch, ascending; one extra entry = number of slots)
__global__ void k_ft_epoch_base(int n_ep, int n_seg, int n_slots, const int* __restrict__ ep_ts,
                                const unsigned long long* __restrict__ key_s, const int* __restrict__ slot_inc,
                                int* __restrict__ ep_g0) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e > n_ep) return;
    if (e == n_ep) { ep_g0[e] = n_slots; return; }
    const unsigned long long want = (unsigned long long)ep_ts[e];
    int lo = 0, hi = n_seg;  // first position whose epoch tile >= want
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if ((key_s[mid] >> 40) < want) lo = mid + 1; else hi = mid;
    }
    ep_g0[e] = lo < n_seg ? slot_inc[lo] - 1 : n_slots;
}
// local slot of every segment, camera of every slot
__global__ void k_ft_slot_assign(int n_seg, int n_ep, const unsigned long long* __restrict__ key_s, const int* __restrict__ order,
                                 const int* __restrict__ slot_inc, const int* __restrict__ ep_ts, const int* __restrict__ ep_g0,
                                 int* __restrict__ seg_slot, unsigned* __restrict__ slot_cam, int* __restrict__ iota) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_seg) return;
    const int g = slot_inc[i] - 1;
    const int ts = (int)(key_s[i] >> 40);
    int lo = 0, hi = n_ep;  // last epoch with ep_ts <= ts
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (ep_ts[mid] <= ts) lo = mid; else hi = mid;
    }
    seg_slot[order[i]] = g - ep_g0[lo];
    if (i == 0 || key_s[i] != key_s[i - 1]) { slot_cam[g] = (unsigned)((key_s[i] >> 8) & 0xffffffffu); iota[g] = g; }
}
// Which observation of a point takes which slot of its track is free (the sums run over the slots in order, any fixed
// assignment is deterministic), and it decides the shared-memory banks of the per-observation scatter of phases A (three
// stores in the mat-vec pass, nine to fifteen in the linearisation): a half-warp of camera-ordered observations writes to
// banks (rs[i] + q) mod 16.  With i = the observation's index in the point-major order the sixteen writes collide 2.8-fold
// on average.  This pass re-deals the slots, one warp per tile: half-warp by half-warp, the observation whose point has
// the fewest free slots first, each taking the free slot of its point whose bank is least loaded so far in the half-warp.
__global__ void __launch_bounds__(32) k_ft_slot_balance(const int* __restrict__ ft_pt, const int* __restrict__ ft_obs,
                                                        const int* __restrict__ pt_off, const int* __restrict__ rs0,
                                                        const int* __restrict__ rs, int2* __restrict__ tc_meta) {
    __shared__ unsigned s_free[FT_PTS][FT_THREADS / 32];  // free slots of every point of the tile
    __shared__ int s_left[FT_PTS];
    __shared__ int s_rs[FT_THREADS];
    const int t = blockIdx.x, lane = threadIdx.x;
    const int p0 = ft_pt[t], m = ft_pt[t + 1] - p0, o0 = ft_obs[t], n = ft_obs[t + 1] - o0;
    if (m <= 0 || n <= 0) return;
    const int maxlen = pt_off[p0 + 1] - pt_off[p0];  // longest track first
    for (int q = lane; q < m; q += 32) {
        const int len = pt_off[p0 + q + 1] - pt_off[p0 + q];
        s_left[q] = len;
#pragma unroll
        for (int w = 0; w < FT_THREADS / 32; ++w) {
            const int k = len - 32 * w;
            s_free[q][w] = k >= 32 ? 0xffffffffu : (k > 0 ? (1u << k) - 1u : 0u);
        }
    }
    for (int i = lane; i < maxlen; i += 32) s_rs[i] = rs[rs0[t] + i];
    __syncwarp();
    const int words = (maxlen + 31) >> 5;
    for (int h = 0; h < n; h += 16) {
        int load = 0;  // lane b < 16: writes of this half-warp that fall on bank b so far
        int2 meta = make_int2(0, 0);
        int myq = 0, mykey = 0x7fffffff, mypos = 0;
        const bool mine = lane < 16 && h + lane < n;
        if (mine) {
            meta = tc_meta[o0 + h + lane];
            myq = meta.y & 1023;
            mykey = s_left[myq] * 16 + lane;
        }
        for (int step = 0; step < 16; ++step) {
            const int kmin = __reduce_min_sync(0xffffffffu, mykey);
            if (kmin == 0x7fffffff) break;
            const int l = kmin & 15;
            const int q = __shfl_sync(0xffffffffu, myq, l);
            int best = 0x7fffffff;
            for (int w = 0; w < words; ++w) {
                const bool cand = (s_free[q][w] >> lane) & 1u;
                const int i = 32 * w + lane;
                const int bank = cand ? (s_rs[i] + q) & 15 : 0;
                const int ld = __shfl_sync(0xffffffffu, load, bank);
                if (cand) best = min(best, ld * 256 + i);
            }
            best = __reduce_min_sync(0xffffffffu, best);
            const int i = best & 255, pos = s_rs[i] + q;
            if (lane == (pos & 15)) ++load;
            if (lane == 0) { s_free[q][i >> 5] &= ~(1u << (i & 31)); --s_left[q]; }
            if (lane == l) { mypos = pos; mykey = 0x7fffffff; }
            __syncwarp();
        }
        if (mine) tc_meta[o0 + h + lane] = make_int2(meta.x, (meta.y & 1023) | (mypos << 10));
    }
}

// ---- packed work items of phase D (mat-vec pass with epoch accumulators) ----------------------------------------------
// A work item = (segment, pair of components): one lane adds the segment's double2 values at positions a .. a+len-1 of plane
// `pair` (plane stride FT_LD = 3 mod 8 sixteen-byte bank groups).  All lanes of a quarter-warp read element j of their item
// in the same LDS.128, so the quarter-warp is conflict-free iff its eight items START in eight different bank groups,
// g = (a + 3 pair) mod 8.  The items of a tile (segments longest first) are therefore dealt into eight queues by g;
// quarter-warp k takes the k-th item of every queue (lane = g; an exhausted queue leaves an empty lane).  The k-th items
// are about equally long, so the lanes of a warp still run equally long loops.  A tile whose longest queue would not fit
// the shared-memory table falls back to the dense numbering (eight consecutive items per quarter-warp).
// item = a | len << 8 | pair << 13 | accumulator slot << 15 (a < 256, len <= 16, slot < 128); 0 = empty lane.
BA_D int ft_item_quarters(int s0, int s1, const int* __restrict__ seg_info, bool& dense) {
    unsigned long long cnt = 0;  // eight 8-bit counters
    for (int s = s0; s < s1; ++s) {
        const int a = __ldg(seg_info + s) & 0xffff;
#pragma unroll
        for (int pr = 0; pr < 3; ++pr) cnt += 1ull << (8 * ((a + 3 * pr) & 7));
    }
    int q = 0;
#pragma unroll
    for (int g = 0; g < 8; ++g) q = max(q, (int)((cnt >> (8 * g)) & 255));
    dense = q > FT_ITEM_QMAX || (s1 - s0) > 84;  // (a counter may not overflow: 3 x 84 < 256)
    return dense ? (3 * (s1 - s0) + 7) / 8 : q;
}
__global__ void k_ft_item_count(int n_tiles, const int* __restrict__ ft_seg, const int* __restrict__ seg_info, int* __restrict__ nq) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tiles) return;
    bool dense;
    nq[t] = ft_item_quarters(__ldg(ft_seg + t), __ldg(ft_seg + t + 1), seg_info, dense);
}
// q_off = exclusive sum of nq; `items` is zeroed by the caller.  Also writes the tile headers of the accumulator pass:
// the second int4 = (first item, item slots, first track row, longest track).
__global__ void k_ft_item_fill(int n_tiles, const int* __restrict__ ft_seg, const int* __restrict__ seg_info,
                               const int* __restrict__ seg_slot, const int* __restrict__ q_off, const int4* __restrict__ hdr,
                               int* __restrict__ items, int4* __restrict__ hdr_acc) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tiles) return;
    const int s0 = __ldg(ft_seg + t), s1 = __ldg(ft_seg + t + 1);
    bool dense;
    const int nq = ft_item_quarters(s0, s1, seg_info, dense);
    int* out = items + 8 * (size_t)q_off[t];
    unsigned long long cnt = 0;
    int n = 0;
    for (int s = s0; s < s1; ++s) {
        const int info = __ldg(seg_info + s), a = info & 0xffff, len = info >> 16;
        const int base = a | (len << 8) | (__ldg(seg_slot + s) << 15);
#pragma unroll
        for (int pr = 0; pr < 3; ++pr) {
            const int g = (a + 3 * pr) & 7;
            const int k = (int)((cnt >> (8 * g)) & 255);
            cnt += 1ull << (8 * g);
            out[dense ? n++ : 8 * k + g] = base | (pr << 13);
        }
    }
    const int4 h0 = hdr[2 * t], h1 = hdr[2 * t + 1];
    hdr_acc[2 * t] = h0;
    hdr_acc[2 * t + 1] = make_int4(8 * q_off[t], 8 * nq, h1.z, h1.w);
}

__global__ void k_iota(int n, int* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = i;
}

}  // namespace ba
