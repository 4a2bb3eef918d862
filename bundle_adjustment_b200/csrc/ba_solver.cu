// ba_solver.cu - host side of libba_b200.so: device buffers, problem upload + device-side
// permutation build, the Levenberg-Marquardt launch sequence, NCCL plumbing and the C ABI
// declared in include/ba_b200.h.
//
// The host never touches problem data during a solve: it enqueues kernels and reads back one small
// status struct per LM iteration (and one int every few PCG iterations).  There is no CPU fallback:
// every entry point fails with BA_ERR_CUDA when no usable device is present.
#include "../../include/ba_b200.h"
#include "ba_kernels.cuh"
#include "ba_tiles.cuh"
#include "ba_peer.cuh"
#include "ba_dense.cuh"
#include "ba_host_math.h"
#include "ba_triangulate.cuh"
#include "ba_map.cuh"

#include <cub/cub.cuh>
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <functional>
#include <vector>

namespace ba {

// ---------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_last_error;

struct BaError : std::runtime_error {
    int code;
    BaError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

[[noreturn]] static void fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    throw BaError(code, buf);
}

#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) fail(BA_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// ---------------------------------------------------------------------------------------------
// NCCL, resolved at run time so that single-GPU use needs no NCCL at all
// ---------------------------------------------------------------------------------------------
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclApi& nccl_api() {
    static NcclApi api;
    if (api.lib) return api;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.lib) break;
    }
    if (!api.lib) fail(BA_ERR_NCCL, "cannot load libnccl.so.2: %s", dlerror());
    auto sym = [&](const char* s) {
        void* p = dlsym(api.lib, s);
        if (!p) fail(BA_ERR_NCCL, "libnccl lacks %s", s);
        return p;
    };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    return api;
}

#define NK(call)                                                                                   \
    do {                                                                                           \
        ncclResult_t r_ = (call);                                                                  \
        if (r_ != ncclSuccess) fail(BA_ERR_NCCL, "%s failed: %s", #call, nccl_api().GetErrorString(r_)); \
    } while (0)

// ---------------------------------------------------------------------------------------------
// device buffer
// ---------------------------------------------------------------------------------------------
template <class T>
struct DBuf {
    T* p = nullptr;
    size_t n = 0;
    DBuf() = default;
    DBuf(const DBuf&) = delete;
    DBuf& operator=(const DBuf&) = delete;
    ~DBuf() { release(); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
    void alloc(size_t count) {
        if (count <= n && p) return;
        release();
        if (count) CK(cudaMalloc(&p, count * sizeof(T)));
        n = count;
    }
    void zero(cudaStream_t s) {
        if (p) CK(cudaMemsetAsync(p, 0, n * sizeof(T), s));
    }
};

// Scratch buffer of one call: comes from the device's stream-ordered memory pool (kept warm: the pool's release
// threshold is raised in create_common), so per-call scratch costs microseconds instead of a cudaMalloc /
// cudaFree pair (each a device-wide synchronisation).  Used on ONE stream, the one current when it is allocated.
static thread_local cudaStream_t g_scratch_stream = nullptr;
struct ScratchScope {
    cudaStream_t prev;
    explicit ScratchScope(cudaStream_t s) : prev(g_scratch_stream) { g_scratch_stream = s; }
    ~ScratchScope() { g_scratch_stream = prev; }
};
template <class T>
struct TBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaStream_t s = nullptr;
    TBuf() = default;
    TBuf(const TBuf&) = delete;
    TBuf& operator=(const TBuf&) = delete;
    ~TBuf() { release(); }
    void release() {
        if (p) cudaFreeAsync(p, s);
        p = nullptr;
        n = 0;
    }
    void alloc(size_t count) {
        if (count <= n && p) return;
        release();
        s = g_scratch_stream;
        if (count) CK(cudaMallocAsync(&p, count * sizeof(T), s));
        n = count;
    }
    void zero(cudaStream_t st) {
        if (p) CK(cudaMemsetAsync(p, 0, n * sizeof(T), st));
    }
};

// ---------------------------------------------------------------------------------------------
// kernel kinds (accounting)
// ---------------------------------------------------------------------------------------------
enum Kind {
    KN_SETUP = 0, KN_CAM_PREP, KN_RESID, KN_JAC, KN_LIN_PTS, KN_PT_PREP, KN_PT_PASS, KN_BACKSUB, KN_LIN_CAM,
    KN_SCHUR_CAM, KN_MATVEC_CAM, KN_CAM_SUM, KN_CAM_FIN, KN_PCG_DOTS, KN_PCG_UPDATE, KN_CAM_UPDATE, KN_REDUCE,
    KN_GRAD, KN_DECIDE, KN_TILE_MV, KN_TILE_BACKSUB, KN_ROWS_FIN, KN_ROWS_XCHG, KN_DOTS_PEER, KN_PCG_TAIL, KN_ALLREDUCE, KN_DENSE, KN_COUNT
};
static const char* kKindNames[KN_COUNT] = {
    "setup", "cam_prepare", "residuals", "jacobian_blocks", "lin_points", "point_prepare", "schur_matvec_points",
    "backsub_points", "lin_cameras", "schur_rhs_cameras", "schur_matvec_cameras", "camera_sum", "camera_finish",
    "pcg_dots", "pcg_update", "camera_update", "reduce_scalars", "gradient_check", "lm_decide",
    "schur_matvec_tiles", "backsub_tiles", "rows_finish", "rows_finish_xchg", "pcg_dots_peer", "pcg_tail", "peer_allreduce", "dense_schur"};

// ---------------------------------------------------------------------------------------------
// setup kernels (device-side construction of the two observation orders)
// ---------------------------------------------------------------------------------------------
// Locality ordering of the points (internal ids): a point's key is the centroid of the cameras that
// observe it (the fixed keyframe is the oldest, so it counts as camera 0 and adjustable camera c as
// c + 1).  Points seen by the same cameras become neighbours, so a point-major tile touches few
// distinct cameras and a camera's observations touch a narrow range of points.  Integer atomics only
// (order-independent, deterministic).
__global__ void k_point_cam_sums(int n_obs, int n_cams, int n_pts, const int* __restrict__ obs_cam,
                                 const int* __restrict__ obs_pt, unsigned long long* __restrict__ sum,
                                 unsigned* __restrict__ cnt, int* __restrict__ err) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    const int c = obs_cam[i], p = obs_pt[i];
    if (c < -1 || c >= n_cams || p < 0 || p >= n_pts) { atomicOr(err, 1); return; }
    atomicAdd(sum + p, (unsigned long long)(c + 1));
    atomicAdd(cnt + p, 1u);
}
__global__ void k_point_keys(int n_pts, const unsigned long long* __restrict__ sum, const unsigned* __restrict__ cnt,
                             unsigned* __restrict__ key, int* __restrict__ iota) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pts) return;
    const unsigned n = cnt[p];
    key[p] = n ? (unsigned)((sum[p] * 8ull) / n) : 0xffffffffu;  // unobserved points last
    iota[p] = p;
}
// track length of every point in the (first) locality order, plus a trailing zero for the scan
__global__ void k_gather_counts(int n_pts, const int* __restrict__ perm, const unsigned* __restrict__ cnt, int* __restrict__ out) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q > n_pts) return;
    out[q] = q < n_pts ? (int)cnt[perm[q]] : 0;
}
// second-level key: (tile of the point, longest track first): inside a tile the points are ordered by
// track length so that the lanes of a warp that sum tracks side by side run equally long loops
__global__ void k_point_tile_keys(int n_pts, int n_tiles, const int* __restrict__ tile_pt, const int* __restrict__ len,
                                  unsigned* __restrict__ key) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_pts) return;
    int lo = 0, hi = n_tiles;  // last tile with tile_pt[t] <= q
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (tile_pt[mid] <= q) lo = mid; else hi = mid;
    }
    key[q] = ((unsigned)lo << 10) | (unsigned)(1023 - min(len[q], 1023));
}
__global__ void k_invert_perm(int n, const int* __restrict__ perm, int* __restrict__ inv) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q < n) inv[perm[q]] = q;
}

__global__ void k_setup_keys(int n_obs, int n_cams, int n_pts, const int* __restrict__ obs_cam,
                             const int* __restrict__ obs_pt, const int* __restrict__ pt_inv, int2* __restrict__ o_idx,
                             unsigned* __restrict__ key_pt, unsigned long long* __restrict__ key_cam,
                             int* __restrict__ iota, int* __restrict__ err) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    const int c = obs_cam[i];
    int p = obs_pt[i];
    if (c < -1 || c >= n_cams || p < 0 || p >= n_pts) {
        atomicOr(err, 1);
        o_idx[i] = make_int2(n_cams, 0);
        key_pt[i] = 0; key_cam[i] = ((unsigned long long)n_cams << 32); iota[i] = i;
        return;
    }
    p = pt_inv[p];  // internal (locality-ordered) point id
    const int slot = c < 0 ? n_cams : c;
    o_idx[i] = make_int2(slot, p);
    key_pt[i] = (unsigned)p;
    key_cam[i] = ((unsigned long long)slot << 32) | (unsigned)p;
    iota[i] = i;
}

__global__ void k_gather_pm(int n, const int* __restrict__ perm, const double2* __restrict__ o_uv,
                            const int2* __restrict__ o_idx, double2* __restrict__ pm_uv, int2* __restrict__ pm_idx) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int j = perm[i];
    pm_uv[i] = o_uv[j];
    pm_idx[i] = o_idx[j];
}

__global__ void k_gather_cm(int n, const int* __restrict__ perm, const double2* __restrict__ o_uv,
                            const int2* __restrict__ o_idx, double2* __restrict__ cm_uv, int* __restrict__ cm_pt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int j = perm[i];
    cm_uv[i] = o_uv[j];
    cm_pt[i] = o_idx[j].y;
}

// out[q] = first index i in sorted keys[0..n) with keys[i] >= q * scale_shifted  (lower bound)
__global__ void k_lower_bound_u32(int n, const unsigned* __restrict__ keys, int nq, int* __restrict__ out) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (keys[mid] < (unsigned)q) lo = mid + 1; else hi = mid;
    }
    out[q] = lo;
}
__global__ void k_lower_bound_cam(int n, const unsigned long long* __restrict__ keys, int nq, int* __restrict__ out) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const unsigned long long target = (unsigned long long)q << 32;
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (keys[mid] < target) lo = mid + 1; else hi = mid;
    }
    out[q] = lo;
}
// pm_tile_pt[t] = first point whose first observation index is >= t * PM_THREADS (clamped to n_pts)
__global__ void k_pm_tiles(int n_pts, const int* __restrict__ pt_off, int n_tiles, int* __restrict__ tile_pt) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > n_tiles) return;
    if (t == n_tiles) { tile_pt[t] = n_pts; return; }
    const int target = t * PM_THREADS;
    int lo = 0, hi = n_pts;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (pt_off[mid] < target) lo = mid + 1; else hi = mid;
    }
    tile_pt[t] = lo;
}
__global__ void k_max_track(int n_pts, const int* __restrict__ pt_off, int* __restrict__ out) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pts) return;
    atomicMax(out, pt_off[p + 1] - pt_off[p]);
}
// compact [n*3] <-> slot `buf` (4 doubles) of the 128-byte point records
// `perm[q]` = caller's index of internal point q; `base` = one parameter buffer (planes (X0,X1) and X2)
__global__ void k_pack_points(int n, int n_pad, const int* __restrict__ perm, const double* __restrict__ in3, double* __restrict__ base) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const size_t p = (size_t)perm[q];
    reinterpret_cast<double2*>(base)[q] = make_double2(in3[3 * p], in3[3 * p + 1]);
    base[2 * (size_t)n_pad + q] = in3[3 * p + 2];
}
__global__ void k_unpack_points(int n, int n_pad, const int* __restrict__ perm, const double* __restrict__ base, double* __restrict__ out3) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const size_t p = (size_t)perm[q];
    const double2 a = reinterpret_cast<const double2*>(base)[q];
    out3[3 * p] = a.x;
    out3[3 * p + 1] = a.y;
    out3[3 * p + 2] = base[2 * (size_t)n_pad + q];
}
__global__ void k_pack_cams(int n_cams, const double* __restrict__ rvec, const double* __restrict__ tvec,
                            double fr0, double fr1, double fr2, double ft0, double ft1, double ft2, double* __restrict__ rt) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c > n_cams) return;
    double* o = rt + (size_t)c * 6;
    if (c == n_cams) { o[0] = fr0; o[1] = fr1; o[2] = fr2; o[3] = ft0; o[4] = ft1; o[5] = ft2; return; }
    for (int k = 0; k < 3; ++k) { o[k] = rvec[3 * (size_t)c + k]; o[3 + k] = tvec[3 * (size_t)c + k]; }
}
__global__ void k_unpack_cams(int n_cams, const double* __restrict__ rt, double* __restrict__ rvec, double* __restrict__ tvec) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_cams) return;
    for (int k = 0; k < 3; ++k) { rvec[3 * (size_t)c + k] = rt[6 * (size_t)c + k]; tvec[3 * (size_t)c + k] = rt[6 * (size_t)c + 3 + k]; }
}
// CSR pattern of the reference's _prepare_sparsity_matrix (src/bundle_adjuster.py:74-120)
__global__ void k_sparsity_counts(int n_obs, int n_cams, const int2* __restrict__ o_idx, long long* __restrict__ cnt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    cnt[i] = (o_idx[i].x < n_cams) ? 18 : 6;
}
__global__ void k_sparsity_fill(int n_obs, int n_cams, const int2* __restrict__ o_idx, const int* __restrict__ pt_perm,
                                const long long* __restrict__ start, long long* __restrict__ indptr,
                                long long* __restrict__ indices) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    const int2 id = o_idx[i];
    const long long s = start[i];
    const bool adj = id.x < n_cams;
    const int per = adj ? 9 : 3;
    indptr[2 * (size_t)i] = s;
    indptr[2 * (size_t)i + 1] = s + per;
    if (i == n_obs - 1) indptr[2 * (size_t)n_obs] = s + 2 * per;
    if (!indices) return;
    for (int row = 0; row < 2; ++row) {
        long long* o = indices + s + row * per;
        int q = 0;
        if (adj) {
            for (int k = 0; k < 3; ++k) o[q++] = 3LL * id.x + k;
            for (int k = 0; k < 3; ++k) o[q++] = 3LL * n_cams + 3LL * id.x + k;
        }
        const long long p = pt_perm[id.y];  // the caller's point index
        for (int k = 0; k < 3; ++k) o[q++] = 6LL * n_cams + 3LL * p + k;
    }
}
__global__ void k_init_state(LMState* st, double lambda) {
    LMState z;
    memset(&z, 0, sizeof z);
    z.lambda = lambda;
    z.nu = 2.0;
    z.cur = st->cur;  // the parameter buffer selector survives across solves
    z.t0 = global_ns();
    *st = z;
}

}  // namespace ba

// ---------------------------------------------------------------------------------------------
// the handle
// ---------------------------------------------------------------------------------------------
using namespace ba;

struct ba_handle {
    int device = 0;
    int rank = 0, nranks = 1;
    cudaStream_t stream = nullptr;
    ncclComm_t comm = nullptr;
    bool have_problem = false;
    Params P{};
    int max_track = 0;
    size_t smem_lin = 0, smem_pass = 0;

    DBuf<double> cam_rt, cam_R, pt_X, pt_z, pt_h, pt_V, pt_g, pt_Vinv, pt_D;
    DBuf<double> cam_U, cam_g, cam_D, cam_Minv, cam_acc, v_b, v_x, v_r, v_u, v_p, v_s, v_w, dot_partial;
    DBuf<double> init_cam_rt, init_pts;  // parameters of the last ba_set_problem (for ba_restore_initial)
    DBuf<int> ticket;
    DBuf<double2> o_uv, pm_uv, cm_uv;
    DBuf<int2> o_idx, pm_idx;
    DBuf<int> pt_off, pm_tile_pt, cm_pt, cam_tile_off;
    DBuf<int> pt_perm;  // internal (locality-ordered) point id -> caller's point index
    // fused point tiles
    bool fused_ok = false;
    // epoch accumulators of the mat-vec tile pass (ba_tiles.cuh); acc_ok = every epoch fits FT_ACC_CAP slots
    bool acc_ok = false;
    bool tile_lin = false;  // point-side linearisation on the tiles (k_tile_lin) instead of the point-major kernel
    bool cm_warp = false;  // camera-major passes: one warp per tile instead of one CTA per tile
    int tail_cluster = 0;  // CTAs of the PCG tail cluster (ba_peer.cuh); 0 = tail kernel not usable for this problem
    int red_cluster = 1;   // CTAs of the scalar-reduction cluster
    int tail_cached = 0;   // last answer of the (slow) cluster occupancy query and the shared-memory size it holds for
    size_t tail_cached_bytes = 0;
    int ep_target = 0, n_epochs = 0, n_slots = 0, max_ep_slots = 0;
    DBuf<int> seg_slot, slot_row, ep_g0, cta_ep0, cam_row_off2;
    DBuf<int> seg_item;      // packed phase-D work items of the accumulator pass (k_ft_item_fill), 8 per quarter-warp
    DBuf<int4> ft_hdr_acc;   // tile headers of the accumulator pass: (first item, item slots) instead of (first segment, segments)
    long long n_item_slots = 0;
    int ft_tiles = 0, n_seg = 0, ft_grid = 0, sm_count = 0;
    std::vector<int> ft_pt_h, ft_obs_h, ft_maxlen_h;  // host copy of the tile cuts and longest tracks (empty: no tiles)
    DBuf<int> ft_pt, ft_obs, ft_seg, ft_rs, seg_cam, seg_info, seg_row, cam_row_off;
    DBuf<double2> tc_uv, tc_J;
    DBuf<int2> tc_meta;
    DBuf<int4> ft_hdr;
    DBuf<int> pm_tc;
    DBuf<double> rows;
    // peer-memory exchange of the PCG mat-vec partials (ba_peer.cuh); multi-GPU handles only
    bool peer_ok = false;
    // two-level preconditioner (BA_PRECOND_TWO_LEVEL): coarse matrix of the current solve
    DBuf<double> coarse_Ainv, coarse_A, coarse_A2, coarse_probe, coarse_w;
    DBuf<unsigned char> coarse_adj;
    DBuf<int> coarse_colour, coarse_nb, coarse_flag;
    DBuf<unsigned long long> coarse_acc;  // direct build: integer sums of the point part of the coarse matrix (k_coarse_direct)
    DBuf<double> coarse_rscale;
    int coarse_agg = 32, coarse_nz = 0;
    int coarse_colours = 0, coarse_graph_agg = 0;  // colouring of the aggregate graph of the uploaded problem (0: not built)
    bool coarse_ready = false;
    // dense reduced system of small problems (ba_dense.cuh)
    bool dense_ok = false;
    int dn_pairs_n = 0, dn_split = 1;
    DBuf<int4> dn_pairs;
    DBuf<int> dn_off, dn_count;   // dn_count: [0] successful factorisations of the last solve, [1] ticket
    DBuf<double> dn_S, dn_extra;
    DBuf<double2> tc_r;
    int xcap = 0;
    int ar_cap = 0;        // capacity (doubles) of the LL all-reduce region behind the tail's slots; 0 = NCCL only
    size_t ar_base = 0;    // its offset in xbuf (doubles)
    int xpush = 2;  // exchange by remote loads (0, pull), remote stores + flags (1, push), remote stores of LL words (2)
    DBuf<double> xbuf;
    DBuf<int> xstate;
    DBuf<double*> d_peers;
    std::vector<double*> peer_ptr;  // exchange buffer of every rank as mapped here (own rank: xbuf.p)
    DBuf<int4> cm_tile;
    DBuf<double> cm_partial, pm_partial, cam_partial;
    DBuf<LMState> st;
    DBuf<TraceRec> trace;  // per-LM-iteration records of the last solve (ba_solve_trace)
    int trace_n = 0;
    DBuf<double> io;  // staging for parameter upload / download
    // device-resident map (ba_map.cuh): observation rows of every keyframe seen so far, map point positions by id
    DBuf<int> ms_kf, ms_mp;
    DBuf<double2> ms_uv;
    size_t ms_rows = 0;
    DBuf<double> mp_xyz;
    DBuf<unsigned char> mp_known;
    DBuf<int> kf_slot;
    DBuf<int> mw_ids;        // map point ids of the last window (sorted), = the caller's point indices of the uploaded problem
    int mw_n_pts = 0;
    bool mw_current = false;  // the uploaded problem is the one ba_map_window built
    long long map_counters[8] = {0};  // see ba_map_stats
    LMState* h_state = nullptr;  // pinned
    int* h_flag = nullptr;       // pinned

    long long launches[KN_COUNT] = {0};
    double kernel_ms[KN_COUNT] = {0};
    bool time_kernels = false;
    std::vector<cudaEvent_t> ev_pool;           // pooled events for per-launch timing
    struct Stamp { int kind; int e0, e1; };
    std::vector<Stamp> stamps;
    cudaEvent_t ev_solve0 = nullptr, ev_solve1 = nullptr, ev_user0 = nullptr, ev_user1 = nullptr;

    ~ba_handle() {
        for (size_t r = 0; r < peer_ptr.size(); ++r)
            if (peer_ptr[r] && (int)r != rank) cudaIpcCloseMemHandle(peer_ptr[r]);
        if (comm) nccl_api().CommDestroy(comm);
        if (h_state) cudaFreeHost(h_state);
        if (h_flag) cudaFreeHost(h_flag);
        for (cudaEvent_t e : ev_pool) cudaEventDestroy(e);
        if (ev_user0) cudaEventDestroy(ev_user0);
        if (ev_user1) cudaEventDestroy(ev_user1);
        if (ev_solve0) cudaEventDestroy(ev_solve0);
        if (ev_solve1) cudaEventDestroy(ev_solve1);
        if (stream) cudaStreamDestroy(stream);
    }
};

namespace ba {

template <class... KArgs, class... Args>
static void launch(ba_handle* h, Kind kind, void (*kernel)(KArgs...), int grid, int block, size_t smem, Args... args) {
    if (grid <= 0) return;
    int e0 = -1, e1 = -1;
    if (h->time_kernels) {
        e0 = (int)h->stamps.size() * 2;
        e1 = e0 + 1;
        while ((int)h->ev_pool.size() <= e1) {
            cudaEvent_t e;
            CK(cudaEventCreate(&e));
            h->ev_pool.push_back(e);
        }
        CK(cudaEventRecord(h->ev_pool[e0], h->stream));
    }
    kernel<<<grid, block, smem, h->stream>>>(args...);
    CK(cudaGetLastError());
    h->launches[kind] += 1;
    if (h->time_kernels) {
        CK(cudaEventRecord(h->ev_pool[e1], h->stream));
        h->stamps.push_back({(int)kind, e0, e1});
    }
}

// one thread-block cluster of `cluster` CTAs
template <class... KArgs, class... Args>
static void launch_cluster(ba_handle* h, Kind kind, void (*kernel)(KArgs...), int cluster, int block, size_t smem, Args... args) {
    int e0 = -1, e1 = -1;
    if (h->time_kernels) {
        e0 = (int)h->stamps.size() * 2;
        e1 = e0 + 1;
        while ((int)h->ev_pool.size() <= e1) {
            cudaEvent_t e;
            CK(cudaEventCreate(&e));
            h->ev_pool.push_back(e);
        }
        CK(cudaEventRecord(h->ev_pool[e0], h->stream));
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(cluster); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = h->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    CK(cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...));
    h->launches[kind] += 1;
    if (h->time_kernels) {
        CK(cudaEventRecord(h->ev_pool[e1], h->stream));
        h->stamps.push_back({(int)kind, e0, e1});
    }
}

// resolve the pooled per-launch events into per-kind milliseconds (after the stream has drained)
static void resolve_stamps(ba_handle* h) {
    for (const auto& st : h->stamps) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, h->ev_pool[st.e0], h->ev_pool[st.e1]));
        h->kernel_ms[st.kind] += ms;
    }
    h->stamps.clear();
}

static inline int cdiv(long long a, int b) { return (int)((a + b - 1) / b); }

// In place, on the solver stream.  Camera-sized vectors travel through the LL words of the peer-mapped exchange buffer
// (k_peer_allreduce: rank-order sums, identical bits on every rank); larger ones, and handles without peer access, use NCCL.
static void allreduce(ba_handle* h, double* buf, size_t count, bool is_max) {
    if (h->nranks <= 1 || count == 0) return;
    if (h->peer_ok && h->ar_cap > 0 && count <= (size_t)h->ar_cap) {
        const int grid = (int)std::min<size_t>((count + AR_THREADS - 1) / AR_THREADS, (size_t)4 * h->sm_count);
        launch(h, KN_ALLREDUCE, k_peer_allreduce, grid, AR_THREADS, 0, buf, (int)count, is_max ? 1 : 0, h->xbuf.p,
               (double* const*)h->d_peers.p, h->nranks, h->rank, h->ar_base, h->ar_cap, h->xstate.p);
        return;
    }
    NK(nccl_api().AllReduce(buf, buf, count, ncclDouble, is_max ? ncclMax : ncclSum, h->comm, h->stream));
}
static void allreduce_sum(ba_handle* h, double* buf, size_t count) { allreduce(h, buf, count, false); }
static void allreduce_max(ba_handle* h, double* buf, size_t count) { allreduce(h, buf, count, true); }

static void create_common(ba_handle* h, int device) {
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        fail(BA_ERR_CUDA, "no CUDA device available (%s); this library has no CPU path",
             e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    if (device < 0 || device >= ndev) fail(BA_ERR_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    h->device = device;
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        fail(BA_ERR_CUDA, "device %d is sm_%d%d; libba_b200 is built for sm_100a only", device, prop.major, prop.minor);
    h->sm_count = prop.multiProcessorCount;
    {
        cudaMemPool_t pool;
        CK(cudaDeviceGetDefaultMemPool(&pool, device));
        unsigned long long keep = ~0ULL;  // freed scratch stays in the pool for the next call
        CK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    }
    CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CK(cudaMallocHost(&h->h_state, sizeof(LMState)));
    CK(cudaMallocHost(&h->h_flag, sizeof(int)));
    CK(cudaEventCreate(&h->ev_user0));
    CK(cudaEventCreate(&h->ev_user1));
    CK(cudaEventCreate(&h->ev_solve0));
    CK(cudaEventCreate(&h->ev_solve1));
    h->st.alloc(1);
    h->st.zero(h->stream);
    h->trace.alloc(TRACE_CAP);
    h->trace.zero(h->stream);
    h->ticket.alloc(1);
    h->ticket.zero(h->stream);
    h->dot_partial.alloc(2 * MAX_VEC_CTAS);
    h->dot_partial.zero(h->stream);
}

static void use(ba_handle* h) {
    if (!h) fail(BA_ERR_ARG, "null handle");
    CK(cudaSetDevice(h->device));
}

// Stage timing of ba_set_problem on stderr when the environment variable BA_B200_TRACE is set (each mark
// synchronises the stream, so it is a diagnostic, not a mode to benchmark in).
struct StageTimer {
    bool on;
    cudaStream_t s;
    std::chrono::steady_clock::time_point t0;
    explicit StageTimer(cudaStream_t st) : on(getenv("BA_B200_TRACE") != nullptr), s(st), t0(std::chrono::steady_clock::now()) {}
    void mark(const char* what) {
        if (!on) return;
        cudaStreamSynchronize(s);
        const auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[ba_b200] set_problem %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

struct MaxOp {
    __host__ __device__ int operator()(int a, int b) const { return a > b ? a : b; }
};

// Build the fused point tiles (ba_tiles.cuh) from the point-major order: greedy cut of the point
// sequence on the host (one pass over the track offsets), everything per-observation on the device.
static void build_tiles(ba_handle* h, int n_cams, int n_pts, int n_obs) {
    cudaStream_t s = h->stream;
    ScratchScope scratch(s);
    const int B = 256;
    h->fused_ok = false;
    h->ft_tiles = 0; h->n_seg = 0;
    if (h->ft_pt_h.empty()) return;  // a track does not fit one tile: the two-pass kernels serve this problem
    const std::vector<int>&tp = h->ft_pt_h, &to = h->ft_obs_h;
    const int T = (int)tp.size() - 1;
    h->ft_pt.alloc(tp.size()); h->ft_obs.alloc(to.size()); h->ft_seg.alloc(tp.size());
    CK(cudaMemcpyAsync(h->ft_pt.p, tp.data(), tp.size() * sizeof(int), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(h->ft_obs.p, to.data(), to.size() * sizeof(int), cudaMemcpyHostToDevice, s));

    // tile-camera order: stable sort of the point-major indices by (tile, camera)
    TBuf<unsigned long long> key, key_s;
    TBuf<int> ia, ib, ic, id;
    key.alloc(n_obs); key_s.alloc(n_obs); ia.alloc(n_obs); ib.alloc(n_obs); ic.alloc(n_obs); id.alloc(n_obs);
    launch(h, KN_SETUP, k_ft_keys, cdiv(n_obs, B), B, 0, n_obs, T, (const int*)h->ft_obs.p, (const int2*)h->pm_idx.p, key.p, ia.p);
    int tile_bits = 1;
    while ((1LL << tile_bits) < T) ++tile_bits;
    size_t tb_sort = 0, tb_max = 0, tb_sum = 0, tb_seg = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tb_sort, key.p, key_s.p, ia.p, ib.p, n_obs, 0, 32 + tile_bits, s));
    CK(cub::DeviceScan::InclusiveScan(nullptr, tb_max, ia.p, ic.p, MaxOp(), n_obs, s));
    CK(cub::DeviceScan::InclusiveSum(nullptr, tb_sum, ia.p, ic.p, n_obs, s));
    TBuf<unsigned char> tmp;
    tmp.alloc(std::max(std::max(tb_sort, tb_max), tb_sum));
    size_t tb = tmp.n;
    CK(cub::DeviceRadixSort::SortPairs(tmp.p, tb, key.p, key_s.p, ia.p, ib.p, n_obs, 0, 32 + tile_bits, s));  // ib = source index
    launch(h, KN_SETUP, k_ft_run_heads, cdiv(n_obs, B), B, 0, n_obs, (const unsigned long long*)key_s.p, ia.p);
    tb = tmp.n;
    CK(cub::DeviceScan::InclusiveScan(tmp.p, tb, ia.p, id.p, MaxOp(), n_obs, s));                             // id = run start
    launch(h, KN_SETUP, k_ft_seg_flags, cdiv(n_obs, B), B, 0, n_obs, (const int*)id.p, ia.p);               // ia = segment head flag
    tb = tmp.n;
    CK(cub::DeviceScan::InclusiveSum(tmp.p, tb, ia.p, ic.p, n_obs, s));                                      // ic = segment id + 1
    h->launches[KN_SETUP] += 3;
    int S = 0;
    CK(cudaMemcpyAsync(&S, ic.p + n_obs - 1, sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));

    h->tc_uv.alloc(n_obs); h->tc_meta.alloc((size_t)n_obs + 2); h->pm_tc.alloc(n_obs); h->tc_J.alloc(6 * (size_t)n_obs);
    h->ft_hdr.alloc(2 * (size_t)T);
    h->seg_cam.alloc(S); h->seg_info.alloc((size_t)S + 4); h->seg_row.alloc((size_t)S + 4);  // + TMA window slack
    h->cam_row_off.alloc((size_t)n_cams + 2);
    h->rows.alloc((size_t)S * 6);
    // track-row tables of the transposed point-side layout: one per tile, padded to 16 bytes
    std::vector<int> rs0((size_t)T + 1, 0);
    for (int t = 0; t < T; ++t) rs0[t + 1] = rs0[t] + ((h->ft_maxlen_h[t] + 3) & ~3);
    TBuf<int> d_rs0, d_maxlen, seg_start, seg_cam0, seg_split0, seg_tile, seg_split;
    d_rs0.alloc((size_t)T + 1); d_maxlen.alloc(T); seg_start.alloc((size_t)S + 1); seg_cam0.alloc(S);
    seg_split0.alloc(S); seg_tile.alloc(S); seg_split.alloc(S);
    h->ft_rs.alloc((size_t)rs0[T] + 4);
    h->ft_rs.zero(s);
    CK(cudaMemcpyAsync(d_rs0.p, rs0.data(), rs0.size() * sizeof(int), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_maxlen.p, h->ft_maxlen_h.data(), (size_t)T * sizeof(int), cudaMemcpyHostToDevice, s));
    launch(h, KN_SETUP, k_ft_rowstart, T, FT_THREADS, 0, (const int*)h->ft_pt.p, (const int*)h->pt_off.p, (const int*)d_rs0.p,
           (const int*)d_maxlen.p, h->ft_rs.p);
    launch(h, KN_SETUP, k_ft_tile_seg, cdiv(T + 1, B), B, 0, T, n_obs, S, (const int*)h->ft_obs.p, (const int*)ic.p, h->ft_seg.p,
           seg_start.p);
    launch(h, KN_SETUP, k_ft_fill, cdiv(n_obs, B), B, 0, n_obs, (const unsigned long long*)key_s.p, (const int*)ib.p,
           (const int*)ia.p, (const int*)ic.p, (const int*)h->ft_pt.p, (const int*)h->pt_off.p, (const int*)d_rs0.p,
           (const int*)h->ft_rs.p, (const double2*)h->pm_uv.p, (const int2*)h->pm_idx.p, h->tc_uv.p, h->tc_meta.p, h->pm_tc.p,
           seg_cam0.p, seg_start.p, (const int*)id.p, seg_split0.p);
    if (!getenv("BA_B200_NO_SLOT_BALANCE"))  // (A/B hook) which observation takes which slot of its point's track: by bank load
        launch(h, KN_SETUP, k_ft_slot_balance, T, 32, 0, (const int*)h->ft_pt.p, (const int*)h->ft_obs.p, (const int*)h->pt_off.p,
               (const int*)d_rs0.p, (const int*)h->ft_rs.p, h->tc_meta.p);
    launch(h, KN_SETUP, k_ft_headers, cdiv(T, B), B, 0, T, (const int*)h->ft_pt.p, (const int*)h->ft_obs.p, (const int*)h->ft_seg.p,
           (const int*)d_rs0.p, (const int*)d_maxlen.p, h->ft_hdr.p);
    // segments longest first inside each tile, then rows camera-major (stable sort of the segments by camera)
    TBuf<unsigned> skey, skey_s;
    TBuf<int> sidx, sidx_s;
    skey.alloc(S); skey_s.alloc(S); sidx.alloc(S); sidx_s.alloc(S);
    int cam_bits = 1;
    while ((1LL << cam_bits) < (long long)n_cams + 1) ++cam_bits;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tb_seg, skey.p, skey_s.p, sidx.p, sidx_s.p, S, 0, 32, s));
    if (tb_seg > tmp.n) { CK(cudaStreamSynchronize(s)); tmp.alloc(tb_seg); }
    launch(h, KN_SETUP, k_ft_seg_keys, cdiv(S, B), B, 0, S, T, (const int*)seg_start.p, (const int*)h->ft_obs.p, skey.p, sidx.p);
    tb = tmp.n;
    CK(cub::DeviceRadixSort::SortPairs(tmp.p, tb, skey.p, skey_s.p, sidx.p, sidx_s.p, S, 0, std::min(32, 5 + tile_bits), s));
    launch(h, KN_SETUP, k_ft_seg_apply, cdiv(S, B), B, 0, S, (const unsigned*)skey_s.p, (const int*)sidx_s.p, (const int*)seg_start.p,
           (const int*)seg_cam0.p, (const int*)h->ft_obs.p, h->seg_info.p, h->seg_cam.p, skey.p, sidx.p, (const int*)seg_split0.p,
           seg_tile.p, seg_split.p);
    tb = tmp.n;
    CK(cub::DeviceRadixSort::SortPairs(tmp.p, tb, skey.p, skey_s.p, sidx.p, sidx_s.p, S, 0, cam_bits, s));
    h->launches[KN_SETUP] += 2;
    launch(h, KN_SETUP, k_ft_seg_rows, cdiv(S, B), B, 0, S, (const int*)sidx_s.p, h->seg_row.p);
    launch(h, KN_SETUP, k_lower_bound_u32, cdiv(n_cams + 2, B), B, 0, S, (const unsigned*)skey_s.p, n_cams + 2, h->cam_row_off.p);
    CK(cudaFuncSetAttribute(k_tile_pass<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ft_smem_bytes(FT_ROWS_MV)));
    CK(cudaFuncSetAttribute(k_tile_pass<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ft_smem_bytes(FT_ROWS_MV)));
    CK(cudaFuncSetAttribute(k_tile_pass<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ft_smem_bytes(FT_ROWS_MV, true)));
    h->ft_tiles = T; h->n_seg = S;
    h->ft_grid = std::min(T, FT_CTAS_PER_SM * h->sm_count);  // persistent CTAs
    if (const char* g = getenv("BA_B200_TILE_GRID"))           // test hook: few CTAs, so that small problems have long tile runs
        h->ft_grid = std::max(1, std::min(h->ft_grid, atoi(g)));

    // ---- epoch accumulators: slots = distinct (epoch, camera, split) of the segments; rows = slots, camera-major ----
    h->acc_ok = false; h->n_slots = 0; h->n_epochs = 0; h->max_ep_slots = 0; h->ep_target = 0;
    if (S > 0 && !getenv("BA_B200_NO_EPOCHS")) {
        const int G = h->ft_grid;
        TBuf<unsigned long long> k64, k64_s;
        TBuf<int> sord, sord_s, sinc, d_ep_ts;
        TBuf<unsigned> scam, scam_s;
        k64.alloc(S); k64_s.alloc(S); sord.alloc(S); sord_s.alloc(S); sinc.alloc(S); scam.alloc(S); scam_s.alloc(S);
        h->seg_slot.alloc((size_t)S + 4); h->slot_row.alloc(S); h->cam_row_off2.alloc((size_t)n_cams + 2); h->cta_ep0.alloc(G);
        size_t tb64 = 0, tbs = 0;
        CK(cub::DeviceRadixSort::SortPairs(nullptr, tb64, k64.p, k64_s.p, sord.p, sord_s.p, S, 0, 40 + tile_bits, s));
        CK(cub::DeviceScan::InclusiveSum(nullptr, tbs, sord.p, sinc.p, S, s));
        if (std::max(tb64, tbs) > tmp.n) { CK(cudaStreamSynchronize(s)); tmp.alloc(std::max(tb64, tbs)); }
        for (int target = 32; target >= 1 && !h->acc_ok; target /= 2) {
            std::vector<int> ep_ts, cta_ep0(G);
            for (int b = 0; b < G; ++b) {
                const int t0 = ft_cta_begin(b, T, G), t1 = ft_cta_begin(b + 1, T, G);
                cta_ep0[b] = (int)ep_ts.size();
                if (t1 <= t0) continue;
                const int el = ft_epoch_len(t1 - t0, target);
                for (int t = t0; t < t1; t += el) ep_ts.push_back(t);
            }
            const int E = (int)ep_ts.size();
            d_ep_ts.alloc(E); h->ep_g0.alloc((size_t)E + 3); h->ep_g0.zero(s);
            CK(cudaMemcpyAsync(d_ep_ts.p, ep_ts.data(), (size_t)E * sizeof(int), cudaMemcpyHostToDevice, s));
            CK(cudaMemcpyAsync(h->cta_ep0.p, cta_ep0.data(), (size_t)G * sizeof(int), cudaMemcpyHostToDevice, s));
            launch(h, KN_SETUP, k_ft_slot_keys, cdiv(S, B), B, 0, S, T, G, target, (const int*)seg_tile.p, (const int*)h->seg_cam.p,
                   (const int*)seg_split.p, k64.p, sord.p);
            tb = tmp.n;
            CK(cub::DeviceRadixSort::SortPairs(tmp.p, tb, k64.p, k64_s.p, sord.p, sord_s.p, S, 0, 40 + tile_bits, s));
            launch(h, KN_SETUP, k_ft_slot_heads, cdiv(S, B), B, 0, S, (const unsigned long long*)k64_s.p, sord.p);
            tb = tmp.n;
            CK(cub::DeviceScan::InclusiveSum(tmp.p, tb, sord.p, sinc.p, S, s));
            int R = 0;
            CK(cudaMemcpyAsync(&R, sinc.p + S - 1, sizeof(int), cudaMemcpyDeviceToHost, s));
            CK(cudaStreamSynchronize(s));
            launch(h, KN_SETUP, k_ft_epoch_base, cdiv(E + 1, B), B, 0, E, S, R, (const int*)d_ep_ts.p, (const unsigned long long*)k64_s.p,
                   (const int*)sinc.p, h->ep_g0.p);
            std::vector<int> g0((size_t)E + 1);
            CK(cudaMemcpyAsync(g0.data(), h->ep_g0.p, g0.size() * sizeof(int), cudaMemcpyDeviceToHost, s));
            CK(cudaStreamSynchronize(s));
            int mx = 0;
            for (int e = 0; e < E; ++e) mx = std::max(mx, g0[e + 1] - g0[e]);
            h->max_ep_slots = mx; h->n_slots = R; h->n_epochs = E; h->ep_target = target;
            if (getenv("BA_B200_TRACE"))
                fprintf(stderr, "[ba_b200] epochs of %d tiles: %d epochs, %d slots, at most %d per epoch (capacity %d)\n", target, E, R, mx, FT_ACC_CAP);
            if (mx > FT_ACC_CAP) continue;
            launch(h, KN_SETUP, k_ft_slot_assign, cdiv(S, B), B, 0, S, E, (const unsigned long long*)k64_s.p, (const int*)sord_s.p,
                   (const int*)sinc.p, (const int*)d_ep_ts.p, (const int*)h->ep_g0.p, h->seg_slot.p, scam.p, sord.p);
            // rows of the slots, camera-major: stable sort of the slot numbers by camera
            tb = tmp.n;
            CK(cub::DeviceRadixSort::SortPairs(tmp.p, tb, scam.p, scam_s.p, sord.p, sinc.p, R, 0, cam_bits, s));
            launch(h, KN_SETUP, k_ft_seg_rows, cdiv(R, B), B, 0, R, (const int*)sinc.p, h->slot_row.p);
            launch(h, KN_SETUP, k_lower_bound_u32, cdiv(n_cams + 2, B), B, 0, R, (const unsigned*)scam_s.p, n_cams + 2, h->cam_row_off2.p);
            h->acc_ok = true;
        }
        if (h->acc_ok) {
            // phase-D work items, dealt so that the lanes of a quarter-warp start in different shared-memory bank groups
            TBuf<int> nq, q_off;
            nq.alloc((size_t)T + 1); q_off.alloc((size_t)T + 1);
            nq.zero(s);
            size_t tbq = 0;
            CK(cub::DeviceScan::ExclusiveSum(nullptr, tbq, nq.p, q_off.p, T + 1, s));
            if (tbq > tmp.n) { CK(cudaStreamSynchronize(s)); tmp.alloc(tbq); }
            launch(h, KN_SETUP, k_ft_item_count, cdiv(T, B), B, 0, T, (const int*)h->ft_seg.p, (const int*)h->seg_info.p, nq.p);
            tb = tmp.n;
            CK(cub::DeviceScan::ExclusiveSum(tmp.p, tb, nq.p, q_off.p, T + 1, s));
            h->launches[KN_SETUP] += 1;
            int Q = 0;
            CK(cudaMemcpyAsync(&Q, q_off.p + T, sizeof(int), cudaMemcpyDeviceToHost, s));
            CK(cudaStreamSynchronize(s));
            h->n_item_slots = 8LL * Q;
            h->seg_item.alloc((size_t)8 * Q + 8);  // + TMA window slack
            h->seg_item.zero(s);
            h->ft_hdr_acc.alloc(2 * (size_t)T);
            launch(h, KN_SETUP, k_ft_item_fill, cdiv(T, B), B, 0, T, (const int*)h->ft_seg.p, (const int*)h->seg_info.p,
                   (const int*)h->seg_slot.p, (const int*)q_off.p, (const int4*)h->ft_hdr.p, h->seg_item.p, h->ft_hdr_acc.p);
        }
    }
    CK(cudaStreamSynchronize(s));  // scratch dies here
    h->fused_ok = true;
    // PCG tail: the largest cluster (16, 8, ... CTAs of 1024 threads) this device can co-schedule with the chunk in shared memory
    h->tail_cluster = 0;
    if (n_cams > 0 && !getenv("BA_B200_NO_TAIL")) {
        if (h->tail_cached > 0 && pt_smem_bytes(n_cams, h->tail_cached) <= h->tail_cached_bytes) h->tail_cluster = h->tail_cached;
        for (int cl = PT_MAX_CLUSTER; cl >= 1 && !h->tail_cluster; cl /= 2) {
            const size_t bytes = (pt_smem_bytes(n_cams, cl) + 16383) & ~(size_t)16383;  // a size class, so the answer can be reused
            if (bytes > 208 * 1024) break;
            bool good = true;
            for (int peer = 0; peer < 2 && good; ++peer) {
                const void* fn = peer ? (const void*)k_pcg_tail<true> : (const void*)k_pcg_tail<false>;
                good = cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess &&
                       cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max(bytes, (size_t)1024)) == cudaSuccess;
                if (!good) break;
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3(cl); cfg.blockDim = dim3(PT_THREADS); cfg.dynamicSmemBytes = bytes;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                cfg.attrs = at; cfg.numAttrs = 1;
                int n = 0;
                good = cudaOccupancyMaxActiveClusters(&n, fn, &cfg) == cudaSuccess && n >= 1;
            }
            cudaGetLastError();
            if (good) { h->tail_cluster = cl; h->tail_cached = cl; h->tail_cached_bytes = bytes; }
        }
    }
    h->red_cluster = std::max(1, h->tail_cluster);  // a 1024-thread kernel with less shared memory: schedulable wherever the tail is
    if (h->red_cluster > 1 && cudaFuncSetAttribute(k_reduce_scalars, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
        cudaGetLastError();
        h->red_cluster = std::min(h->red_cluster, 8);
    }
    if (getenv("BA_B200_TRACE")) {
        fprintf(stderr, "[ba_b200] PCG tail cluster: %d CTAs\n", h->tail_cluster);
        int o0 = 0, o1 = 0, o2 = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o0, k_tile_pass<0, false>, FT_THREADS, ft_smem_bytes(FT_ROWS_MV));
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o1, k_tile_pass<0, true>, FT_THREADS, ft_smem_bytes(FT_ROWS_MV, true));
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o2, k_tile_pass<1, false>, FT_THREADS, ft_smem_bytes(FT_ROWS_MV));
        fprintf(stderr, "[ba_b200] tile pass CTAs/SM: mat-vec %d, mat-vec with accumulators %d, back-substitution %d; grid %d, %d tiles, "
                "%d segments, %d slots in %d epochs (max %d per epoch)\n", o0, o1, o2, h->ft_grid, T, S, h->n_slots, h->n_epochs, h->max_ep_slots);
    }
}

// Parameter block of the mat-vec tile pass and its row finish: per-segment rows, or epoch accumulator slots
static Params matvec_params(const ba_handle* h, bool acc) {
    Params Q = h->P;
    if (acc) {
        Q.seg_row = h->seg_slot.p; Q.cam_row_off = h->cam_row_off2.p; Q.ep_target = h->ep_target;
        Q.seg_info = h->seg_item.p; Q.ft_hdr = h->ft_hdr_acc.p;  // packed work items instead of the segment table
        Q.cta_ep0 = h->cta_ep0.p; Q.ep_g0 = h->ep_g0.p; Q.slot_row = h->slot_row.p;
    }
    return Q;
}
static void launch_tile_matvec(ba_handle* h, const Params& Q, bool acc, const double* vec) {
    if (acc) launch(h, KN_TILE_MV, k_tile_pass<0, true>, h->ft_grid, FT_THREADS, ft_smem_bytes(FT_ROWS_MV, true), Q, vec);
    else launch(h, KN_TILE_MV, k_tile_pass<0, false>, h->ft_grid, FT_THREADS, ft_smem_bytes(FT_ROWS_MV), Q, vec);
}

// Unmap the peers' exchange buffers and wait until every rank has done the same: freeing memory that a peer still has
// mapped through cudaIpc is undefined behaviour.  Collective.
static void close_peer_maps(ba_handle* h) {
    bool any = false;
    for (size_t r = 0; r < h->peer_ptr.size(); ++r)
        if (h->peer_ptr[r] && (int)r != h->rank) { cudaIpcCloseMemHandle(h->peer_ptr[r]); any = true; }
    h->peer_ptr.assign(h->nranks, nullptr);
    if (!any || !h->comm || !h->xstate.p) return;
    // xstate[3] is a spare word; the all-reduce is only a barrier across the ranks
    NK(nccl_api().AllReduce(h->xstate.p + 3, h->xstate.p + 3, 1, ncclInt, ncclSum, h->comm, h->stream));
    CK(cudaStreamSynchronize(h->stream));
}

// Map every rank's exchange buffer into this process (collective: all ranks call it with the same n_cams).
// Falls back to the NCCL all-reduce (peer_ok = false on EVERY rank) if any rank cannot map a peer.
static void setup_peer_exchange(ba_handle* h, int n_cams) {
    if (h->nranks <= 1) return;
    const int need = ((std::max(n_cams, 1) * 6 + 15) / 16) * 16;
    if (h->xcap >= need) return;  // same decision on every rank: cameras are replicated
    cudaStream_t s = h->stream;
    ScratchScope scratch(s);
    CK(cudaStreamSynchronize(s));
    close_peer_maps(h);
    h->peer_ok = false;
    h->xbuf.release();
    // behind the tail's slots: the LL all-reduce region, [2][nranks][ar_cap] pairs of words
    const int ar_need = ((std::max(n_cams, 1) * 27 + 15) / 16) * 16;
    h->ar_cap = (getenv("BA_B200_NCCL_ALLREDUCE") || ar_need > AR_MAX_DOUBLES) ? 16 : ar_need;
    h->ar_base = (xb_doubles(h->nranks, need) + 1) & ~(size_t)1;
    h->xbuf.alloc(h->ar_base + 4 * (size_t)h->nranks * h->ar_cap);
    h->xbuf.zero(s);
    h->xstate.alloc(8);
    h->xstate.zero(s);
    h->xcap = need;
    h->d_peers.alloc(h->nranks);
    int good = (getenv("BA_B200_NO_PEER") || h->nranks > XB_MAX_RANKS) ? 0 : 1;
    {
        const char* e = getenv("BA_B200_XCHG");  // "pull": peers read this rank's buffer; default: this rank stores into theirs
        h->xpush = (e && strcmp(e, "pull") == 0) ? 0 : (e && strcmp(e, "push") == 0) ? 1 : 2;  // 2 = push, LL words in the tail
    }
    struct Card { cudaIpcMemHandle_t mem; char uuid[16]; } mine;
    memset(&mine, 0, sizeof mine);
    if (good && cudaIpcGetMemHandle(&mine.mem, h->xbuf.p) != cudaSuccess) { cudaGetLastError(); good = 0; }
    {
        int dev = 0;
        cudaDeviceProp prop;
        CK(cudaGetDevice(&dev));
        CK(cudaGetDeviceProperties(&prop, dev));
        memcpy(mine.uuid, prop.uuid.bytes, 16);
    }
    TBuf<unsigned char> d_h;
    d_h.alloc((size_t)h->nranks * sizeof mine);
    CK(cudaMemcpyAsync(d_h.p + (size_t)h->rank * sizeof mine, &mine, sizeof mine, cudaMemcpyHostToDevice, s));
    NK(nccl_api().AllGather(d_h.p + (size_t)h->rank * sizeof mine, d_h.p, sizeof mine, ncclChar, h->comm, s));
    std::vector<Card> all(h->nranks);
    CK(cudaMemcpyAsync(all.data(), d_h.p, all.size() * sizeof mine, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    // ranks that share a GPU must not wait for each other inside kernels: NCCL path for them
    for (int a = 0; a < h->nranks; ++a)
        for (int b = a + 1; b < h->nranks; ++b)
            if (memcmp(all[a].uuid, all[b].uuid, 16) == 0) good = 0;
    for (int r = 0; r < h->nranks && good; ++r) {
        if (r == h->rank) { h->peer_ptr[r] = h->xbuf.p; continue; }
        void* p = nullptr;
        if (cudaIpcOpenMemHandle(&p, all[r].mem, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); good = 0; break; }
        h->peer_ptr[r] = static_cast<double*>(p);
    }
    // consensus: everybody or nobody
    TBuf<double> d_flag;
    d_flag.alloc(1);
    const double mine_ok = good ? 1.0 : 0.0;
    double all_ok = 0.0;
    CK(cudaMemcpyAsync(d_flag.p, &mine_ok, sizeof(double), cudaMemcpyHostToDevice, s));
    NK(nccl_api().AllReduce(d_flag.p, d_flag.p, 1, ncclDouble, ncclMin, h->comm, s));
    CK(cudaMemcpyAsync(&all_ok, d_flag.p, sizeof(double), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    h->peer_ok = all_ok > 0.5;
    if (h->peer_ok) CK(cudaMemcpyAsync(h->d_peers.p, h->peer_ptr.data(), h->nranks * sizeof(double*), cudaMemcpyHostToDevice, s));
    CK(cudaStreamSynchronize(s));
}

// Pair lists of the dense reduced system (ba_dense.cuh): only for single-GPU problems with at most DN_MAX_CAMS adjustable cameras
// and a tile layout (the Jacobian cache is indexed in tile-camera order).
static void build_dense_pairs(ba_handle* h) {
    const Params& P = h->P;
    h->dense_ok = false;
    h->dn_pairs_n = 0;
    h->P.tc_r = nullptr;
    if (h->nranks != 1 || !h->fused_ok || !h->tile_lin || P.n_cams < 1 || P.n_cams > DN_MAX_CAMS || P.n_obs > DN_MAX_OBS || getenv("BA_B200_NO_DENSE")) return;
    cudaStream_t s = h->stream;
    ScratchScope scratch(s);
    const int B = 256, n = P.n_obs, nb = P.n_cams * P.n_cams;
    TBuf<int> cnt, off;
    cnt.alloc((size_t)n + 1); off.alloc((size_t)n + 1);
    launch(h, KN_SETUP, k_dn_pair_count, cdiv(n + 1, B), B, 0, n, P.n_cams, P.pm_idx, P.pt_off, cnt.p);
    size_t tb = 0, tb2 = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt.p, off.p, n + 1, s));
    TBuf<unsigned char> tmp;
    tmp.alloc(tb);
    CK(cub::DeviceScan::ExclusiveSum(tmp.p, tb, cnt.p, off.p, n + 1, s));
    int total = 0;
    CK(cudaMemcpyAsync(&total, off.p + n, sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    if (total <= 0 || total > DN_MAX_PAIRS) return;
    TBuf<unsigned> key, key_s;
    TBuf<int> idx, idx_s;
    TBuf<int4> raw;
    key.alloc(total); key_s.alloc(total); idx.alloc(total); idx_s.alloc(total); raw.alloc(total);
    launch(h, KN_SETUP, k_dn_pair_fill, cdiv(n, B), B, 0, n, P.n_cams, P.pm_idx, P.pt_off, P.pm_tc, (const int*)off.p, key.p, idx.p, raw.p);
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tb2, key.p, key_s.p, idx.p, idx_s.p, total, 0, 8, s));
    TBuf<unsigned char> tmp2;
    tmp2.alloc(tb2);
    CK(cub::DeviceRadixSort::SortPairs(tmp2.p, tb2, key.p, key_s.p, idx.p, idx_s.p, total, 0, 8, s));  // stable: point order inside a bin
    // CTAs per bin: fill the device (a window has 16 bins), but no part shorter than one trip of a CTA
    h->dn_split = std::max(1, std::min(std::min(DN_MAX_SPLIT, h->sm_count / nb), cdiv(cdiv(total, nb), DN_THREADS)));
    h->dn_pairs.alloc(total); h->dn_off.alloc((size_t)nb + 1); h->dn_S.alloc((size_t)nb * h->dn_split * 36); h->dn_count.alloc(2);
    h->dn_extra.alloc((size_t)P.n_cams * h->dn_split * DN_EXTRA); h->tc_r.alloc(2 * (size_t)n);
    h->dn_count.zero(s);
    launch(h, KN_SETUP, k_dn_pair_gather, cdiv(total, B), B, 0, total, (const int*)idx_s.p, (const int4*)raw.p, h->dn_pairs.p);
    launch(h, KN_SETUP, k_lower_bound_u32, cdiv(nb + 1, B), B, 0, total, (const unsigned*)key_s.p, nb + 1, h->dn_off.p);
    if (cudaFuncSetAttribute(k_dense_schur, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dn_solve_smem(P.n_cams)) != cudaSuccess) {
        cudaGetLastError();
        CK(cudaStreamSynchronize(s));
        return;
    }
    CK(cudaStreamSynchronize(s));  // scratch dies here
    h->dn_pairs_n = total;
    h->dense_ok = true;
    h->P.tc_r = h->tc_r.p;  // from now on the tile linearisation also stores the weighted residuals
}

// Distributed handles: a validation error on ONE rank must fail the call on ALL ranks, or the others block forever in the next
// collective.  Every rank contributes its error code (0 = fine) to one all-reduce (max); a rank that is fine itself reports
// that another rank failed.  Single-GPU handles: fails at once.
static void agree_or_fail(ba_handle* h, int code, const char* fmt, ...) {
    char buf[512] = "";
    if (code) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
    }
    int any = code;
    if (h->nranks > 1 && h->comm) {
        TBuf<int> d;
        d.alloc(1);
        const int mine = code ? 1 : 0;
        CK(cudaMemcpyAsync(d.p, &mine, sizeof(int), cudaMemcpyHostToDevice, h->stream));
        NK(nccl_api().AllReduce(d.p, d.p, 1, ncclInt, ncclMax, h->comm, h->stream));
        int all = 0;
        CK(cudaMemcpyAsync(&all, d.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        if (all && !code) fail(BA_ERR_ARG, "ba_set_problem failed on another rank (this rank's arguments were accepted)");
        any = all ? code : 0;
    }
    if (any) fail(BA_ERR_ARG, "%s", buf);
}

static void set_problem(ba_handle* h, double fx, double fy, double cx, double cy, int n_cams, int n_pts, int n_obs,
                        const double* rvec, const double* tvec, const double* frv, const double* ftv, const double* pts,
                        const int* obs_cam, const int* obs_pt, const double* obs_uv) {
    use(h);
    cudaStream_t s = h->stream;
    ScratchScope scratch(s);
    {
        const bool bad_n = n_cams < 0 || n_pts <= 0 || n_obs <= 0;
        const bool bad_p = !frv || !ftv || !pts || !obs_cam || !obs_pt || !obs_uv || (n_cams > 0 && (!rvec || !tvec));
        agree_or_fail(h, (bad_n || bad_p) ? 1 : 0, "%s", bad_n ? "need n_adj_cams >= 0, n_pts > 0, n_obs > 0" : "null array argument");
    }
    h->have_problem = false;
    h->mw_current = false;
    h->coarse_colours = 0; h->coarse_graph_agg = 0;  // the aggregate graph belongs to the problem
    const size_t nc1 = (size_t)n_cams + 1;
    StageTimer tm(s);

    h->cam_rt.alloc(2 * nc1 * 6); h->cam_R.alloc(2 * nc1 * 12);
    h->pt_X.alloc(2 * 3 * (size_t)((n_pts + 1) & ~1) + 4); h->pt_z.alloc((size_t)n_pts * 4); h->pt_h.alloc((size_t)n_pts * 4);
    h->pt_V.alloc(2 * (size_t)n_pts * 6); h->pt_g.alloc(2 * (size_t)n_pts * 4);
    h->pt_Vinv.alloc((size_t)n_pts * 6);
    h->pt_D.alloc((size_t)n_pts * 4);
    h->pt_X.zero(s); h->pt_z.zero(s); h->pt_h.zero(s);
    h->init_cam_rt.alloc(nc1 * 6); h->init_pts.alloc((size_t)n_pts * 3);
    if (cdiv(std::max(n_cams, 1), RS_WARPS) > MAX_VEC_CTAS) fail(BA_ERR_ARG, "too many cameras (%d)", n_cams);  // same on every rank
    const size_t ncz = std::max<size_t>(n_cams, 1);
    h->cam_U.alloc(ncz * 21); h->cam_g.alloc(ncz * 6); h->cam_D.alloc(ncz * 6); h->cam_Minv.alloc(ncz * 36);
    h->cam_acc.alloc(ncz * 27); h->cam_partial.alloc(ncz * NSCAL);
    h->v_b.alloc(nc1 * 6); h->v_x.alloc(nc1 * 6); h->v_r.alloc(nc1 * 6); h->v_u.alloc(nc1 * 6);
    h->v_p.alloc(nc1 * 6); h->v_s.alloc(nc1 * 6); h->v_w.alloc(nc1 * 6);
    h->v_b.zero(s); h->v_x.zero(s); h->v_r.zero(s); h->v_u.zero(s); h->v_p.zero(s); h->v_s.zero(s); h->v_w.zero(s);
    h->cam_partial.zero(s); h->cam_g.zero(s); h->cam_D.zero(s); h->cam_U.zero(s);
    h->o_uv.alloc(n_obs); h->o_idx.alloc(n_obs); h->pm_uv.alloc(n_obs); h->pm_idx.alloc(n_obs);
    h->cm_uv.alloc(n_obs); h->cm_pt.alloc(n_obs);
    h->pt_off.alloc((size_t)n_pts + 1 + 4);  // + TMA window slack
    h->cam_tile_off.alloc(nc1);

    // ---- upload -------------------------------------------------------------------------------
    TBuf<int> d_cam, d_pt, d_iota, d_perm, d_err, d_camoff;
    TBuf<unsigned> d_kpt, d_kpt_s;
    TBuf<unsigned long long> d_kcam, d_kcam_s;
    d_cam.alloc(n_obs); d_pt.alloc(n_obs); d_iota.alloc(n_obs); d_perm.alloc(n_obs); d_err.alloc(2);
    d_kpt.alloc(n_obs); d_kpt_s.alloc(n_obs); d_kcam.alloc(n_obs); d_kcam_s.alloc(n_obs); d_camoff.alloc(nc1 + 1);
    d_err.zero(s);
    // cudaMemcpyDefault: the observation arrays are the caller's host arrays (ba_set_problem) or device arrays built from the
    // resident map (ba_map_window)
    CK(cudaMemcpyAsync(d_cam.p, obs_cam, (size_t)n_obs * sizeof(int), cudaMemcpyDefault, s));
    CK(cudaMemcpyAsync(d_pt.p, obs_pt, (size_t)n_obs * sizeof(int), cudaMemcpyDefault, s));
    CK(cudaMemcpyAsync(h->o_uv.p, obs_uv, (size_t)n_obs * sizeof(double2), cudaMemcpyDefault, s));

    tm.mark("allocate + upload");
    Params& P = h->P;
    memset(&P, 0, sizeof P);
    P.K = Intr{fx, fy, cx, cy};
    P.delta = 1.0; P.mode = ROBUST_IRLS;
    P.n_cams = n_cams; P.n_pts = n_pts; P.n_obs = n_obs;
    P.n_pad = (n_pts + 1) & ~1; P.pt_stride = 3LL * P.n_pad;

    const int B = 256;
    int pt_bits = 1;
    while ((1LL << pt_bits) < n_pts) ++pt_bits;
    int cam_bits = 1;
    while ((1LL << cam_bits) < (long long)n_cams + 1) ++cam_bits;
    size_t tmp_bytes = 0, tb2 = 0, tb3 = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_kpt.p, d_kpt_s.p, d_iota.p, d_perm.p, n_obs, 0, pt_bits, s));
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tb2, d_kcam.p, d_kcam_s.p, d_iota.p, d_perm.p, n_obs, 0, 32 + cam_bits, s));
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tb3, d_kpt.p, d_kpt_s.p, d_iota.p, d_perm.p, n_pts, 0, 32, s));
    TBuf<unsigned char> d_tmp;
    d_tmp.alloc(std::max(std::max(tmp_bytes, tb2), tb3));

    // ---- locality order of the points: stable sort by camera centroid ----------------------------
    h->pt_perm.alloc(n_pts);
    {
        TBuf<unsigned long long> d_sum;
        TBuf<unsigned> d_cnt, d_key, d_key_s;
        TBuf<int> d_pi;
        d_sum.alloc(n_pts); d_cnt.alloc(n_pts); d_key.alloc(n_pts); d_key_s.alloc(n_pts); d_pi.alloc(n_pts);
        d_sum.zero(s); d_cnt.zero(s);
        launch(h, KN_SETUP, k_point_cam_sums, cdiv(n_obs, B), B, 0, n_obs, n_cams, n_pts, (const int*)d_cam.p,
               (const int*)d_pt.p, d_sum.p, d_cnt.p, d_err.p);
        launch(h, KN_SETUP, k_point_keys, cdiv(n_pts, B), B, 0, n_pts, (const unsigned long long*)d_sum.p,
               (const unsigned*)d_cnt.p, d_key.p, d_pi.p);
        tmp_bytes = d_tmp.n;
        CK(cub::DeviceRadixSort::SortPairs(d_tmp.p, tmp_bytes, d_key.p, d_key_s.p, d_pi.p, h->pt_perm.p, n_pts, 0, 32, s));
        h->launches[KN_SETUP] += 1;
        // tile cuts: greedy on the host over the track offsets of this order (<= FT_THREADS observations and
        // <= FT_PTS points per tile); then the points of each tile are re-ordered by track length
        {
            TBuf<int> d_len, d_off;
            d_len.alloc((size_t)n_pts + 1); d_off.alloc((size_t)n_pts + 1);
            launch(h, KN_SETUP, k_gather_counts, cdiv(n_pts + 1, B), B, 0, n_pts, (const int*)h->pt_perm.p, (const unsigned*)d_cnt.p, d_len.p);
            size_t tbs = 0;
            CK(cub::DeviceScan::ExclusiveSum(nullptr, tbs, d_len.p, d_off.p, n_pts + 1, s));
            if (tbs > d_tmp.n) { CK(cudaStreamSynchronize(s)); d_tmp.alloc(tbs); }
            tbs = d_tmp.n;
            CK(cub::DeviceScan::ExclusiveSum(d_tmp.p, tbs, d_len.p, d_off.p, n_pts + 1, s));
            std::vector<int> off((size_t)n_pts + 1);
            CK(cudaMemcpyAsync(off.data(), d_off.p, off.size() * sizeof(int), cudaMemcpyDeviceToHost, s));
            CK(cudaStreamSynchronize(s));
            h->ft_pt_h.assign(1, 0); h->ft_obs_h.assign(1, 0); h->ft_maxlen_h.clear();
            bool fits = true;
            for (int p = 0; p < n_pts && fits;) {
                const int o_begin = off[p];
                if (off[p + 1] - o_begin > FT_THREADS) { fits = false; break; }
                int q = p + 1, longest = off[p + 1] - o_begin;
                while (q < n_pts && off[q + 1] - o_begin <= FT_THREADS && q - p < FT_PTS) {
                    longest = std::max(longest, off[q + 1] - off[q]);
                    ++q;
                }
                h->ft_pt_h.push_back(q); h->ft_obs_h.push_back(off[q]); h->ft_maxlen_h.push_back(longest);
                p = q;
            }
            if (!fits) { h->ft_pt_h.clear(); h->ft_obs_h.clear(); h->ft_maxlen_h.clear(); }
            if (fits) {
                const int T = (int)h->ft_pt_h.size() - 1;
                TBuf<int> d_tp, d_perm2;
                d_tp.alloc(h->ft_pt_h.size()); d_perm2.alloc(n_pts);
                CK(cudaMemcpyAsync(d_tp.p, h->ft_pt_h.data(), h->ft_pt_h.size() * sizeof(int), cudaMemcpyHostToDevice, s));
                launch(h, KN_SETUP, k_point_tile_keys, cdiv(n_pts, B), B, 0, n_pts, T, (const int*)d_tp.p, (const int*)d_len.p, d_key.p);
                int bits = 11;
                while ((1LL << (bits - 10)) < T) ++bits;
                tmp_bytes = d_tmp.n;
                CK(cub::DeviceRadixSort::SortPairs(d_tmp.p, tmp_bytes, d_key.p, d_key_s.p, h->pt_perm.p, d_perm2.p, n_pts, 0,
                                                   std::min(bits, 32), s));
                CK(cudaMemcpyAsync(h->pt_perm.p, d_perm2.p, (size_t)n_pts * sizeof(int), cudaMemcpyDeviceToDevice, s));
                CK(cudaStreamSynchronize(s));
                h->launches[KN_SETUP] += 2;
            }
        }
        launch(h, KN_SETUP, k_invert_perm, cdiv(n_pts, B), B, 0, n_pts, (const int*)h->pt_perm.p, d_pi.p);
        // d_pi = caller's index -> internal id
        launch(h, KN_SETUP, k_setup_keys, cdiv(n_obs, B), B, 0, n_obs, n_cams, n_pts, (const int*)d_cam.p, (const int*)d_pt.p,
               (const int*)d_pi.p, h->o_idx.p, d_kpt.p, d_kcam.p, d_iota.p, d_err.p);
        CK(cudaStreamSynchronize(s));  // the scratch buffers of this block die here
    }

    tm.mark("locality order + tile cuts");
    // ---- point-major order: stable radix sort of (internal point, original index) ----------------
    tmp_bytes = d_tmp.n;
    CK(cub::DeviceRadixSort::SortPairs(d_tmp.p, tmp_bytes, d_kpt.p, d_kpt_s.p, d_iota.p, d_perm.p, n_obs, 0, pt_bits, s));
    h->launches[KN_SETUP] += 1;
    launch(h, KN_SETUP, k_gather_pm, cdiv(n_obs, B), B, 0, n_obs, (const int*)d_perm.p, (const double2*)h->o_uv.p,
           (const int2*)h->o_idx.p, h->pm_uv.p, h->pm_idx.p);
    launch(h, KN_SETUP, k_lower_bound_u32, cdiv(n_pts + 1, B), B, 0, n_obs, (const unsigned*)d_kpt_s.p, n_pts + 1, h->pt_off.p);
    const int pm_tiles = cdiv(n_obs, PM_THREADS);
    h->pm_tile_pt.alloc((size_t)pm_tiles + 1);
    launch(h, KN_SETUP, k_pm_tiles, cdiv(pm_tiles + 1, B), B, 0, n_pts, (const int*)h->pt_off.p, pm_tiles, h->pm_tile_pt.p);
    launch(h, KN_SETUP, k_max_track, cdiv(n_pts, B), B, 0, n_pts, (const int*)h->pt_off.p, d_err.p + 1);

    tm.mark("point-major order");
    // ---- camera-major order: stable sort of ((camera, point), original index); fixed camera last ----
    tmp_bytes = d_tmp.n;
    CK(cub::DeviceRadixSort::SortPairs(d_tmp.p, tmp_bytes, d_kcam.p, d_kcam_s.p, d_iota.p, d_perm.p, n_obs, 0, 32 + cam_bits, s));
    h->launches[KN_SETUP] += 1;
    launch(h, KN_SETUP, k_lower_bound_cam, cdiv((int)nc1 + 1, B), B, 0, n_obs, (const unsigned long long*)d_kcam_s.p, (int)nc1 + 1,
           d_camoff.p);
    std::vector<int> camoff(nc1 + 1);
    int err_track[2] = {0, 0};
    CK(cudaMemcpyAsync(camoff.data(), d_camoff.p, (nc1 + 1) * sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(err_track, d_err.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    agree_or_fail(h, err_track[0], "observation index out of range (obs_cam must be in [-1, n_adj_cams), obs_pt in [0, n_pts))");
    const int n_obs_adj = camoff[n_cams];
    launch(h, KN_SETUP, k_gather_cm, cdiv(n_obs_adj, B), B, 0, n_obs_adj, (const int*)d_perm.p, (const double2*)h->o_uv.p,
           (const int2*)h->o_idx.p, h->cm_uv.p, h->cm_pt.p);
    std::vector<int4> tiles;
    std::vector<int> tile_off(nc1, 0);
    // cameras with thousands of observations here: long tiles, one CTA each (few 27-value CTA reductions, about four
    // tiles per SM); otherwise tiles of <= CM_TILE observations, one WARP each (shuffle-tree reductions only)
    h->cm_warp = n_cams > 0 && n_obs_adj / n_cams < 2 * CM_TILE;
    if (const char* e = getenv("BA_B200_CM_WARP")) h->cm_warp = atoi(e) != 0;
    // warp tiles: short enough that the device is full of warps (16 per SM) also on a window-sized problem - 4 cameras x 1500
    // observations as 8 warps of 1024 took 45 us per pass, as 48 warps of 128 a few - but at least four trips of a warp
    const int cm_len = h->cm_warp ? std::min(CM_TILE, std::max(128, cdiv(cdiv(n_obs_adj, 16 * h->sm_count), 32) * 32))
                                  : std::min(8 * CM_TILE, std::max(CM_TILE, cdiv(cdiv(n_obs_adj, 4 * h->sm_count), CM_THREADS) * CM_THREADS));
    for (int c = 0; c < n_cams; ++c) {
        tile_off[c] = (int)tiles.size();
        for (int o = camoff[c]; o < camoff[c + 1]; o += cm_len)
            tiles.push_back(make_int4(c, o, std::min(cm_len, camoff[c + 1] - o), 0));
    }
    tile_off[n_cams] = (int)tiles.size();
    const int cm_tiles = (int)tiles.size();
    h->cm_tile.alloc(std::max<size_t>(cm_tiles, 1));
    if (cm_tiles) CK(cudaMemcpyAsync(h->cm_tile.p, tiles.data(), (size_t)cm_tiles * sizeof(int4), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(h->cam_tile_off.p, tile_off.data(), nc1 * sizeof(int), cudaMemcpyHostToDevice, s));
    h->cm_partial.alloc(std::max<size_t>((size_t)cm_tiles * 27, 27));
    h->max_track = err_track[1];
    tm.mark("camera-major order");
    build_tiles(h, n_cams, n_pts, n_obs);
    if (h->nranks > 1) {
        // every rank must run the same kernels per PCG iteration (they exchange inside them) and add the partial sums
        // in the same order: agree on "tiles usable" and on the size of the tail cluster (minimum over the ranks)
        TBuf<double> d_agree;
        d_agree.alloc(2);
        double mine[2] = {h->fused_ok ? 1.0 : 0.0, (double)h->tail_cluster}, all[2] = {0.0, 0.0};
        CK(cudaMemcpyAsync(d_agree.p, mine, sizeof mine, cudaMemcpyHostToDevice, s));
        NK(nccl_api().AllReduce(d_agree.p, d_agree.p, 2, ncclDouble, ncclMin, h->comm, s));
        CK(cudaMemcpyAsync(all, d_agree.p, sizeof all, cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        if (all[0] < 0.5) h->fused_ok = false;
        if ((int)all[1] < h->tail_cluster) {  // a smaller cluster means larger chunks per CTA
            h->tail_cluster = (int)all[1];
            if (h->tail_cluster > 0) {
                const int bytes = (int)std::max(pt_smem_bytes(n_cams, h->tail_cluster), (size_t)1024);
                if (cudaFuncSetAttribute(k_pcg_tail<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess ||
                    cudaFuncSetAttribute(k_pcg_tail<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess) {
                    cudaGetLastError();
                    fail(BA_ERR_CUDA, "the ranks disagree on the PCG tail cluster and %d CTAs do not fit this device", h->tail_cluster);
                }
            }
            h->tail_cached = 0;
        }
        h->red_cluster = std::max(1, std::min(h->red_cluster, std::max(1, (int)all[1])));
    }
    tm.mark("point tiles");
    const int n_partial_rows = std::max(pm_tiles, h->ft_tiles);
    h->pm_partial.alloc((size_t)n_partial_rows * NSCAL);
    h->pm_partial.zero(s);
    const int cap = (PM_THREADS + h->max_track + 1) & ~1;  // even: keeps the cp.async stage 16-byte aligned
    h->smem_lin = pm_smem_bytes<9, 6>(cap);
    h->smem_pass = pm_smem_bytes<3, 9>(cap);
    const size_t smem_max = 227 * 1024 - 1024;  // static shared memory of the kernels comes on top
    agree_or_fail(h, (h->smem_lin > smem_max || h->smem_pass > smem_max) ? 1 : 0,
                  "a point is observed %d times; this build supports track lengths up to %d", h->max_track,
                  (int)((smem_max - pm_smem_bytes<9, 6>(0)) / 72) - PM_THREADS);
    CK(cudaFuncSetAttribute(k_lin_points, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_lin));
    CK(cudaFuncSetAttribute(k_point_pass<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_pass));
    CK(cudaFuncSetAttribute(k_point_pass<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_pass));

    // ---- parameters -----------------------------------------------------------------------------
    h->io.alloc(std::max<size_t>((size_t)n_pts * 3, nc1 * 6));
    P.n_obs_adj = n_obs_adj; P.pm_tiles = pm_tiles; P.cm_tiles = cm_tiles; P.pm_cap = cap;
    P.vec_ctas = cdiv(std::max(n_cams, 1), VEC_THREADS);
    P.cam_rt = h->cam_rt.p; P.cam_R = h->cam_R.p; P.pt_X = h->pt_X.p; P.pt_z = h->pt_z.p; P.pt_h = h->pt_h.p; P.pt_V = h->pt_V.p; P.pt_g = h->pt_g.p;
    P.pt_Vinv = h->pt_Vinv.p; P.pt_D = h->pt_D.p;
    P.cam_U = h->cam_U.p; P.cam_g = h->cam_g.p; P.cam_D = h->cam_D.p; P.cam_Minv = h->cam_Minv.p; P.cam_acc = h->cam_acc.p;
    P.v_b = h->v_b.p; P.v_x = h->v_x.p; P.v_r = h->v_r.p; P.v_u = h->v_u.p; P.v_p = h->v_p.p; P.v_s = h->v_s.p;
    P.v_w = h->v_w.p; P.dot_partial = h->dot_partial.p;
    P.o_uv = h->o_uv.p; P.o_idx = h->o_idx.p; P.pm_uv = h->pm_uv.p; P.pm_idx = h->pm_idx.p; P.pt_off = h->pt_off.p;
    P.pm_tile_pt = h->pm_tile_pt.p; P.cm_uv = h->cm_uv.p; P.cm_pt = h->cm_pt.p; P.cm_tile = h->cm_tile.p;
    P.cam_tile_off = h->cam_tile_off.p; P.cm_partial = h->cm_partial.p; P.pm_partial = h->pm_partial.p;
    P.cam_partial = h->cam_partial.p; P.st = h->st.p; P.trace = h->trace.p;
    P.ft_tiles = h->ft_tiles; P.n_seg = h->n_seg; P.n_partial_rows = n_partial_rows;
    P.ft_pt = h->ft_pt.p; P.ft_obs = h->ft_obs.p; P.ft_seg = h->ft_seg.p; P.tc_uv = h->tc_uv.p;
    P.ft_hdr = h->ft_hdr.p; P.tc_meta = h->tc_meta.p; P.pm_tc = h->fused_ok ? h->pm_tc.p : nullptr;
    P.tc_J = h->fused_ok ? h->tc_J.p : nullptr;
    h->tile_lin = h->fused_ok && !getenv("BA_B200_NO_TILE_LIN");
    P.seg_cam = h->seg_cam.p; P.seg_info = h->seg_info.p; P.ft_rs = h->ft_rs.p; P.seg_row = h->seg_row.p; P.cam_row_off = h->cam_row_off.p;
    P.rows = h->rows.p;
    P.ep_target = 0; P.cta_ep0 = nullptr; P.ep_g0 = nullptr; P.slot_row = nullptr;
    h->st.zero(s);
    CK(cudaStreamSynchronize(s));  // the sort scratch buffers die at scope exit
    setup_peer_exchange(h, n_cams);
    build_dense_pairs(h);
    tm.mark("finish");
    h->have_problem = true;
}

static void set_params(ba_handle* h, const double* rvec, const double* tvec, const double* frv, const double* ftv,
                       const double* pts, bool save_initial = false) {
    // writes parameter buffer 0 and makes it current
    use(h);
    cudaStream_t s = h->stream;
    const Params& P = h->P;
    const int B = 256;
    const size_t nc1 = (size_t)P.n_cams + 1;
    CK(cudaMemsetAsync(&h->st.p->cur, 0, sizeof(int), s));
    if (pts) {
        CK(cudaMemcpyAsync(h->io.p, pts, (size_t)P.n_pts * 3 * sizeof(double), cudaMemcpyDefault, s));  // host, or device (ba_map_window)
        launch(h, KN_SETUP, k_pack_points, cdiv(P.n_pts, B), B, 0, P.n_pts, P.n_pad, (const int*)h->pt_perm.p, (const double*)h->io.p,
               h->pt_X.p);
        if (save_initial)
            CK(cudaMemcpyAsync(h->init_pts.p, h->io.p, (size_t)P.n_pts * 3 * sizeof(double), cudaMemcpyDeviceToDevice, s));
    }
    if (frv) {
        DBuf<double> tmp;
        tmp.alloc(std::max<size_t>((size_t)P.n_cams * 6, 1));
        if (P.n_cams) {
            CK(cudaMemcpyAsync(tmp.p, rvec, (size_t)P.n_cams * 3 * sizeof(double), cudaMemcpyHostToDevice, s));
            CK(cudaMemcpyAsync(tmp.p + (size_t)P.n_cams * 3, tvec, (size_t)P.n_cams * 3 * sizeof(double), cudaMemcpyHostToDevice, s));
        }
        launch(h, KN_SETUP, k_pack_cams, cdiv((int)nc1, B), B, 0, P.n_cams, (const double*)tmp.p,
               (const double*)(tmp.p + (size_t)P.n_cams * 3), frv[0], frv[1], frv[2], ftv[0], ftv[1], ftv[2], h->cam_rt.p);
        if (save_initial)
            CK(cudaMemcpyAsync(h->init_cam_rt.p, h->cam_rt.p, nc1 * 6 * sizeof(double), cudaMemcpyDeviceToDevice, s));
        CK(cudaStreamSynchronize(s));
    }
    launch(h, KN_CAM_PREP, k_cam_prepare, cdiv((int)nc1, B), B, 0, P, 0);
    CK(cudaStreamSynchronize(s));
}

static void read_state(ba_handle* h) {
    CK(cudaMemcpyAsync(h->h_state, h->st.p, sizeof(LMState), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
}

static void need_problem(ba_handle* h) {
    use(h);
    if (!h->have_problem) fail(BA_ERR_STATE, "no problem uploaded: call ba_set_problem first");
}

static void apply_robust(ba_handle* h, const ba_options* opt) {
    h->P.delta = opt->huber_delta;
    h->P.mode = opt->robust_mode;
    if (!(opt->huber_delta > 0.0)) fail(BA_ERR_ARG, "huber_delta must be positive");
    if (opt->robust_mode < 0 || opt->robust_mode > 2) fail(BA_ERR_ARG, "unknown robust_mode %d", opt->robust_mode);
}

// linearise the point side at buffer `which` and reduce the cost scalars
// `linesearch`: the trial point (which = 3) with the per-point line search fused into the tile kernel
static void lin_points_and_cost(ba_handle* h, int which, bool linesearch = false) {
    const Params& P = h->P;
    if (h->tile_lin && linesearch && which == 3) launch(h, KN_LIN_PTS, k_tile_lin<true>, h->ft_grid, FT_THREADS, 0, P, which);
    else if (h->tile_lin) launch(h, KN_LIN_PTS, k_tile_lin<false>, h->ft_grid, FT_THREADS, 0, P, which);
    else launch(h, KN_LIN_PTS, k_lin_points, P.pm_tiles, PM_THREADS, h->smem_lin, P, which);
}

static const unsigned MASK_LIN = (1u << S_COST) | (1u << S_RAW) | (1u << S_BAD) | (1u << S_GPMAX);
static const unsigned MASK_STEP_PM = (1u << S_GDP) | (1u << S_DDP) | (1u << S_DP2) | (1u << S_XP2);
static const unsigned MASK_STEP_CAM = (1u << S_GDC) | (1u << S_DDC) | (1u << S_RDC) | (1u << S_DC2) | (1u << S_XC2);

template <int KIND>
static void launch_camera_pass(ba_handle* h, Kind kind, const double* vec, int with_schur) {
    const Params& P = h->P;
    if constexpr (KIND == 1) {
        // the Schur-assembly pass holds 27 accumulators, the point record of two trips and a dozen 3x3 temporaries: 192-thread
        // CTAs leave it 170 registers (no spills; at 256 threads / 128 registers it spilled 56-88 bytes and ran 160-180 us on
        // the venice shape against 121 us), gathers of the next trip - X, h and V^-1 - in flight
        constexpr int T1 = 192;
        if (h->cm_warp) launch(h, kind, k_camera_pass<KIND, true, 2, T1>, cdiv(P.cm_tiles, T1 / 32), T1, 0, P, vec, with_schur);
        else launch(h, kind, k_camera_pass<KIND, false, 2, T1>, P.cm_tiles, T1, 0, P, vec, with_schur);
    } else {
        if (h->cm_warp) launch(h, kind, k_camera_pass<KIND, true>, cdiv(P.cm_tiles, CM_THREADS / 32), CM_THREADS, 0, P, vec, with_schur);
        else launch(h, kind, k_camera_pass<KIND, false>, P.cm_tiles, CM_THREADS, 0, P, vec, with_schur);
    }
}

static const double kDiagMin = 1e-6, kDiagMax = 1e32;

// `decide`: single-GPU solves fold the accept / reject decision into the reduction kernel (returns true if it did)
static bool reduce_scalars(ba_handle* h, unsigned pm_mask, unsigned cam_mask, const DecideArgs* decide = nullptr, bool prep = false) {
    const Params& P = h->P;
    const bool fused = decide && h->nranks == 1;
    launch_cluster(h, KN_REDUCE, k_reduce_scalars, std::max(1, h->red_cluster), RED_THREADS, 0, P, pm_mask, cam_mask, fused ? 1 : 0,
                   fused ? *decide : DecideArgs{0.0, 0.0, 0.0, 0.0, 0, 0ull}, (fused && prep) ? 1 : 0, kDiagMin, kDiagMax);
    if (h->nranks > 1) {
        double* scal = reinterpret_cast<double*>(reinterpret_cast<char*>(h->st.p) + offsetof(LMState, scal));
        allreduce_sum(h, scal, N_PT_SCAL);
        if (pm_mask & (1u << S_GPMAX)) allreduce_max(h, scal + S_GPMAX, 1);
    }
    return fused;
}

static void lin_cameras(ba_handle* h, double dmin, double dmax) {
    const Params& P = h->P;
    if (P.n_cams == 0) return;
    launch_camera_pass<0>(h, KN_LIN_CAM, nullptr, 0);
    launch(h, KN_CAM_SUM, k_camera_sum, cdiv((long long)P.n_cams * 27, 256), 256, 0, P, 27, 2);
    allreduce_sum(h, h->cam_acc.p, (size_t)P.n_cams * 27);
    launch(h, KN_CAM_FIN, k_camera_finish_lin, cdiv(P.n_cams, 128), 128, 0, P, dmin, dmax);
}

static const int kDeviceLoopMaxObs = 262144;  // largest problem the device-side LM / PCG loops are the default for

static void solve(ba_handle* h, const ba_options* opt, ba_stats* out) {
    need_problem(h);
    apply_robust(h, opt);
    if (opt->max_iters < 0 || opt->pcg_max_iters < 1) fail(BA_ERR_ARG, "bad iteration limits");
    const Params& P = h->P;
    cudaStream_t s = h->stream;
    for (int k = 0; k < KN_COUNT; ++k) { h->launches[k] = 0; h->kernel_ms[k] = 0.0; }
    h->time_kernels = opt->verbose >= 2;
    const int check_every = std::max(1, opt->pcg_check_every);
    // small problems: explicit reduced system + dense Cholesky instead of PCG (ba_dense.cuh); no preconditioner needed then
    const bool dense = h->dense_ok && opt->dense_schur != 0 && opt->schur_path == BA_PATH_TILES && h->nranks == 1 && h->P.n_cams > 0;
    struct FlagScope {  // the flag rides in the kernel parameters of this solve only (also when it throws)
        int& f;
        FlagScope(int& flag, int v) : f(flag) { f = v; }
        ~FlagScope() { f = 0; }
    } cam_flag(h->P.cam_in_backsub, dense ? 1 : 0);
    const int with_schur = (opt->precond >= BA_PRECOND_SCHUR_JACOBI && !dense) ? 1 : 0;
    if (opt->precond < BA_PRECOND_JACOBI || opt->precond > BA_PRECOND_AUTO) fail(BA_ERR_ARG, "unknown preconditioner %d", opt->precond);
    const int nc1 = P.n_cams + 1;
    if (opt->schur_path < BA_PATH_TILES || opt->schur_path > BA_PATH_TILES_ROWS) fail(BA_ERR_ARG, "unknown schur_path %d", opt->schur_path);
    const bool tiles = opt->schur_path != BA_PATH_TWO_PASS && h->fused_ok;
    const bool acc = tiles && opt->schur_path == BA_PATH_TILES && h->acc_ok;
    const Params PA = matvec_params(h, acc);  // what the mat-vec tile pass and its row finish see
    const bool tail = tiles && h->tail_cluster > 0 && (h->nranks == 1 || h->peer_ok) && opt->pcg_tail != 0;
    // the PCG iterations of a solve are replayed from a CUDA graph (no NCCL call inside, no per-kernel event timing)
    const bool use_graph = (h->nranks == 1 || (h->peer_ok && tiles)) && !h->time_kernels && check_every > 1;
    cudaGraphExec_t pcg_graph[2] = {nullptr, nullptr};  // [0] Schur-Jacobi tail, [1] tail with the coarse correction
    long long graph_launches[2][KN_COUNT] = {{0}, {0}};

    CK(cudaEventRecord(h->ev_solve0, s));
    // the two kernel paths write the per-tile step sums into different rows of pm_partial and k_reduce_scalars adds ALL rows:
    // leftovers of a solve on the other path must not survive into this one
    h->pm_partial.zero(s);
    launch(h, KN_SETUP, k_init_state, 1, 1, 0, h->st.p, opt->initial_lambda);
    launch(h, KN_CAM_PREP, k_cam_prepare, cdiv(nc1, 256), 256, 0, P, 2);
    lin_points_and_cost(h, 2);
    reduce_scalars(h, MASK_LIN, 0);
    launch(h, KN_DECIDE, k_store_cost, 1, 1, 0, P, 2);
    read_state(h);
    if (h->h_state->bad) fail(BA_ERR_NONFINITE, "Residuals are not finite in the initial point.");
    const double cost0 = h->h_state->cost[h->h_state->cur], raw0 = h->h_state->raw[h->h_state->cur];

    bool need_lin = true;
    int status = opt->max_iters == 0 ? 6 : 0;
    unsigned long long pcg_loop = 0, lm_loop = 0;  // conditional handles of the device-side loops (0: host loops)

    // ---- two-level preconditioner (BA_PRECOND_TWO_LEVEL / AUTO): default kernel path (tiles + cluster tail, peer exchange on
    //      N > 1 GPUs), host loops.  The coarse matrix A0 = Z^T S Z is probed with the tile pass once the PCG of an LM iteration
    //      has needed more than 15 iterations (the damping has reached its floor by then and S changes little), inverted on the
    //      device and kept; it is rebuilt when the iteration count climbs past 30 again.
    const bool want_two_level = opt->precond == BA_PRECOND_TWO_LEVEL || (opt->precond == BA_PRECOND_AUTO && P.n_cams >= 128);
    bool two_level = want_two_level && tiles && tail && P.n_cams > 0 && !dense;
    h->coarse_ready = false;
    int coarse_builds = 0, coarse_matvecs = 0, coarse_built_at = -1000;
    // 24: Schur-Jacobi has to be struggling.  On the 50 M-observation problem (10 k cameras) it never needs more than 18 iterations and a
    // coarse matrix, stale after a few LM iterations of a far-from-converged problem, cost more than it saved (profiles/r02j_*)
    const int coarse_after = opt->coarse_after == 0 ? 24 : opt->coarse_after;
    if (two_level) {
        h->coarse_agg = opt->coarse_aggregate > 0 ? opt->coarse_aggregate : std::max(32, cdiv(P.n_cams, 64));
        h->coarse_agg = std::max(h->coarse_agg, cdiv(P.n_cams, 128));  // at most 128 aggregates = 768 coarse unknowns (the inverse is cubic)
        h->coarse_nz = 6 * cdiv(P.n_cams, h->coarse_agg);
        const size_t cb = pt_smem_bytes(P.n_cams, h->tail_cluster, h->coarse_nz, h->coarse_agg);
        const void* fn = h->nranks > 1 ? (const void*)k_pcg_tail<true, true> : (const void*)k_pcg_tail<false, true>;
        if (cb > 200 * 1024 || cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess ||
            cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max(cb, (size_t)1024)) != cudaSuccess) {
            cudaGetLastError();
            two_level = false;  // the coarse tail does not fit this device: plain Schur-Jacobi (reported in ba_stats)
        }
    }
    // Colour the aggregates so that no aggregate is coupled (through a shared point) to two aggregates of one colour: a probe
    // vector can then carry a whole colour.  Local to this rank's points - every rank probes its own part of S.
    auto coarse_colouring = [&]() {
        const int agg = h->coarse_agg, na = h->coarse_nz / 6;
        if (h->coarse_graph_agg == agg && h->coarse_colours > 0) return;
        h->coarse_adj.alloc((size_t)na * na);
        h->coarse_adj.zero(s);
        launch(h, KN_SETUP, k_agg_adjacency, cdiv(P.n_pts, 128), 128, 0, P.n_pts, P.n_cams, agg, na, P.pt_off, P.pm_idx, h->coarse_adj.p);
        std::vector<unsigned char> adj((size_t)na * na);
        CK(cudaMemcpyAsync(adj.data(), h->coarse_adj.p, adj.size(), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        for (int a = 0; a < na; ++a) adj[(size_t)a * na + a] = 1;
        // columns a, b conflict when some row is coupled to both
        std::vector<int> colour(na, -1);
        int ncol = 0;
        std::vector<char> used;
        for (int a = 0; a < na; ++a) {
            used.assign(ncol + 1, 0);
            for (int r = 0; r < na; ++r) {
                if (!adj[(size_t)r * na + a]) continue;
                for (int b = 0; b < a; ++b)
                    if (adj[(size_t)r * na + b]) used[colour[b]] = 1;
            }
            int c = 0;
            while (used[c]) ++c;
            colour[a] = c;
            ncol = std::max(ncol, c + 1);
        }
        // nb[col][b] = the aggregate of colour col that b is coupled to (at most one), -1 if none
        std::vector<int> nb((size_t)ncol * na, -1);
        for (int b = 0; b < na; ++b)
            for (int a = 0; a < na; ++a)
                if (adj[(size_t)b * na + a]) nb[(size_t)colour[a] * na + b] = a;
        h->coarse_colour.alloc(na); h->coarse_nb.alloc(nb.size());
        CK(cudaMemcpyAsync(h->coarse_colour.p, colour.data(), (size_t)na * sizeof(int), cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(h->coarse_nb.p, nb.data(), nb.size() * sizeof(int), cudaMemcpyHostToDevice, s));
        CK(cudaStreamSynchronize(s));
        h->coarse_colours = ncol; h->coarse_graph_agg = agg;
    };
    // direct = the point part of A0 in one pass over the point-major observations (k_coarse_direct); otherwise (a point whose
    // track changes aggregate more than CD_POOL times, or BA_B200_COARSE_PROBE set: the A/B and test hook) by coloured probing
    std::function<bool(bool)> build_coarse_with = [&](bool direct) -> bool {  // between k_camera_finish_schur and the first PCG iteration: V^-1, lambda and D are current
        const int nz = h->coarse_nz, agg = h->coarse_agg, na = nz / 6;
        h->coarse_A.alloc((size_t)nz * nz); h->coarse_A2.alloc((size_t)nz * nz); h->coarse_Ainv.alloc((size_t)nz * nz);
        h->coarse_flag.alloc(2);
        h->coarse_flag.zero(s);
        if (direct) {
            h->coarse_acc.alloc((size_t)nz * nz); h->coarse_rscale.alloc((size_t)nz + 1);
            h->coarse_acc.zero(s);
            launch(h, KN_SETUP, k_coarse_rscale, cdiv(nz, 128), 128, 0, P.n_cams, agg, nz, (const double*)P.cam_U, h->coarse_rscale.p);
            launch(h, KN_SETUP, k_coarse_direct, 5 * h->sm_count, CD_WARPS * 32, 0, P, agg, nz, (const double*)h->coarse_rscale.p,
                   h->coarse_acc.p, h->coarse_flag.p + 1);
            launch(h, KN_SETUP, k_coarse_from_acc, cdiv((long long)nz * nz, 256), 256, 0, nz, (const double*)h->coarse_rscale.p,
                   (const unsigned long long*)h->coarse_acc.p, h->coarse_A.p, (const int*)(h->coarse_flag.p + 1), h->coarse_rscale.p + nz);
            // a track that did not fit the pool on ANY rank sends all ranks to the probing build (same collectives everywhere)
            if (h->nranks > 1) allreduce_max(h, h->coarse_rscale.p + nz, 1);
            double ovf = 0.0;
            CK(cudaMemcpyAsync(&ovf, h->coarse_rscale.p + nz, sizeof(double), cudaMemcpyDeviceToHost, s));
            CK(cudaStreamSynchronize(s));
            if (ovf != 0.0) return build_coarse_with(false);
        } else {
        coarse_colouring();
        h->coarse_probe.alloc((size_t)nc1 * 6); h->coarse_w.alloc((size_t)nc1 * 6);
        h->coarse_A.zero(s);
        Params PQ = PA;  // the row finish leaves this rank's part of S * probe (without the damping term) in a scratch vector
        PQ.cam_acc = h->coarse_w.p;
        for (int col = 0; col < h->coarse_colours; ++col)
            for (int d = 0; d < 6; ++d) {
                launch(h, KN_SETUP, k_coarse_probe, cdiv(nc1 * 6, 256), 256, 0, P.n_cams, agg, (const int*)h->coarse_colour.p, col, d,
                       h->coarse_probe.p);
                launch_tile_matvec(h, PA, acc, h->coarse_probe.p);
                launch(h, KN_ROWS_FIN, k_rows_finish<true>, cdiv(P.n_cams, RS_WARPS), RS_WARPS * 32, 0, PQ);
                launch(h, KN_SETUP, k_coarse_scatter, cdiv(nz, 128), 128, 0, P.n_cams, agg, nz, d,
                       (const int*)(h->coarse_nb.p + (size_t)col * na), (const double*)h->coarse_w.p, h->coarse_A.p);
                ++coarse_matvecs;
            }
        }
        allreduce_sum(h, h->coarse_A.p, (size_t)nz * nz);
        launch(h, KN_SETUP, k_coarse_finish, cdiv((long long)nz * nz, 256), 256, 0, P.n_cams, agg, nz, (const double*)P.cam_D,
               (const LMState*)h->st.p, (const double*)h->coarse_A.p, h->coarse_A2.p, direct ? (const double*)P.cam_U : (const double*)nullptr);
        ++coarse_builds;
        // inverse on the device by the tail's cluster when its rows fit shared memory, else on the host
        int cl = h->tail_cluster;
        // 6 x 6 pivot blocks (one cluster barrier per aggregate); the scalar elimination behind BA_B200_SCALAR_INVERSE (A/B hook)
        const bool blocked = !getenv("BA_B200_SCALAR_INVERSE");
        const void* inv_fn = blocked ? (const void*)k_spd_inverse6 : (const void*)k_spd_inverse;
        const size_t sb = blocked ? spd6_smem_bytes(nz, std::max(cl, 1)) : spd_smem_bytes(nz, std::max(cl, 1));
        bool on_device = cl >= 1 && sb <= 200 * 1024 && !getenv("BA_B200_HOST_INVERSE") &&
                         cudaFuncSetAttribute(inv_fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess &&
                         cudaFuncSetAttribute(inv_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sb) == cudaSuccess;
        cudaGetLastError();
        if (on_device) {
            if (blocked)
                launch_cluster(h, KN_SETUP, k_spd_inverse6, cl, PT_THREADS, sb, nz, (const double*)h->coarse_A2.p, h->coarse_Ainv.p,
                               h->coarse_flag.p);
            else
            launch_cluster(h, KN_SETUP, k_spd_inverse, cl, PT_THREADS, sb, nz, (const double*)h->coarse_A2.p, h->coarse_Ainv.p,
                           h->coarse_flag.p);
            int bad = 0;
            CK(cudaMemcpyAsync(&bad, h->coarse_flag.p, sizeof(int), cudaMemcpyDeviceToHost, s));
            CK(cudaStreamSynchronize(s));
            return bad == 0;
        }
        std::vector<double> A((size_t)nz * nz), Ainv;
        CK(cudaMemcpyAsync(A.data(), h->coarse_A2.p, A.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        if (!spd_inverse_host(A, nz, Ainv)) return false;  // not positive definite (degenerate aggregate): stay with Schur-Jacobi
        CK(cudaMemcpyAsync(h->coarse_Ainv.p, Ainv.data(), Ainv.size() * sizeof(double), cudaMemcpyHostToDevice, s));
        CK(cudaStreamSynchronize(s));
        return true;
    };
    int last_pcg_iters = 0, lm_iter_host = 0, pcg_at_build = 0, pcg_after_build = 0;
    bool coarse_failed = false, use_coarse = false, used_coarse_last = false;
    double lambda_now = opt->initial_lambda, lambda_at_build = 0.0;

    // ---- the three parts of one LM iteration ------------------------------------------------------------------
    auto before_pcg = [&](bool lin) {
        if (lin) {
            lin_cameras(h, kDiagMin, kDiagMax);
            launch(h, KN_GRAD, k_gradient_check, 1, RED_THREADS, 0, P, opt->gtol);
        }
        launch(h, KN_PT_PREP, k_point_prepare, cdiv(P.n_pts, 256), 256, 0, P, kDiagMin, kDiagMax);
        if (P.n_cams > 0) {
            launch_camera_pass<1>(h, KN_SCHUR_CAM, nullptr, with_schur);
            launch(h, KN_CAM_SUM, k_camera_sum, cdiv((long long)P.n_cams * 27, 256), 256, 0, P, 27, 0);
            allreduce_sum(h, h->cam_acc.p, (size_t)P.n_cams * 27);
            launch(h, KN_CAM_FIN, k_camera_finish_schur, P.vec_ctas, VEC_THREADS, 0, P, with_schur, h->ticket.p, pcg_loop);
            // --- two-level preconditioner: when to build, use, rebuild and give up (all from replicated state: every rank decides alike)
            if (two_level && !coarse_failed) {
                if (h->coarse_ready && used_coarse_last && lm_iter_host == coarse_built_at + 1) pcg_after_build = last_pcg_iters;
                const bool first_build = !h->coarse_ready && last_pcg_iters > coarse_after;
                const bool rebuild = h->coarse_ready && used_coarse_last && lm_iter_host - coarse_built_at >= 3 &&
                                     last_pcg_iters > std::max(30, 2 * pcg_after_build);
                if (first_build || rebuild) {
                    h->coarse_ready = build_coarse_with(!getenv("BA_B200_COARSE_PROBE"));
                    coarse_failed = !h->coarse_ready;
                    coarse_built_at = lm_iter_host;
                    pcg_at_build = last_pcg_iters;
                    lambda_at_build = lambda_now;
                }
            }
            // the coarse matrix belongs to the damping it was probed with ((V + lambda D)^-1 inside S): after a rejected step has
            // pushed lambda out of [1/4, 4] x that value the lagged correction does harm - this LM iteration runs Schur-Jacobi
            use_coarse = two_level && h->coarse_ready && !coarse_failed &&
                         lambda_now <= 4.0 * lambda_at_build && lambda_now >= 0.25 * lambda_at_build;
            if (use_coarse)
                launch(h, KN_CAM_FIN, k_coarse_first, 1, PT_THREADS, 2 * (size_t)h->coarse_nz * sizeof(double), P,
                       Coarse{h->coarse_Ainv.p, h->coarse_agg, h->coarse_nz}, pcg_loop);
        }
    };
    auto pcg_iteration = [&]() {
        int n_dots = P.vec_ctas;
        if (tiles && tail) {  // two kernels per iteration: the tile pass, then everything else in one cluster
            launch_tile_matvec(h, PA, acc, h->v_u.p);
            const size_t bytes = pt_smem_bytes(P.n_cams, h->tail_cluster);
            const Coarse none{nullptr, 1, 0};
            if (h->nranks > 1 && use_coarse)
                launch_cluster(h, KN_PCG_TAIL, k_pcg_tail<true, true>, h->tail_cluster, PT_THREADS,
                               pt_smem_bytes(P.n_cams, h->tail_cluster, h->coarse_nz, h->coarse_agg), PA, opt->pcg_tol, opt->pcg_max_iters,
                               h->xbuf.p, (double* const*)h->d_peers.p, h->nranks, h->rank, h->xpush, h->xcap, h->xstate.p, pcg_loop,
                               Coarse{h->coarse_Ainv.p, h->coarse_agg, h->coarse_nz});
            else if (h->nranks > 1)
                launch_cluster(h, KN_PCG_TAIL, k_pcg_tail<true, false>, h->tail_cluster, PT_THREADS, bytes, PA, opt->pcg_tol, opt->pcg_max_iters,
                               h->xbuf.p, (double* const*)h->d_peers.p, h->nranks, h->rank, h->xpush, h->xcap, h->xstate.p, pcg_loop, none);
            else if (use_coarse)
                launch_cluster(h, KN_PCG_TAIL, k_pcg_tail<false, true>, h->tail_cluster, PT_THREADS,
                               pt_smem_bytes(P.n_cams, h->tail_cluster, h->coarse_nz, h->coarse_agg), PA, opt->pcg_tol, opt->pcg_max_iters,
                               (double*)nullptr, (double* const*)nullptr, 1, 0, 0, 0, (int*)nullptr, pcg_loop,
                               Coarse{h->coarse_Ainv.p, h->coarse_agg, h->coarse_nz});
            else
                launch_cluster(h, KN_PCG_TAIL, k_pcg_tail<false, false>, h->tail_cluster, PT_THREADS, bytes, PA, opt->pcg_tol, opt->pcg_max_iters,
                               (double*)nullptr, (double* const*)nullptr, 1, 0, 0, 0, (int*)nullptr, pcg_loop, none);
            return;
        }
        if (tiles) {
            launch_tile_matvec(h, PA, acc, h->v_u.p);
            if (h->nranks > 1 && h->peer_ok) {  // exchange through peer memory, fused into both neighbours
                // the separate kernels keep the pull protocol unless BA_B200_XCHG=push asks for remote stores
                const int xp = h->xpush == 1 ? 1 : 0;
                launch(h, KN_ROWS_XCHG, k_rows_finish_xchg, cdiv(P.n_cams, RS_WARPS), RS_WARPS * 32, 0, PA, h->xbuf.p,
                       (double* const*)h->d_peers.p, h->nranks, h->rank, xp, h->xcap, h->xstate.p);
                launch(h, KN_DOTS_PEER, k_pcg_dots_peer, P.vec_ctas, VEC_THREADS, 0, P, h->xbuf.p, (double* const*)h->d_peers.p,
                       h->nranks, xp, h->xcap, h->xstate.p);
            } else if (h->nranks > 1) {
                launch(h, KN_ROWS_FIN, k_rows_finish<true>, cdiv(P.n_cams, RS_WARPS), RS_WARPS * 32, 0, PA);
                allreduce_sum(h, h->cam_acc.p, (size_t)P.n_cams * 6);
                launch(h, KN_PCG_DOTS, k_pcg_dots<false>, P.vec_ctas, VEC_THREADS, 0, P);
            } else {
                n_dots = cdiv(P.n_cams, RS_WARPS);
                launch(h, KN_ROWS_FIN, k_rows_finish<false>, n_dots, RS_WARPS * 32, 0, PA);
            }
        } else {
            launch(h, KN_PT_PASS, k_point_pass<0>, P.pm_tiles, PM_THREADS, h->smem_pass, P, (const double*)h->v_u.p);
            launch_camera_pass<2>(h, KN_MATVEC_CAM, h->v_u.p, 0);
            if (h->nranks > 1) {
                launch(h, KN_CAM_SUM, k_camera_sum, cdiv((long long)P.n_cams * 6, 256), 256, 0, P, 6, 1);
                allreduce_sum(h, h->cam_acc.p, (size_t)P.n_cams * 6);
                launch(h, KN_PCG_DOTS, k_pcg_dots<false>, P.vec_ctas, VEC_THREADS, 0, P);
            } else {
                launch(h, KN_PCG_DOTS, k_pcg_dots<true>, P.vec_ctas, VEC_THREADS, 0, P);
            }
        }
        launch(h, KN_PCG_UPDATE, k_pcg_update, P.vec_ctas, VEC_THREADS, 0, P, opt->pcg_tol, opt->pcg_max_iters, n_dots, h->ticket.p);
    };
    // the whole camera side of an LM iteration: bins, gradient, S, Cholesky, exact step, trial poses.  The points (V^-1, h) are
    // prepared by the reduction kernel that closed the previous iteration; once here for the first.
    if (dense && status == 0) launch(h, KN_PT_PREP, k_point_prepare, cdiv(P.n_pts, 256), 256, 0, P, kDiagMin, kDiagMax);
    auto dense_solve = [&]() {
        launch(h, KN_DENSE, k_dense_schur, P.n_cams * P.n_cams * h->dn_split, DN_THREADS, dn_solve_smem(P.n_cams), P,
               (const int4*)h->dn_pairs.p, (const int*)h->dn_off.p, h->dn_S.p, h->dn_extra.p, h->dn_count.p + 1, h->dn_count.p, kDiagMin,
               kDiagMax, opt->gtol, h->dn_split);
    };
    if (dense) CK(cudaMemsetAsync(h->dn_count.p, 0, sizeof(int), s));
    auto after_pcg = [&]() {
        if (tiles)  // dense: one more CTA, which moves the cameras
            launch(h, KN_TILE_BACKSUB, k_tile_pass<1>, h->ft_grid + (dense ? 1 : 0), FT_THREADS, ft_smem_bytes(FT_ROWS_MV), P,
                   (const double*)h->v_x.p);
        else
            launch(h, KN_BACKSUB, k_point_pass<1>, P.pm_tiles, PM_THREADS, h->smem_pass, P, (const double*)h->v_x.p);
        if (!dense) launch(h, KN_CAM_UPDATE, k_camera_update, cdiv(nc1, 128), 128, 0, P);
        lin_points_and_cost(h, 3, opt->point_linesearch != 0);
        const DecideArgs da{opt->ftol, opt->xtol, opt->lambda_min, opt->lambda_max, opt->max_iters, lm_loop};
        if (!reduce_scalars(h, MASK_LIN | MASK_STEP_PM, MASK_STEP_CAM, &da, dense))
            launch(h, KN_DECIDE, k_lm_decide, 1, 1, 0, P, opt->ftol, opt->xtol, opt->lambda_min, opt->lambda_max, opt->max_iters, lm_loop);
    };

    // ---- device-side loops: the WHOLE solve is one CUDA graph - a conditional WHILE node over the LM iterations whose body
    //      holds a second WHILE node over the PCG iterations; k_lm_decide / k_camera_finish_schur / k_pcg_tail set the two
    //      conditions (cudaGraphSetConditional).  One graph launch and one state read-back per solve: no host round trip per
    //      LM iteration or per batch of PCG iterations, which is what a window-sized problem spends its time on.
    //      Single GPU, default kernel path; the host loops below serve everything else (and verbose / per-kernel timing).
    //      Measured (profiles/r01j_ab_device_loop.txt): a WHILE trip costs about 2.8 us more than the same two kernels replayed
    //      from a flat graph, so the device loops win only where the host round trips outweigh that - small problems
    //      (window: 13.3 -> 12.5 ms, 100 k observations: 17.0 -> 16.4 ms; 680 k: 84.7 -> 89.5 ms, 5 M: 337 -> 345 ms).
    //      device_loop = 1 therefore means "up to kDeviceLoopMaxObs observations", 2 = always.
    const bool device_loop = status == 0 && (opt->device_loop == 2 || (opt->device_loop == 1 && P.n_obs <= kDeviceLoopMaxObs)) &&
                             h->nranks == 1 && tiles && (tail || dense) && P.n_cams > 0 && !h->time_kernels && !(opt->verbose & 1) &&
                             !two_level && !getenv("BA_B200_HOST_LOOP");
    if (device_loop) {
        long long before[KN_COUNT], per_lm[KN_COUNT], per_pcg[KN_COUNT];
        memcpy(before, h->launches, sizeof before);
        cudaGraph_t g = nullptr;
        cudaGraphExec_t gx = nullptr;
        CK(cudaGraphCreate(&g, 0));
        try {
            cudaGraphConditionalHandle hl = 0, hp = 0;
            CK(cudaGraphConditionalHandleCreate(&hl, g, 1, cudaGraphCondAssignDefault));
            cudaGraphNodeParams np = {};
            np.type = cudaGraphNodeTypeConditional;
            np.conditional.handle = hl; np.conditional.type = cudaGraphCondTypeWhile; np.conditional.size = 1;
            cudaGraphNode_t lm_node = nullptr;
            CK(cudaGraphAddNode(&lm_node, g, nullptr, 0, &np));
            cudaGraph_t lm_body = np.conditional.phGraph_out[0];
            h->P.lin_gate = 1;  // the linearisation kernels are always in the graph and return at once after a rejected step
            if (dense) {
                // no inner loop: linearisation, pair bins, Cholesky, back-substitution and decision are one straight body
                lm_loop = (unsigned long long)hl; pcg_loop = 0;
                CK(cudaStreamBeginCaptureToGraph(s, lm_body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
                dense_solve();
                after_pcg();
                CK(cudaStreamEndCapture(s, nullptr));
                for (int k = 0; k < KN_COUNT; ++k) { per_lm[k] = h->launches[k] - before[k]; per_pcg[k] = 0; }
            } else {
            CK(cudaGraphConditionalHandleCreate(&hp, lm_body, 0, 0));
            lm_loop = (unsigned long long)hl; pcg_loop = (unsigned long long)hp;
            // first part of the body; the PCG loop hangs off its last node
            CK(cudaStreamBeginCaptureToGraph(s, lm_body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
            before_pcg(true);
            cudaStreamCaptureStatus cs;
            const cudaGraphNode_t* deps = nullptr;
            size_t n_deps = 0;
            CK(cudaStreamGetCaptureInfo(s, &cs, nullptr, nullptr, &deps, &n_deps));
            std::vector<cudaGraphNode_t> last(deps, deps + n_deps);
            CK(cudaStreamEndCapture(s, nullptr));
            cudaGraphNodeParams pp = {};
            pp.type = cudaGraphNodeTypeConditional;
            pp.conditional.handle = hp; pp.conditional.type = cudaGraphCondTypeWhile; pp.conditional.size = 1;
            cudaGraphNode_t pcg_node = nullptr;
            CK(cudaGraphAddNode(&pcg_node, lm_body, last.data(), last.size(), &pp));
            cudaGraph_t pcg_body = pp.conditional.phGraph_out[0];
            for (int k = 0; k < KN_COUNT; ++k) per_lm[k] = h->launches[k] - before[k];
            CK(cudaStreamBeginCaptureToGraph(s, pcg_body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
            pcg_iteration();
            CK(cudaStreamEndCapture(s, nullptr));
            for (int k = 0; k < KN_COUNT; ++k) per_pcg[k] = h->launches[k] - before[k] - per_lm[k];
            CK(cudaStreamBeginCaptureToGraph(s, lm_body, &pcg_node, nullptr, 1, cudaStreamCaptureModeThreadLocal));
            after_pcg();
            CK(cudaStreamEndCapture(s, nullptr));
            for (int k = 0; k < KN_COUNT; ++k) per_lm[k] = h->launches[k] - before[k] - per_pcg[k];
            }
            h->P.lin_gate = 0;
            lm_loop = pcg_loop = 0;
            CK(cudaGraphInstantiate(&gx, g, 0));
            CK(cudaGraphLaunch(gx, s));
            read_state(h);
        } catch (...) {
            h->P.lin_gate = 0;
            cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
            if (cudaStreamIsCapturing(s, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone) cudaStreamEndCapture(s, nullptr);
            cudaGetLastError();
            if (gx) cudaGraphExecDestroy(gx);
            cudaGraphDestroy(g);
            throw;
        }
        CK(cudaGraphExecDestroy(gx));
        CK(cudaGraphDestroy(g));
        const LMState& st = *h->h_state;
        status = st.status;
        // a pass that ended on the gradient test ran its kernels without counting as an iteration
        const long long passes = st.iter + (st.status == 1 ? 1 : 0);
        for (int k = 0; k < KN_COUNT; ++k) h->launches[k] = before[k] + per_lm[k] * passes + per_pcg[k] * st.pcg_total;
    }
    while (status == 0) {
        if (dense) {
            dense_solve();
        } else {
        before_pcg(need_lin);
        if (P.n_cams > 0) {
            auto pcg_converged = [&]() {
                CK(cudaMemcpyAsync(h->h_flag, &h->st.p->pcg_done, sizeof(int), cudaMemcpyDeviceToHost, s));
                CK(cudaStreamSynchronize(s));
                return *h->h_flag != 0;
            };
            if (use_graph) {
                // `check_every` PCG iterations = one CUDA graph launch (kernels early-exit once the device flag is set)
                const int gi = use_coarse ? 1 : 0;
                if (!pcg_graph[gi]) {
                    long long before[KN_COUNT];
                    memcpy(before, h->launches, sizeof before);
                    cudaGraph_t g = nullptr;
                    CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
                    for (int k = 0; k < check_every; ++k) pcg_iteration();
                    CK(cudaStreamEndCapture(s, &g));
                    CK(cudaGraphInstantiate(&pcg_graph[gi], g, 0));
                    CK(cudaGraphDestroy(g));
                    for (int k = 0; k < KN_COUNT; ++k) { graph_launches[gi][k] = h->launches[k] - before[k]; h->launches[k] = before[k]; }
                }
                for (int it = 0; it < opt->pcg_max_iters; it += check_every) {
                    CK(cudaGraphLaunch(pcg_graph[gi], s));
                    for (int k = 0; k < KN_COUNT; ++k) h->launches[k] += graph_launches[gi][k];
                    if (it + check_every < opt->pcg_max_iters && pcg_converged()) break;
                }
            } else {
                for (int it = 0; it < opt->pcg_max_iters; ++it) {
                    pcg_iteration();
                    if ((it + 1) % check_every == 0 && it + 1 < opt->pcg_max_iters && pcg_converged()) break;
                }
            }
        }
        }
        after_pcg();
        read_state(h);
        const LMState& st = *h->h_state;
        status = st.status;
        need_lin = st.last_accepted != 0;
        last_pcg_iters = st.pcg_iter;
        lm_iter_host = st.iter;
        lambda_now = st.lambda;
        used_coarse_last = use_coarse;
        if (opt->verbose & 1)
            fprintf(stderr, "[ba_b200] it %3d %s cost %.9e -> trial %.9e  rho %+.3e  lambda %.3e  pcg %d  |g| %.3e  |dx| %.3e\n",
                    st.iter, st.last_accepted ? "acc" : "REJ", st.cost[st.cur ^ (st.last_accepted ? 1 : 0)], st.scal[S_COST], st.rho,
                    st.lambda, st.pcg_iter, st.gmax, st.step_norm);
    }
    if (opt->max_iters == 0) read_state(h);
    for (int gi = 0; gi < 2; ++gi)
        if (pcg_graph[gi]) CK(cudaGraphExecDestroy(pcg_graph[gi]));
    if (h->nranks > 1 && h->peer_ok) {
        int xs[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        CK(cudaMemcpyAsync(xs, h->xstate.p, sizeof xs, cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        if (xs[2]) {
            CK(cudaMemsetAsync(h->xstate.p + 2, 0, sizeof(int), s));
            CK(cudaMemsetAsync(h->xstate.p + 6, 0, sizeof(int), s));
            fail(BA_ERR_NCCL, "peer-memory exchange: rank %d waited more than %.0f s for the words of rank %d", h->rank,
                 1e-9 * (double)XB_WAIT_NS, xs[6] - 1);
        }
    }
    CK(cudaEventRecord(h->ev_solve1, s));
    CK(cudaEventSynchronize(h->ev_solve1));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, h->ev_solve0, h->ev_solve1));
    if (h->time_kernels) resolve_stamps(h);
    h->time_kernels = false;
    h->trace_n = std::min(h->h_state->iter + 1, TRACE_CAP);
    if (out) {
        const LMState& st = *h->h_state;
        memset(out, 0, sizeof *out);
        out->status = status == 6 ? 0 : status;
        out->iterations = st.iter; out->accepted = st.accepted; out->rejected = st.rejected;
        out->pcg_iterations = st.pcg_total;
        long long L = 0;
        for (int k = 0; k < KN_COUNT; ++k) L += h->launches[k];
        out->launches = (int)L;
        out->initial_raw_cost = raw0; out->initial_cost = cost0;
        out->final_raw_cost = st.raw[st.cur]; out->final_cost = st.cost[st.cur];
        out->gradient_max = st.gmax; out->final_lambda = st.lambda; out->solve_ms = ms;
        out->precond_used = coarse_builds > 0 && !coarse_failed ? BA_PRECOND_TWO_LEVEL : with_schur ? BA_PRECOND_SCHUR_JACOBI : BA_PRECOND_JACOBI;
        out->coarse_builds = coarse_builds; out->coarse_matvecs = coarse_matvecs;
        out->coarse_unknowns = coarse_builds > 0 ? h->coarse_nz : 0;
        if (dense) {
            CK(cudaMemcpyAsync(&out->dense_solves, h->dn_count.p, sizeof(int), cudaMemcpyDeviceToHost, s));
            CK(cudaStreamSynchronize(s));
        }
    }
}

}  // namespace ba

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
#define API_BEGIN try {
#define API_END                                  \
    }                                            \
    catch (const ba::BaError& e) {               \
        ba::g_last_error = e.what();             \
        return e.code;                           \
    }                                            \
    catch (const std::exception& e) {            \
        ba::g_last_error = e.what();             \
        return BA_ERR_STATE;                     \
    }                                            \
    return BA_OK;

extern "C" {

const char* ba_last_error(void) { return ba::g_last_error.c_str(); }

void ba_default_options(ba_options* o) {
    if (!o) return;
    memset(o, 0, sizeof *o);
    o->max_iters = 50;
    o->gtol = 1e-8; o->ftol = 1e-5; o->xtol = 1e-5;  // the reference's call: src/bundle_adjuster.py:172
    o->huber_delta = 1.0; o->robust_mode = BA_ROBUST_IRLS;
    o->initial_lambda = 1e-4; o->lambda_min = 1e-6; o->lambda_max = 1e16;
    o->precond = BA_PRECOND_AUTO;
    o->pcg_max_iters = 100; o->pcg_tol = 0.1; o->pcg_check_every = 8;
    o->verbose = 0;
    o->schur_path = BA_PATH_TILES;
    o->pcg_tail = 1;
    o->device_loop = 1;
    o->point_linesearch = 1;
    o->dense_schur = 1;
}

int ba_create(int device, ba_handle** out) {
    API_BEGIN
    if (!out) ba::fail(BA_ERR_ARG, "null out pointer");
    *out = nullptr;
    ba_handle* h = new ba_handle();
    try {
        ba::create_common(h, device);
    } catch (...) {
        delete h;
        throw;
    }
    *out = h;
    API_END
}

int ba_nccl_unique_id(char id_out[128]) {
    API_BEGIN
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    ncclUniqueId id;
    NK(ba::nccl_api().GetUniqueId(&id));
    memcpy(id_out, &id, 128);
    API_END
}

int ba_create_distributed(int device, int rank, int nranks, const char nccl_id[128], ba_handle** out) {
    API_BEGIN
    if (!out || !nccl_id) ba::fail(BA_ERR_ARG, "null pointer");
    if (nranks < 1 || rank < 0 || rank >= nranks) ba::fail(BA_ERR_ARG, "bad rank %d of %d", rank, nranks);
    *out = nullptr;
    ba_handle* h = new ba_handle();
    try {
        ba::create_common(h, device);
        h->rank = rank; h->nranks = nranks;
        if (nranks > 1) {
            ncclUniqueId id;
            memcpy(&id, nccl_id, 128);
            NK(ba::nccl_api().CommInitRank(&h->comm, nranks, id, rank));
        }
    } catch (...) {
        delete h;
        throw;
    }
    *out = h;
    API_END
}

int ba_destroy(ba_handle* h) {
    API_BEGIN
    if (h) {
        cudaSetDevice(h->device);
        cudaStreamSynchronize(h->stream);
        // distributed handles: collective - the exchange buffer is freed only after every rank has unmapped it
        try { ba::close_peer_maps(h); } catch (...) { delete h; throw; }
        delete h;
    }
    API_END
}

int ba_set_problem(ba_handle* h, double fx, double fy, double cx, double cy, int32_t n_adj_cams, int32_t n_pts, int32_t n_obs,
                   const double* rvec, const double* tvec, const double fixed_rvec[3], const double fixed_tvec[3],
                   const double* pts, const int32_t* obs_cam, const int32_t* obs_pt, const double* obs_uv) {
    API_BEGIN
    ba::set_problem(h, fx, fy, cx, cy, n_adj_cams, n_pts, n_obs, rvec, tvec, fixed_rvec, fixed_tvec, pts, obs_cam, obs_pt, obs_uv);
    ba::set_params(h, rvec, tvec, fixed_rvec, fixed_tvec, pts, /*save_initial=*/true);
    API_END
}

int ba_set_params(ba_handle* h, const double* rvec, const double* tvec, const double* pts) {
    API_BEGIN
    ba::need_problem(h);
    const ba::Params& P = h->P;
    if (!pts || (P.n_cams > 0 && (!rvec || !tvec))) ba::fail(BA_ERR_ARG, "null array argument");
    // the fixed pose stays what ba_set_problem uploaded: fetch it from the current buffer
    ba::read_state(h);
    const int b = h->h_state->cur;
    double f[6];
    CK(cudaMemcpy(f, h->cam_rt.p + (size_t)b * (P.n_cams + 1) * 6 + (size_t)P.n_cams * 6, sizeof f, cudaMemcpyDeviceToHost));
    ba::set_params(h, rvec, tvec, f, f + 3, pts);
    API_END
}

int ba_get_params(ba_handle* h, double* rvec, double* tvec, double* pts) {
    API_BEGIN
    ba::need_problem(h);
    const ba::Params& P = h->P;
    cudaStream_t s = h->stream;
    ba::read_state(h);
    const int b = h->h_state->cur;
    if (pts) {
        ba::launch(h, ba::KN_SETUP, ba::k_unpack_points, ba::cdiv(P.n_pts, 256), 256, 0, P.n_pts, P.n_pad, (const int*)h->pt_perm.p,
                   (const double*)(h->pt_X.p + (size_t)b * P.pt_stride), h->io.p);
        CK(cudaMemcpyAsync(pts, h->io.p, (size_t)P.n_pts * 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
    }
    if ((rvec || tvec) && P.n_cams > 0) {
        ba::DBuf<double> tmp;
        tmp.alloc((size_t)P.n_cams * 6);
        ba::launch(h, ba::KN_SETUP, ba::k_unpack_cams, ba::cdiv(P.n_cams, 256), 256, 0, P.n_cams,
                   (const double*)(h->cam_rt.p + (size_t)b * (P.n_cams + 1) * 6), tmp.p, tmp.p + (size_t)P.n_cams * 3);
        if (rvec) CK(cudaMemcpyAsync(rvec, tmp.p, (size_t)P.n_cams * 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
        if (tvec) CK(cudaMemcpyAsync(tvec, tmp.p + (size_t)P.n_cams * 3, (size_t)P.n_cams * 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
    }
    API_END
}

int ba_restore_initial(ba_handle* h) {
    API_BEGIN
    ba::need_problem(h);
    const ba::Params& P = h->P;
    cudaStream_t s = h->stream;
    const size_t nc1 = (size_t)P.n_cams + 1;
    CK(cudaMemsetAsync(&h->st.p->cur, 0, sizeof(int), s));
    CK(cudaMemcpyAsync(h->cam_rt.p, h->init_cam_rt.p, nc1 * 6 * sizeof(double), cudaMemcpyDeviceToDevice, s));
    ba::launch(h, ba::KN_SETUP, ba::k_pack_points, ba::cdiv(P.n_pts, 256), 256, 0, P.n_pts, P.n_pad, (const int*)h->pt_perm.p,
               (const double*)h->init_pts.p, h->pt_X.p);
    ba::launch(h, ba::KN_CAM_PREP, ba::k_cam_prepare, ba::cdiv((int)nc1, 256), 256, 0, P, 0);
    API_END
}

int ba_timer_start(ba_handle* h) {
    API_BEGIN
    ba::use(h);
    CK(cudaEventRecord(h->ev_user0, h->stream));
    API_END
}

int ba_timer_stop(ba_handle* h, double* ms_out) {
    API_BEGIN
    ba::use(h);
    CK(cudaEventRecord(h->ev_user1, h->stream));
    CK(cudaEventSynchronize(h->ev_user1));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, h->ev_user0, h->ev_user1));
    if (ms_out) *ms_out = ms;
    API_END
}

int ba_fp64_peak(ba_handle* h, double* tflops_out) {
    API_BEGIN
    ba::use(h);
    if (!tflops_out) ba::fail(BA_ERR_ARG, "null pointer");
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
    const int grid = sms * 8, block = 256, iters = 4096;
    ba::DBuf<double> sink;
    sink.alloc((size_t)grid * block);
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {  // the first one warms up
        CK(cudaEventRecord(h->ev_user0, h->stream));
        ba::k_fp64_peak<<<grid, block, 0, h->stream>>>(sink.p, iters, 1.0 + 1e-9 * rep);
        CK(cudaGetLastError());
        CK(cudaEventRecord(h->ev_user1, h->stream));
        CK(cudaEventSynchronize(h->ev_user1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, h->ev_user0, h->ev_user1));
        const double flops = 2.0 * 16.0 * iters * (double)grid * block;
        if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    *tflops_out = best;
    API_END
}

int ba_residuals(ba_handle* h, double* r_out) {
    API_BEGIN
    ba::need_problem(h);
    if (!r_out) ba::fail(BA_ERR_ARG, "null output");
    const ba::Params& P = h->P;
    ba::DBuf<double2> d;
    d.alloc(P.n_obs);
    ba::launch(h, ba::KN_RESID, ba::k_residuals, ba::cdiv(P.n_obs, 256), 256, 0, P, d.p);
    CK(cudaMemcpyAsync(r_out, d.p, (size_t)P.n_obs * sizeof(double2), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    API_END
}

int ba_cost(ba_handle* h, double huber_delta, double* raw_out, double* robust_out) {
    API_BEGIN
    ba::need_problem(h);
    if (!(huber_delta > 0.0)) ba::fail(BA_ERR_ARG, "huber_delta must be positive");
    h->P.delta = huber_delta;
    h->P.mode = ba::ROBUST_IRLS;
    ba::lin_points_and_cost(h, 2);
    ba::reduce_scalars(h, ba::MASK_LIN, 0);
    ba::read_state(h);
    if (raw_out) *raw_out = h->h_state->scal[ba::S_RAW];
    if (robust_out) *robust_out = h->h_state->scal[ba::S_COST];
    API_END
}

int ba_jacobian_blocks(ba_handle* h, double* jc_out, double* jp_out) {
    API_BEGIN
    ba::need_problem(h);
    if (!jc_out || !jp_out) ba::fail(BA_ERR_ARG, "null output");
    const ba::Params& P = h->P;
    ba::DBuf<double> jc, jp;
    jc.alloc((size_t)P.n_obs * 12);
    jp.alloc((size_t)P.n_obs * 6);
    ba::launch(h, ba::KN_JAC, ba::k_jacobian_blocks, ba::cdiv(P.n_obs, 256), 256, 0, P, jc.p, jp.p);
    CK(cudaMemcpyAsync(jc_out, jc.p, jc.n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(jp_out, jp.p, jp.n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    API_END
}

int ba_sparsity_csr(ba_handle* h, int64_t* indptr, int64_t* indices, int64_t* nnz_out) {
    API_BEGIN
    ba::need_problem(h);
    static_assert(sizeof(long long) == sizeof(int64_t), "int64");
    const ba::Params& P = h->P;
    cudaStream_t s = h->stream;
    ba::DBuf<long long> cnt, start, d_indptr, d_indices;
    cnt.alloc(P.n_obs); start.alloc(P.n_obs); d_indptr.alloc(2 * (size_t)P.n_obs + 1);
    ba::launch(h, ba::KN_SETUP, ba::k_sparsity_counts, ba::cdiv(P.n_obs, 256), 256, 0, P.n_obs, P.n_cams, P.o_idx, cnt.p);
    size_t tb = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt.p, start.p, P.n_obs, s));
    ba::DBuf<unsigned char> tmp;
    tmp.alloc(tb);
    CK(cub::DeviceScan::ExclusiveSum(tmp.p, tb, cnt.p, start.p, P.n_obs, s));
    long long last_start = 0, last_cnt = 0;
    CK(cudaMemcpyAsync(&last_start, start.p + P.n_obs - 1, sizeof(long long), cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(&last_cnt, cnt.p + P.n_obs - 1, sizeof(long long), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    const long long nnz = last_start + last_cnt;
    if (nnz_out) *nnz_out = nnz;
    if (indptr || indices) {
        if (indices) d_indices.alloc((size_t)nnz);
        ba::launch(h, ba::KN_SETUP, ba::k_sparsity_fill, ba::cdiv(P.n_obs, 256), 256, 0, P.n_obs, P.n_cams, P.o_idx,
                   (const int*)h->pt_perm.p, (const long long*)start.p, d_indptr.p, indices ? d_indices.p : (long long*)nullptr);
        if (indptr) CK(cudaMemcpyAsync(indptr, d_indptr.p, d_indptr.n * sizeof(long long), cudaMemcpyDeviceToHost, s));
        if (indices) CK(cudaMemcpyAsync(indices, d_indices.p, (size_t)nnz * sizeof(long long), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
    }
    API_END
}

int ba_normal_blocks(ba_handle* h, const ba_options* opt, double* gc, double* gp, double* U, double* V) {
    API_BEGIN
    ba::need_problem(h);
    ba_options o;
    if (opt) o = *opt; else ba_default_options(&o);
    ba::apply_robust(h, &o);
    const ba::Params& P = h->P;
    cudaStream_t s = h->stream;
    ba::launch(h, ba::KN_CAM_PREP, ba::k_cam_prepare, ba::cdiv(P.n_cams + 1, 256), 256, 0, P, 2);
    ba::lin_points_and_cost(h, 2);
    ba::lin_cameras(h, ba::kDiagMin, ba::kDiagMax);
    ba::read_state(h);
    const int b = h->h_state->cur;
    if (gc && P.n_cams) CK(cudaMemcpyAsync(gc, h->cam_g.p, (size_t)P.n_cams * 6 * sizeof(double), cudaMemcpyDeviceToHost, s));
    std::vector<double> u21, v6, g4;
    std::vector<int> perm((size_t)P.n_pts);
    CK(cudaMemcpyAsync(perm.data(), h->pt_perm.p, perm.size() * sizeof(int), cudaMemcpyDeviceToHost, s));
    if (U && P.n_cams) {
        u21.resize((size_t)P.n_cams * 21);
        CK(cudaMemcpyAsync(u21.data(), h->cam_U.p, u21.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
    }
    if (V) {
        v6.resize((size_t)P.n_pts * 6);
        CK(cudaMemcpyAsync(v6.data(), h->pt_V.p + (size_t)b * P.n_pts * 6, v6.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
    }
    if (gp) {
        g4.resize((size_t)P.n_pts * 4);
        CK(cudaMemcpyAsync(g4.data(), h->pt_g.p + (size_t)b * P.n_pts * 4, g4.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
    }
    CK(cudaStreamSynchronize(s));
    if (U)
        for (int c = 0; c < P.n_cams; ++c)
            for (int i = 0; i < 6; ++i)
                for (int j = i; j < 6; ++j) {
                    const double v = u21[(size_t)c * 21 + ba::sym6(i, j)];
                    U[(size_t)c * 36 + 6 * i + j] = v;
                    U[(size_t)c * 36 + 6 * j + i] = v;
                }
    if (V)
        for (int q = 0; q < P.n_pts; ++q)
            for (int i = 0; i < 3; ++i)
                for (int j = i; j < 3; ++j) {
                    const double v = v6[(size_t)q * 6 + ba::sym3(i, j)];
                    V[(size_t)perm[q] * 9 + 3 * i + j] = v;
                    V[(size_t)perm[q] * 9 + 3 * j + i] = v;
                }
    if (gp)
        for (int q = 0; q < P.n_pts; ++q)
            for (int k = 0; k < 3; ++k) gp[(size_t)perm[q] * 3 + k] = g4[(size_t)q * 4 + k];
    API_END
}

int ba_schur_matvec(ba_handle* h, const ba_options* opt, double lambda, const double* u, double* w_out) {
    API_BEGIN
    ba::need_problem(h);
    if (!u || !w_out) ba::fail(BA_ERR_ARG, "null array argument");
    if (h->nranks > 1) ba::fail(BA_ERR_STATE, "ba_schur_matvec is a single-GPU test hook");
    ba_options o;
    if (opt) o = *opt; else ba_default_options(&o);
    ba::apply_robust(h, &o);
    const ba::Params& P = h->P;
    cudaStream_t s = h->stream;
    if (P.n_cams == 0) return BA_OK;
    const bool tiles = o.schur_path != BA_PATH_TWO_PASS && h->fused_ok;
    const bool acc = tiles && o.schur_path == BA_PATH_TILES && h->acc_ok;
    const ba::Params PA = ba::matvec_params(h, acc);
    ba::launch(h, ba::KN_SETUP, ba::k_init_state, 1, 1, 0, h->st.p, lambda);
    ba::launch(h, ba::KN_CAM_PREP, ba::k_cam_prepare, ba::cdiv(P.n_cams + 1, 256), 256, 0, P, 2);
    ba::lin_points_and_cost(h, 2);
    ba::lin_cameras(h, ba::kDiagMin, ba::kDiagMax);
    ba::launch(h, ba::KN_PT_PREP, ba::k_point_prepare, ba::cdiv(P.n_pts, 256), 256, 0, P, ba::kDiagMin, ba::kDiagMax);
    CK(cudaMemcpyAsync(h->v_u.p, u, (size_t)P.n_cams * 6 * sizeof(double), cudaMemcpyHostToDevice, s));
    if (tiles) {
        ba::launch_tile_matvec(h, PA, acc, h->v_u.p);
        ba::launch(h, ba::KN_ROWS_FIN, ba::k_rows_finish<false>, ba::cdiv(P.n_cams, ba::RS_WARPS), ba::RS_WARPS * 32, 0, PA);
    } else {
        ba::launch(h, ba::KN_PT_PASS, ba::k_point_pass<0>, P.pm_tiles, ba::PM_THREADS, h->smem_pass, P, (const double*)h->v_u.p);
        ba::launch_camera_pass<2>(h, ba::KN_MATVEC_CAM, h->v_u.p, 0);
        ba::launch(h, ba::KN_PCG_DOTS, ba::k_pcg_dots<true>, P.vec_ctas, ba::VEC_THREADS, 0, P);
    }
    CK(cudaMemcpyAsync(w_out, h->v_w.p, (size_t)P.n_cams * 6 * sizeof(double), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    API_END
}

int ba_solve(ba_handle* h, const ba_options* opt, ba_stats* stats_out) {
    API_BEGIN
    ba_options o;
    if (opt) o = *opt; else ba_default_options(&o);
    ba::solve(h, &o, stats_out);
    API_END
}

int ba_layout_info(ba_handle* h, int64_t out[12]) {
    API_BEGIN
    ba::need_problem(h);
    if (!out) ba::fail(BA_ERR_ARG, "null output");
    const ba::Params& P = h->P;
    const bool acc = h->fused_ok && h->acc_ok;
    const int64_t v[12] = {P.n_obs, P.n_obs_adj, P.n_pts, P.n_cams, h->fused_ok ? h->ft_tiles : 0, h->n_seg, P.pm_tiles, P.cm_tiles,
                           acc ? h->n_slots : h->n_seg, acc ? h->n_epochs : 0, h->max_ep_slots, acc ? h->ep_target : 0};
    for (int k = 0; k < 12; ++k) out[k] = v[k];
    API_END
}

int ba_tile_items(ba_handle* h, int64_t* n_slots, int32_t* items, int32_t* hdr_acc) {
    API_BEGIN
    ba::need_problem(h);
    if (!n_slots) ba::fail(BA_ERR_ARG, "null output");
    const bool acc = h->fused_ok && h->acc_ok;
    *n_slots = acc ? h->n_item_slots : 0;
    if (acc) {
        cudaStream_t s = h->stream;
        if (items) CK(cudaMemcpyAsync(items, h->seg_item.p, (size_t)h->n_item_slots * sizeof(int), cudaMemcpyDeviceToHost, s));
        if (hdr_acc) CK(cudaMemcpyAsync(hdr_acc, h->ft_hdr_acc.p, (size_t)h->ft_tiles * 8 * sizeof(int), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
    }
    API_END
}

int ba_tile_layout(ba_handle* h, int32_t* ft_pt, int32_t* ft_obs, int32_t* ft_seg, int32_t* tc_meta, double* tc_uv,
                   int32_t* seg_cam, int32_t* seg_info, int32_t* seg_row, int32_t* cam_row_off, int32_t* pt_perm) {
    API_BEGIN
    ba::need_problem(h);
    if (!h->fused_ok) ba::fail(BA_ERR_STATE, "no point tiles for this problem (a track is longer than a tile)");
    const ba::Params& P = h->P;
    cudaStream_t s = h->stream;
    const size_t T = (size_t)h->ft_tiles, S = (size_t)h->n_seg;
    auto get = [&](void* dst, const void* src, size_t bytes) {
        if (dst) CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s));
    };
    get(ft_pt, h->ft_pt.p, (T + 1) * 4); get(ft_obs, h->ft_obs.p, (T + 1) * 4); get(ft_seg, h->ft_seg.p, (T + 1) * 4);
    get(tc_meta, h->tc_meta.p, (size_t)P.n_obs * 8); get(tc_uv, h->tc_uv.p, (size_t)P.n_obs * 16);
    get(seg_cam, h->seg_cam.p, S * 4); get(seg_info, h->seg_info.p, S * 4); get(seg_row, h->seg_row.p, S * 4);
    get(cam_row_off, h->cam_row_off.p, ((size_t)P.n_cams + 2) * 4); get(pt_perm, h->pt_perm.p, (size_t)P.n_pts * 4);
    CK(cudaStreamSynchronize(s));
    API_END
}

int ba_triangulate(ba_handle* h, const double P1[12], const double P2[12], const double* R_rel, const double* t_rel, int32_t n,
                   const double* pts1, const double* pts2, double* hom, double* xyz, uint8_t* valid) {
    API_BEGIN
    ba::use(h);
    if (n < 0 || !P1 || !P2 || (n > 0 && (!pts1 || !pts2))) ba::fail(BA_ERR_ARG, "null or negative argument");
    if (valid && (!R_rel || !t_rel)) ba::fail(BA_ERR_ARG, "the cheirality test needs R_rel and t_rel");
    if (n == 0) return BA_OK;
    cudaStream_t s = h->stream;
    ba::ScratchScope scratch(s);
    ba::TriParams T;
    memcpy(T.P1, P1, sizeof T.P1); memcpy(T.P2, P2, sizeof T.P2);
    for (int k = 0; k < 9; ++k) T.R[k] = R_rel ? R_rel[k] : (k % 4 == 0 ? 1.0 : 0.0);
    for (int k = 0; k < 3; ++k) T.t[k] = t_rel ? t_rel[k] : 0.0;
    ba::TBuf<double> d1, d2, dh, dx;
    ba::TBuf<unsigned char> dv;
    d1.alloc((size_t)n * 2); d2.alloc((size_t)n * 2);
    if (hom) dh.alloc((size_t)n * 4);
    if (xyz) dx.alloc((size_t)n * 3);
    if (valid) dv.alloc(n);
    CK(cudaMemcpyAsync(d1.p, pts1, (size_t)n * 2 * sizeof(double), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d2.p, pts2, (size_t)n * 2 * sizeof(double), cudaMemcpyHostToDevice, s));
    ba::launch(h, ba::KN_SETUP, ba::k_triangulate, ba::cdiv(n, 128), 128, 0, T, (int)n, (const double*)d1.p, (const double*)d2.p, dh.p, dx.p,
               dv.p);
    if (hom) CK(cudaMemcpyAsync(hom, dh.p, (size_t)n * 4 * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (xyz) CK(cudaMemcpyAsync(xyz, dx.p, (size_t)n * 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (valid) CK(cudaMemcpyAsync(valid, dv.p, (size_t)n, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    API_END
}

// ---------------------------------------------------------------------------------------------
// ba_triangulate_static: ba_triangulate without any allocation (requested by the run's operators; no counterpart in the
// reference).  ba_triangulate stages pts1 / pts2 and the three outputs in scratch buffers from the stream-ordered pool
// (cudaMallocAsync / cudaFreeAsync); here the same scratch is ONE file-scope __device__ array of fixed size, carved into
// [pts1 2M | pts2 2M | hom 4M | xyz 3M | valid M bytes], M = BA_TRIANGULATE_STATIC_MAX correspondences (1.46 MB, held by every
// GPU that loads the library until the process exits).  In host code `g_tri_scratch` is only a stand-in for the symbol, not a
// device address: the address comes from cudaGetSymbolAddress, once per device with that device current, and is what the
// kernel is given.  Every call shares the array: it keeps the previous call's contents (the output part is cleared before the
// kernel runs), and two calls that overlap on the GPU - two handles / streams of one device - race.
// ---------------------------------------------------------------------------------------------
namespace ba {
constexpr size_t TRI_M = BA_TRIANGULATE_STATIC_MAX;
constexpr size_t TRI_SCRATCH_DOUBLES = 11 * TRI_M + TRI_M / 8;  // valid: M bytes
static_assert(TRI_M % 32 == 0, "the carve-up keeps every part 256-byte aligned");
static __device__ __align__(256) double g_tri_scratch[TRI_SCRATCH_DOUBLES];

constexpr int TRI_MAX_DEVICES = 64;
static double* g_tri_ptr[TRI_MAX_DEVICES] = {nullptr};  // device address of g_tri_scratch on each device (0: not resolved yet)
static std::mutex g_tri_mutex;
// set-up, once per device: the caller has made `device` current (ba::use)
static double* tri_scratch_ptr(int device) {
    if (device < 0 || device >= TRI_MAX_DEVICES) fail(BA_ERR_ARG, "device ordinal %d beyond %d", device, TRI_MAX_DEVICES - 1);
    std::lock_guard<std::mutex> lock(g_tri_mutex);
    if (!g_tri_ptr[device]) {
        double* ptr = nullptr;
        CK(cudaGetSymbolAddress((void**)&ptr, g_tri_scratch));
        if (!ptr) fail(BA_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the triangulation scratch");
        g_tri_ptr[device] = ptr;
    }
    return g_tri_ptr[device];
}
}  // namespace ba

int ba_triangulate_static(ba_handle* h, const double P1[12], const double P2[12], const double* R_rel, const double* t_rel,
                          int32_t n, const double* pts1, const double* pts2, double* hom, double* xyz, uint8_t* valid) {
    API_BEGIN
    ba::use(h);
    if (n < 0 || !P1 || !P2 || (n > 0 && (!pts1 || !pts2))) ba::fail(BA_ERR_ARG, "null or negative argument");
    if (valid && (!R_rel || !t_rel)) ba::fail(BA_ERR_ARG, "the cheirality test needs R_rel and t_rel");
    if (n > BA_TRIANGULATE_STATIC_MAX)
        ba::fail(BA_ERR_ARG, "ba_triangulate_static: %d correspondences exceed the static scratch (BA_TRIANGULATE_STATIC_MAX = %d); "
                             "split the batch or call ba_triangulate", (int)n, (int)BA_TRIANGULATE_STATIC_MAX);
    if (n == 0) return BA_OK;
    cudaStream_t s = h->stream;
    double* base = ba::tri_scratch_ptr(h->device);
    const size_t M = ba::TRI_M;
    double* d1 = base;
    double* d2 = base + 2 * M;
    double* dh = base + 4 * M;
    double* dx = base + 8 * M;
    unsigned char* dv = reinterpret_cast<unsigned char*>(base + 11 * M);
    ba::TriParams T;
    memcpy(T.P1, P1, sizeof T.P1); memcpy(T.P2, P2, sizeof T.P2);
    for (int k = 0; k < 9; ++k) T.R[k] = R_rel ? R_rel[k] : (k % 4 == 0 ? 1.0 : 0.0);
    for (int k = 0; k < 3; ++k) T.t[k] = t_rel ? t_rel[k] : 0.0;
    // the array keeps the last call's results: clear the output part this call will read back
    CK(cudaMemsetAsync(dh, 0, (size_t)n * 4 * sizeof(double), s));
    CK(cudaMemsetAsync(dx, 0, (size_t)n * 3 * sizeof(double), s));
    CK(cudaMemsetAsync(dv, 0, (size_t)n, s));
    CK(cudaMemcpyAsync(d1, pts1, (size_t)n * 2 * sizeof(double), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d2, pts2, (size_t)n * 2 * sizeof(double), cudaMemcpyHostToDevice, s));
    ba::k_triangulate<<<ba::cdiv(n, 128), 128, 0, s>>>(T, (int)n, (const double*)d1, (const double*)d2, hom ? dh : nullptr,
                                                       xyz ? dx : nullptr, valid ? dv : nullptr);
    CK(cudaGetLastError());
    if (hom) CK(cudaMemcpyAsync(hom, dh, (size_t)n * 4 * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (xyz) CK(cudaMemcpyAsync(xyz, dx, (size_t)n * 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (valid) CK(cudaMemcpyAsync(valid, dv, (size_t)n, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    API_END
}

int ba_solve_trace(ba_handle* h, int32_t cap, double* cost, double* raw_cost, double* ms, int32_t* pcg_total, int32_t* accepted,
                   double* lambda, int32_t* n_out) {
    API_BEGIN
    ba::need_problem(h);
    const int n = std::min<int>(h->trace_n, std::max(cap, 0));
    std::vector<ba::TraceRec> rec((size_t)std::max(n, 1));
    if (n > 0) {
        CK(cudaMemcpyAsync(rec.data(), h->trace.p, (size_t)n * sizeof(ba::TraceRec), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    for (int i = 0; i < n; ++i) {
        if (cost) cost[i] = rec[i].cost;
        if (raw_cost) raw_cost[i] = rec[i].raw;
        if (ms) ms[i] = rec[i].ms;
        if (pcg_total) pcg_total[i] = rec[i].pcg_total;
        if (accepted) accepted[i] = rec[i].accepted;
        if (lambda) lambda[i] = rec[i].lambda;
    }
    if (n_out) *n_out = h->trace_n;
    API_END
}

int ba_kernel_count(void) { return ba::KN_COUNT; }
const char* ba_kernel_name(int k) { return (k >= 0 && k < ba::KN_COUNT) ? ba::kKindNames[k] : ""; }
int ba_kernel_stats(ba_handle* h, int64_t* launches, double* ms) {
    API_BEGIN
    if (!h) ba::fail(BA_ERR_ARG, "null handle");
    for (int k = 0; k < ba::KN_COUNT; ++k) {
        if (launches) launches[k] = h->launches[k];
        if (ms) ms[k] = h->kernel_ms[k];
    }
    API_END
}

// ---- device-resident map (kernels in ba_map.cuh) ---------------------------------------------
}  // extern "C"

namespace ba {

static const int kMapMaxId = 1 << 24;               // dense tables: ids of keyframes and map points below this
static const long long kMapMaxPairs = 1LL << 28;    // (window keyframes) x (window points) entries of the pair table

// grow a resident buffer, keeping its first `used` elements; `fill` is the byte value of the new part (-1: leave it)
template <class T>
static void map_grow(DBuf<T>& b, size_t used, size_t need, cudaStream_t s, int fill) {
    if (need <= b.n) return;
    const size_t cap = std::max(need, std::max<size_t>(2 * b.n, 4096));
    T* q = nullptr;
    CK(cudaMalloc(&q, cap * sizeof(T)));
    if (used) CK(cudaMemcpyAsync(q, b.p, used * sizeof(T), cudaMemcpyDeviceToDevice, s));
    if (fill >= 0) CK(cudaMemsetAsync(q + used, fill, (cap - used) * sizeof(T), s));
    CK(cudaStreamSynchronize(s));
    if (b.p) cudaFree(b.p);
    b.p = q;
    b.n = cap;
}

static void map_need_single(ba_handle* h) {
    if (!h) fail(BA_ERR_ARG, "null handle");
    if (h->nranks != 1) fail(BA_ERR_STATE, "the resident map belongs to single-GPU handles (windows are far below one GPU)");
    use(h);
}

static void map_window(ba_handle* h, double fx, double fy, double cx, double cy, int n_kfs, const int* kf_id, const double* rvec,
                       const double* tvec, int* n_pts_out, int* n_obs_out) {
    map_need_single(h);
    if (n_kfs < 1 || !kf_id || !rvec || !tvec || !n_pts_out || !n_obs_out) fail(BA_ERR_ARG, "need n_kfs >= 1 and non-null arrays");
    cudaStream_t s = h->stream;
    ScratchScope scratch(s);
    *n_pts_out = 0; *n_obs_out = 0;
    h->mw_current = false;
    int max_kf = 0;
    for (int i = 0; i < n_kfs; ++i) {
        if (kf_id[i] < 0 || kf_id[i] >= kMapMaxId) fail(BA_ERR_ARG, "keyframe id %d outside [0, %d)", kf_id[i], kMapMaxId);
        max_kf = std::max(max_kf, kf_id[i]);
    }
    const int n_rows = (int)h->ms_rows, mp_cap = (int)h->mp_known.n;
    if (n_rows == 0 || mp_cap == 0) return;
    const int B = 256;
    map_grow(h->kf_slot, 0, (size_t)max_kf + 1, s, -1);
    const int kf_cap = (int)h->kf_slot.n;
    CK(cudaMemsetAsync(h->kf_slot.p, 0xff, (size_t)kf_cap * sizeof(int), s));
    TBuf<int> d_kf, d_val, d_order, d_flag, d_local, d_cnt;
    TBuf<unsigned> d_key, d_key_s;
    d_kf.alloc(n_kfs); d_val.alloc(n_rows); d_order.alloc(n_rows); d_key.alloc(n_rows); d_key_s.alloc(n_rows);
    d_flag.alloc((size_t)mp_cap + 1); d_local.alloc((size_t)mp_cap + 1); d_cnt.alloc(1);
    d_flag.zero(s); d_cnt.zero(s);
    CK(cudaMemcpyAsync(d_kf.p, kf_id, (size_t)n_kfs * sizeof(int), cudaMemcpyHostToDevice, s));
    launch(h, KN_SETUP, k_mw_slots, cdiv(n_kfs, B), B, 0, n_kfs, (const int*)d_kf.p, h->kf_slot.p);
    launch(h, KN_SETUP, k_mw_keys, cdiv(n_rows, B), B, 0, n_rows, n_kfs, kf_cap, mp_cap, (const int*)h->ms_kf.p, (const int*)h->ms_mp.p,
           (const int*)h->kf_slot.p, (const unsigned char*)h->mp_known.p, d_key.p, d_val.p, d_flag.p, d_cnt.p);
    int bits = 1;
    while ((1LL << bits) < (long long)n_kfs + 1) ++bits;
    size_t tb1 = 0, tb2 = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tb1, d_key.p, d_key_s.p, d_val.p, d_order.p, n_rows, 0, bits, s));
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tb2, d_flag.p, d_local.p, mp_cap + 1, s));
    TBuf<unsigned char> d_tmp;
    d_tmp.alloc(std::max(tb1, tb2));
    tb1 = d_tmp.n;
    CK(cub::DeviceRadixSort::SortPairs(d_tmp.p, tb1, d_key.p, d_key_s.p, d_val.p, d_order.p, n_rows, 0, bits, s));  // stable
    tb2 = d_tmp.n;
    CK(cub::DeviceScan::ExclusiveSum(d_tmp.p, tb2, d_flag.p, d_local.p, mp_cap + 1, s));
    h->launches[KN_SETUP] += 2;
    int n_obs = 0, n_pts = 0;
    CK(cudaMemcpyAsync(&n_obs, d_cnt.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(&n_pts, d_local.p + mp_cap, sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    if (n_obs == 0 || n_pts == 0) return;  // "No points in the local window" (src/bundle_adjuster.py:150-152)
    if ((long long)n_kfs * n_pts > kMapMaxPairs)
        fail(BA_ERR_ARG, "window of %d keyframes x %d points exceeds the pair table (%lld entries)", n_kfs, n_pts, kMapMaxPairs);
    TBuf<int> d_last, d_cam, d_pt;
    TBuf<double2> d_uv;
    TBuf<double> d_pts;
    d_last.alloc((size_t)n_kfs * n_pts); d_cam.alloc(n_obs); d_pt.alloc(n_obs); d_uv.alloc(n_obs); d_pts.alloc((size_t)n_pts * 3);
    h->mw_ids.alloc(n_pts);
    CK(cudaMemsetAsync(d_last.p, 0xff, (size_t)n_kfs * n_pts * sizeof(int), s));
    launch(h, KN_SETUP, k_mw_points, cdiv(mp_cap, B), B, 0, mp_cap, (const int*)d_flag.p, (const int*)d_local.p, (const double*)h->mp_xyz.p,
           h->mw_ids.p, d_pts.p);
    launch(h, KN_SETUP, k_mw_last, cdiv(n_obs, B), B, 0, n_obs, n_pts, (const unsigned*)d_key_s.p, (const int*)d_order.p,
           (const int*)h->ms_mp.p, (const int*)d_local.p, d_last.p);
    launch(h, KN_SETUP, k_mw_fill, cdiv(n_obs, B), B, 0, n_obs, n_pts, (const unsigned*)d_key_s.p, (const int*)d_order.p,
           (const int*)h->ms_mp.p, (const double2*)h->ms_uv.p, (const int*)d_local.p, (const int*)d_last.p, d_cam.p, d_pt.p, d_uv.p);
    // slot 0 is the fixed keyframe (src/bundle_adjuster.py:141-142): its pose is rvec[0..3) / tvec[0..3)
    set_problem(h, fx, fy, cx, cy, n_kfs - 1, n_pts, n_obs, rvec + 3, tvec + 3, rvec, tvec, d_pts.p, d_cam.p, d_pt.p,
                reinterpret_cast<const double*>(d_uv.p));
    set_params(h, rvec + 3, tvec + 3, rvec, tvec, d_pts.p, /*save_initial=*/true);
    h->mw_n_pts = n_pts;
    h->mw_current = true;
    *n_pts_out = n_pts; *n_obs_out = n_obs;
}

}  // namespace ba

extern "C" {

int ba_map_clear(ba_handle* h) {
    API_BEGIN
    ba::map_need_single(h);
    h->ms_rows = 0;
    h->mw_current = false;
    if (h->mp_known.p) h->mp_known.zero(h->stream);
    CK(cudaStreamSynchronize(h->stream));
    h->map_counters[0] = 0;  // the other counters are cumulative over the handle's life
    API_END
}

int ba_map_append_observations(ba_handle* h, int32_t kf_id, int32_t n, const int32_t* mp_id, const double* uv) {
    API_BEGIN
    ba::map_need_single(h);
    if (kf_id < 0 || kf_id >= ba::kMapMaxId) ba::fail(BA_ERR_ARG, "keyframe id %d outside [0, %d)", kf_id, ba::kMapMaxId);
    if (n < 0 || (n > 0 && (!mp_id || !uv))) ba::fail(BA_ERR_ARG, "need n >= 0 and non-null arrays");
    if (n == 0) return BA_OK;
    cudaStream_t s = h->stream;
    const size_t r0 = h->ms_rows, r1 = r0 + (size_t)n;
    if (r1 > (size_t)INT32_MAX) ba::fail(BA_ERR_ARG, "more than 2^31 resident observation rows");
    ba::map_grow(h->ms_kf, r0, r1, s, -1);
    ba::map_grow(h->ms_mp, r0, r1, s, -1);
    ba::map_grow(h->ms_uv, r0, r1, s, -1);
    CK(cudaMemcpyAsync(h->ms_mp.p + r0, mp_id, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(h->ms_uv.p + r0, uv, (size_t)n * sizeof(double2), cudaMemcpyHostToDevice, s));
    ba::launch(h, ba::KN_SETUP, ba::k_map_fill_rows, ba::cdiv(n, 256), 256, 0, (int)n, (int)kf_id, h->ms_kf.p + r0);
    CK(cudaStreamSynchronize(s));
    h->ms_rows = r1;
    h->map_counters[0] = (long long)r1;
    h->map_counters[1] += n;
    h->map_counters[4] += (long long)n * 20;
    API_END
}

int ba_map_set_points(ba_handle* h, int32_t n, const int32_t* mp_id, const double* xyz) {
    API_BEGIN
    ba::map_need_single(h);
    if (n < 0 || (n > 0 && (!mp_id || !xyz))) ba::fail(BA_ERR_ARG, "need n >= 0 and non-null arrays");
    if (n == 0) return BA_OK;
    int max_id = 0;
    for (int i = 0; i < n; ++i) {
        if (mp_id[i] < 0 || mp_id[i] >= ba::kMapMaxId) ba::fail(BA_ERR_ARG, "map point id %d outside [0, %d)", mp_id[i], ba::kMapMaxId);
        max_id = std::max(max_id, (int)mp_id[i]);
    }
    cudaStream_t s = h->stream;
    ba::ScratchScope scratch(s);
    const size_t cap0 = h->mp_known.n;
    ba::map_grow(h->mp_known, cap0, (size_t)max_id + 1, s, 0);
    ba::map_grow(h->mp_xyz, 3 * cap0, 3 * h->mp_known.n, s, 0);
    ba::TBuf<int> d_id;
    ba::TBuf<double> d_xyz;
    d_id.alloc(n); d_xyz.alloc((size_t)n * 3);
    CK(cudaMemcpyAsync(d_id.p, mp_id, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_xyz.p, xyz, (size_t)n * 3 * sizeof(double), cudaMemcpyHostToDevice, s));
    ba::launch(h, ba::KN_SETUP, ba::k_map_set_points, ba::cdiv(n, 256), 256, 0, (int)n, (const int*)d_id.p, (const double*)d_xyz.p,
               h->mp_xyz.p, h->mp_known.p);
    CK(cudaStreamSynchronize(s));
    h->mw_current = false;  // the uploaded window may hold older positions
    h->map_counters[2] += n;
    h->map_counters[4] += (long long)n * 28;
    API_END
}

int ba_map_forget_points(ba_handle* h, int32_t n, const int32_t* mp_id) {
    API_BEGIN
    ba::map_need_single(h);
    if (n < 0 || (n > 0 && !mp_id)) ba::fail(BA_ERR_ARG, "need n >= 0 and a non-null array");
    if (n == 0 || h->mp_known.n == 0) return BA_OK;
    cudaStream_t s = h->stream;
    ba::ScratchScope scratch(s);
    ba::TBuf<int> d_id;
    d_id.alloc(n);
    CK(cudaMemcpyAsync(d_id.p, mp_id, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, s));
    ba::launch(h, ba::KN_SETUP, ba::k_map_forget_points, ba::cdiv(n, 256), 256, 0, (int)n, (int)h->mp_known.n, (const int*)d_id.p,
               h->mp_known.p);
    CK(cudaStreamSynchronize(s));
    h->mw_current = false;
    API_END
}

int ba_map_window(ba_handle* h, double fx, double fy, double cx, double cy, int32_t n_kfs, const int32_t* kf_id, const double* rvec,
                  const double* tvec, int32_t* n_pts_out, int32_t* n_obs_out) {
    API_BEGIN
    ba::map_window(h, fx, fy, cx, cy, n_kfs, kf_id, rvec, tvec, n_pts_out, n_obs_out);
    h->map_counters[3] += 1;
    h->map_counters[4] += (long long)n_kfs * 52;
    API_END
}

int ba_map_window_points(ba_handle* h, int32_t* mp_id_out) {
    API_BEGIN
    ba::map_need_single(h);
    if (!h->mw_current || !h->have_problem) ba::fail(BA_ERR_STATE, "the uploaded problem was not built by ba_map_window");
    if (!mp_id_out) ba::fail(BA_ERR_ARG, "null array argument");
    CK(cudaMemcpyAsync(mp_id_out, h->mw_ids.p, (size_t)h->mw_n_pts * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    API_END
}

int ba_map_commit(ba_handle* h) {
    API_BEGIN
    ba::map_need_single(h);
    if (!h->mw_current || !h->have_problem) ba::fail(BA_ERR_STATE, "the uploaded problem was not built by ba_map_window");
    const ba::Params& P = h->P;
    ba::read_state(h);
    const int b = h->h_state->cur;
    ba::launch(h, ba::KN_SETUP, ba::k_map_commit, ba::cdiv(P.n_pts, 256), 256, 0, P.n_pts, P.n_pad, (const int*)h->pt_perm.p,
               (const double*)(h->pt_X.p + (size_t)b * P.pt_stride), (const int*)h->mw_ids.p, h->mp_xyz.p);
    CK(cudaStreamSynchronize(h->stream));
    API_END
}

int ba_map_stats(ba_handle* h, int64_t out[8]) {
    API_BEGIN
    if (!h || !out) ba::fail(BA_ERR_ARG, "null argument");
    for (int k = 0; k < 8; ++k) out[k] = h->map_counters[k];
    API_END
}

}  // extern "C"

#ifdef BA_DN_STAMPS
extern "C" int ba_ls_moved(unsigned long long out[8]) {  // reads and clears (scripts/ls_moved.py)
    unsigned long long z[8] = {0};
    if (cudaMemcpyFromSymbol(out, ba::g_ls_moved, sizeof z) != cudaSuccess) return -2;
    return cudaMemcpyToSymbol(ba::g_ls_moved, z, sizeof z) == cudaSuccess ? 0 : -2;
}
extern "C" int ba_dn_stamps(unsigned long long out[16]) {
    return cudaMemcpyFromSymbol(out, ba::g_dn_stamps, 16 * sizeof(unsigned long long)) == cudaSuccess ? 0 : -2;
}
#endif
