// ba_peer.cuh - the PCG iteration's camera-vector exchange between the GPUs of one box, fused into the
// kernels on either side of it (no NCCL call inside the PCG loop).
//
// One process per GPU; every rank owns a slice of the points and ALL cameras.  After the tile pass each
// rank holds its partial of the Schur mat-vec (6 doubles per camera).  Instead of  rows_finish ->
// ncclAllReduce -> pcg_dots  the ranks exchange through PEER MEMORY (cudaIpc-mapped buffers, NVLink /
// NVSwitch loads):
//
//   k_rows_finish_xchg   adds this rank's rows per camera straight into the rank's EXCHANGE SLOT (device
//                        memory that every peer has mapped); the last CTA (integer ticket) publishes the
//                        slot: sequence number -> flag, system-scope release.
//   k_pcg_dots_peer      waits for every rank's flag (system-scope acquire, bounded spin), then each thread
//                        reads its camera's six partials from every rank's slot IN RANK ORDER, adds them
//                        (so all ranks get bit-identical sums, and the same bits run to run), and goes on
//                        with what k_pcg_dots does: w = sum + lambda D u, partial of w.u.
//
// PUSH (default) or PULL.  Pull is what is described above: a rank writes its own buffer, the peers poll and read
// it over NVLink (two round trips after the slowest rank has published: flag, then data).  Push turns both
// around: a rank STORES its partials into the slot it owns in EVERY rank's buffer, fences, and stores the flag into
// every rank's buffer; the receivers poll and read LOCAL memory.  Stores are one-way and pipelined, so the exchange
// costs one fence (the round trip that acknowledges the remote stores) instead of two dependent remote reads.
//
// Two slot sets alternate by sequence parity: a rank can only be one exchange ahead of the slowest rank
// (exchange n+1 needs everybody's flag n+1, which a rank sets after it has finished reading n), so the set
// (n & 1) is never rewritten while a rank still reads exchange n.
//
// Exchange buffer of a rank (doubles): [parity * XB_MAX_RANKS + src] 64-bit flags = sequence number of the last
// exchange whose partials of rank `src` are complete, then payload slots [parity][src][cap].  Pull uses only the
// slots src = own rank of the own buffer.
#pragma once
#include "ba_tiles.cuh"

#include <cooperative_groups.h>

namespace ba {

constexpr int XB_MAX_RANKS = 16;           // ranks that can exchange through peer memory (more: NCCL all-reduce)
constexpr int XB_HDR = 2 * XB_MAX_RANKS;   // doubles reserved in front of the payloads (the flags)
// A rank waits for a peer's words for at most XB_WAIT_NS of device time (%globaltimer, looked at every 256 polls), then gives
// up, records the rank it was waiting for (xstate[6] = rank + 1) and raises the error flag xstate[2]: the solve fails with
// BA_ERR_NCCL naming the rank instead of hanging on a dead peer.
constexpr unsigned long long XB_WAIT_NS = 4000000000ull;  // 4 s: three orders of magnitude above a whole LM iteration
struct SpinGuard {
    unsigned long long t0 = 0;
    long long spins = 0;
    BA_D bool expired() {  // call once per failed poll
        if ((++spins & 255) != 0) return false;
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t0 == 0) { t0 = t; return false; }
        return t - t0 > XB_WAIT_NS;
    }
};
BA_D void spin_failed(int* xstate, int waited_for_rank) {
    atomicExch(xstate + 2, 1);
    atomicCAS(xstate + 6, 0, waited_for_rank + 1);
}
__host__ __device__ inline size_t xb_slot(long long seq, int src, int nranks, int cap) {
    return (size_t)XB_HDR + ((size_t)(seq & 1) * nranks + src) * cap;
}
// LL slots (k_pcg_tail, push == 2): every double travels as two 8-byte words (32 data bits | 32-bit sequence number), so
// the data is its own flag - no fence, no flag store, no barrier between them (the scheme of NCCL's LL protocol; 8-byte
// stores are single-copy atomic).  [parity][src][cap] pairs of words, behind the flag-synchronised slots.
__host__ __device__ inline size_t xb_ll_base(int nranks, int cap) { return (size_t)XB_HDR + 2 * (size_t)nranks * cap; }
__host__ __device__ inline size_t xb_doubles(int nranks, int cap) { return xb_ll_base(nranks, cap) + 4 * (size_t)nranks * cap; }
BA_D ulonglong2* xb_ll_slot(double* buf, long long seq, int src, int nranks, int cap) {
    return reinterpret_cast<ulonglong2*>(buf + xb_ll_base(nranks, cap)) + ((size_t)(seq & 1) * nranks + src) * cap;
}
BA_D void st_relaxed_sys_v2(ulonglong2* p, unsigned long long a, unsigned long long b) {
    asm volatile("st.relaxed.sys.global.v2.b64 [%0], {%1, %2};\n" ::"l"(p), "l"(a), "l"(b) : "memory");
}
BA_D ulonglong2 ld_relaxed_sys_v2(const ulonglong2* p) {
    ulonglong2 v;
    asm volatile("ld.relaxed.sys.global.v2.b64 {%0, %1}, [%2];\n" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    return v;
}
BA_D long long* xb_flag(double* buf, long long seq, int src) {
    return reinterpret_cast<long long*>(buf) + (seq & 1) * XB_MAX_RANKS + src;
}

BA_D void st_release_sys(long long* p, long long v) {
    asm volatile("st.release.sys.global.s64 [%0], %1;\n" ::"l"(p), "l"(v) : "memory");
}
BA_D long long ld_acquire_sys(const long long* p) {
    long long v;
    asm volatile("ld.acquire.sys.global.s64 %0, [%1];\n" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// xstate: [0] sequence number of the last exchange this rank published, [1] ticket, [2] error flag, [3] spare (host barrier),
// [4] / [5] sequence number / ticket of the LL all-reduce, [6] rank + 1 a timed-out wait was for
__global__ void __launch_bounds__(RS_WARPS * 32) k_rows_finish_xchg(Params P, double* __restrict__ xbuf, double* const* __restrict__ peers,
                                                                    int nranks, int rank, int push, int cap, int* __restrict__ xstate) {
    __shared__ int last;
    if (P.st->pcg_done) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int c = blockIdx.x * RS_WARPS + wid;
    const long long seq = (long long)xstate[0] + 1;  // only the last CTA changes xstate[0], after everybody has read it
    const size_t slot = xb_slot(seq, rank, nranks, cap);
    if (c < P.n_cams) {
        const int r0 = __ldg(P.cam_row_off + c), r1 = __ldg(P.cam_row_off + c + 1);
        double a[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        for (int r = r0 + lane; r < r1; r += 32) {
            const double2* q = reinterpret_cast<const double2*>(P.rows + (size_t)r * 6);
            const double2 v0 = q[0], v1 = q[1], v2 = q[2];
            a[0] += v0.x; a[1] += v0.y; a[2] += v1.x; a[3] += v1.y; a[4] += v2.x; a[5] += v2.y;
        }
#pragma unroll
        for (int k = 0; k < 6; ++k) a[k] = __shfl_sync(0xffffffffu, warp_sum(a[k]), 0);  // the total is lane 0's
        // push: lane d stores the camera's sums into rank d's buffer; pull: lane 0 into the own buffer
        if (lane < (push ? nranks : 1)) {
            double2* o = reinterpret_cast<double2*>((push ? peers[lane] : xbuf) + slot + (size_t)c * 6);
            o[0] = make_double2(a[0], a[1]); o[1] = make_double2(a[2], a[3]); o[2] = make_double2(a[4], a[5]);
        }
    }
    __threadfence_system();  // this thread's stores, before the CTA's ticket
    __syncthreads();
    if (threadIdx.x == 0) last = (atomicAdd(xstate + 1, 1) == (int)gridDim.x - 1);
    __syncthreads();
    if (last) {
        if (threadIdx.x == 0) {
            __threadfence_system();
            xstate[1] = 0;
            xstate[0] = (int)seq;
        }
        if (threadIdx.x < (push ? nranks : 1)) {
            __threadfence_system();
            st_release_sys(xb_flag(push ? peers[threadIdx.x] : xbuf, seq, rank), seq);  // publish
        }
    }
}

// peers[r] = exchange buffer of rank r as mapped into this process (peers[own rank] = the local buffer)
__global__ void __launch_bounds__(VEC_THREADS) k_pcg_dots_peer(Params P, double* __restrict__ xbuf, double* const* __restrict__ peers,
                                                               int nranks, int push, int cap, int* __restrict__ xstate) {
    __shared__ double sm[VEC_THREADS / 32];
    __shared__ int ok;
    if (P.st->pcg_done) return;
    const long long seq = xstate[0];
    if (threadIdx.x == 0) ok = 1;
    __syncthreads();
    if (threadIdx.x < nranks) {  // one thread per rank polls that rank's flag
        const long long* flag = xb_flag(push ? xbuf : peers[threadIdx.x], seq, threadIdx.x);
        SpinGuard g;
        while (ld_acquire_sys(flag) < seq)
            if (g.expired()) { spin_failed(xstate, threadIdx.x); ok = 0; break; }
    }
    __syncthreads();
    const int c = blockIdx.x * VEC_THREADS + threadIdx.x;
    double part = 0.0;
    if (c < P.n_cams && ok) {
        const double lam = P.st->lambda;
        double w[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        for (int r = 0; r < nranks; ++r) {  // rank order: identical bits on every rank
            const double2* q = reinterpret_cast<const double2*>((push ? xbuf : peers[r]) + xb_slot(seq, r, nranks, cap) + (size_t)c * 6);
            const double2 v0 = __ldcv(q), v1 = __ldcv(q + 1), v2 = __ldcv(q + 2);
            w[0] += v0.x; w[1] += v0.y; w[2] += v1.x; w[3] += v1.y; w[4] += v2.x; w[5] += v2.y;
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

// ---------------------------------------------------------------------------------------------
// All-reduce (sum or max) of a camera-sized vector over the ranks through the LL words of the exchange buffer, for the sums
// that are needed once per LM iteration (U / g_c, Schur right-hand side / preconditioner blocks, cost scalars): every rank
// stores its vector into its slot of every rank's buffer - data and sequence number in the same 8-byte words - and adds the
// ranks' vectors in rank order, so every rank gets identical bits.  One kernel of a few microseconds instead of an
// ncclAllReduce of ~25 us on eight GPUs, and nothing in an LM iteration calls into NCCL any more (capturable).
// The region sits behind the tail's slots in the same peer-mapped allocation: [parity][src][ar_cap] pairs of words.
// xstate[4] = sequence number of the last all-reduce, xstate[5] = ticket.
// ---------------------------------------------------------------------------------------------
constexpr int AR_THREADS = 256;
constexpr int AR_MAX_DOUBLES = 65536;  // larger vectors (10 k cameras x 27) go through NCCL: all-to-all traffic grows with N
__global__ void __launch_bounds__(AR_THREADS) k_peer_allreduce(double* __restrict__ data, int n, int is_max, double* __restrict__ xbuf,
                                                               double* const* __restrict__ peers, int nranks, int rank,
                                                               size_t ar_base, int ar_cap, int* __restrict__ xstate) {
    const long long seq = (long long)xstate[4] + 1;  // written by the last CTA only, after every CTA has taken its ticket
    const unsigned long long f = (unsigned long long)(unsigned)seq << 32;
    const int stride = gridDim.x * AR_THREADS;
    const size_t slot_off = (size_t)(seq & 1) * nranks;
    for (int e = blockIdx.x * AR_THREADS + threadIdx.x; e < n; e += stride) {
        const unsigned long long bits = (unsigned long long)__double_as_longlong(data[e]);
        for (int d = 0; d < nranks; ++d)
            st_relaxed_sys_v2(reinterpret_cast<ulonglong2*>(peers[d] + ar_base) + (slot_off + rank) * ar_cap + e, (bits & 0xffffffffull) | f,
                              (bits >> 32) | f);
    }
    for (int e = blockIdx.x * AR_THREADS + threadIdx.x; e < n; e += stride) {
        double w = 0.0;
        for (int r = 0; r < nranks; ++r) {
            const ulonglong2* src = reinterpret_cast<const ulonglong2*>(xbuf + ar_base) + (slot_off + r) * ar_cap + e;
            ulonglong2 v = ld_relaxed_sys_v2(src);
            SpinGuard g;
            while ((unsigned)(v.x >> 32) != (unsigned)seq || (unsigned)(v.y >> 32) != (unsigned)seq) {
                if (g.expired()) { spin_failed(xstate, r); break; }
                v = ld_relaxed_sys_v2(src);
            }
            const double x = __longlong_as_double((long long)((v.x & 0xffffffffull) | (v.y << 32)));
            w = (is_max && r > 0) ? fmax(w, x) : (is_max ? x : w + x);
        }
        data[e] = w;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(xstate + 5, 1) == (int)gridDim.x - 1) {
            xstate[5] = 0;
            xstate[4] = (int)seq;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// The TAIL of a PCG iteration in one kernel: everything between two tile passes.
//
//   rows (one per accumulator slot / segment)  ->  per-camera sums  [-> exchange with the other GPUs]
//   -> w = sum + lambda D u,  delta = w.u  ->  alpha, beta  ->  p, s, x, r  ->  u = Minv r,  gamma = r.u
//   -> PCG state (iteration count, convergence flag)
//
// One thread-block CLUSTER (8 or 16 CTAs on neighbouring SMs) does all of it: the two inner products need a
// barrier across all cameras, and a cluster has one in hardware (barrier.cluster) plus distributed shared
// memory for the partial sums - no second and third kernel launch, no ticket, and the camera vectors of the
// iteration stay in shared memory between the two halves.  Each CTA owns a contiguous chunk of cameras.
// All sums have a fixed order (lane order, warp order, CTA rank order, GPU rank order): bit-reproducible,
// and identical on every GPU of a multi-GPU run.
//
// PEER: the per-camera sums of this GPU go to its exchange slot (see above), the cluster publishes the slot,
// waits for the flags of all ranks and adds the ranks' partials in rank order.
// ---------------------------------------------------------------------------------------------
namespace cg = cooperative_groups;
constexpr int PT_THREADS = 1024;
constexpr int PT_MAX_CLUSTER = 16;
__host__ __device__ inline size_t pt_smem_bytes(int n_cams, int cluster, int coarse_nz = 0, int coarse_agg = 1) {
    const int chunk = (n_cams + cluster - 1) / cluster;
    // two-level preconditioner: + the restricted residual [nz] and the correction of the aggregates that meet a chunk
    const size_t coarse = coarse_nz ? (size_t)coarse_nz + (size_t)(chunk / coarse_agg + 2) * 6 : 0;
    return ((size_t)chunk * 6 * 2 + coarse) * sizeof(double);
}

// Two-level preconditioner (BA_PRECOND_TWO_LEVEL; DESIGN.md section 8, oracle/lm_schur.py: pcg with Z / A0_inv):
//     M^-1 r = blockdiag(Minv) r + Z A0^-1 Z^T r,
// Z = one rigid 6-vector per aggregate of `agg` consecutive cameras, A0 = Z^T S Z (built by probing, solver side).
struct Coarse {
    const double* Ainv;  // [nz * nz] inverse of the coarse matrix (symmetric), nz = 6 * number of aggregates
    int agg, nz;
};
// (Z^T r)[j], j < nz, into shared memory: thread j adds its aggregate's residual component in camera order (fixed order ->
// the same bits in every CTA).  Callers synchronise afterwards.
BA_D void coarse_restrict(const Coarse& C, const double* __restrict__ r, int n_cams, double* __restrict__ s_rc) {
    for (int j = threadIdx.x; j < C.nz; j += blockDim.x) {
        const int a = j / 6, d = j - 6 * a;
        const int c0 = a * C.agg, c1 = min(c0 + C.agg, n_cams);
        double rc = 0.0;
        for (int c = c0; c < c1; ++c) rc += r[(size_t)c * 6 + d];
        s_rc[j] = rc;
    }
}
// y[row] = sum_j Ainv[row][j] * rc[j] by one warp (lane-strided products, shuffle tree); all lanes return the value.
BA_D double coarse_row(const Coarse& C, const double* __restrict__ s_rc, int row, int lane) {
    double acc = 0.0;
    for (int j = lane; j < C.nz; j += 32) acc += __ldg(C.Ainv + (size_t)row * C.nz + j) * s_rc[j];
    return __shfl_sync(0xffffffffu, warp_sum(acc), 0);
}

template <bool PEER, bool COARSE = false>
__global__ void __launch_bounds__(PT_THREADS, 1) k_pcg_tail(Params P, double tol, int max_iters, double* __restrict__ xbuf,
                                                            double* const* __restrict__ peers, int nranks, int rank, int push,
                                                            int cap, int* __restrict__ xstate, unsigned long long pcg_loop,
                                                            Coarse C = Coarse{nullptr, 0, 0}) {
    extern __shared__ __align__(16) double dsm[];
    __shared__ double part_d[PT_MAX_CLUSTER];  // delta partials of all CTAs (every CTA gets all of them)
    __shared__ double part_g[PT_MAX_CLUSTER];  // gamma partials (CTA 0 only)
    __shared__ double red[33];
    __shared__ int ok;
    LMState* st = P.st;
    if (st->pcg_done) return;  // same value in every CTA: it is only written after the last cluster barrier
    cg::cluster_group cl = cg::this_cluster();
    const int R = (int)cl.block_rank(), CL = (int)cl.num_blocks();
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int chunk = (P.n_cams + CL - 1) / CL;
    const int c0 = min(R * chunk, P.n_cams), c1 = min(c0 + chunk, P.n_cams);
    const int n_el = (c1 - c0) * 6;
    double* sW = dsm;              // [chunk*6] mat-vec result of the chunk, later s
    double* sR = dsm + chunk * 6;  // [chunk*6] new residual of the chunk
    const double gamma = st->gamma, alpha_prev = st->alpha, gamma_prev = st->gamma_prev, lam = st->lambda;
    const int first = (st->pcg_iter == 0);
    const long long seq = PEER ? (long long)xstate[0] + 1 : 0;

    // ---- per-camera sums of this GPU's rows (camera-major, so a chunk's rows are one contiguous range):
    //      a group of `lpc` lanes per camera, all cameras of the chunk side by side --------------------------
    {
        int lpc = 32;  // largest power of two with chunk * lpc <= PT_THREADS
        while (lpc > 1 && chunk * lpc > PT_THREADS) lpc >>= 1;
        const int li = tid & (lpc - 1);
        const int gpw = 32 / lpc;  // groups (cameras) per warp
        for (int cb = c0 + wid * gpw; cb < c1; cb += (PT_THREADS / 32) * gpw) {  // warp-uniform; one trip unless chunk > PT_THREADS
            const int c = cb + lane / lpc;
            const bool valid = c < c1;
            const int r0 = valid ? __ldg(P.cam_row_off + c) : 0, r1 = valid ? __ldg(P.cam_row_off + c + 1) : 0;
            double a[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
            for (int r = r0 + li; r < r1; r += 2 * lpc) {  // two rows in flight per lane
                const double2* q = reinterpret_cast<const double2*>(P.rows + (size_t)r * 6);
                const bool two = r + lpc < r1;
                const double2 v0 = q[0], v1 = q[1], v2 = q[2];
                const double2 z = make_double2(0.0, 0.0);
                const double2 w0 = two ? q[3 * lpc] : z, w1 = two ? q[3 * lpc + 1] : z, w2 = two ? q[3 * lpc + 2] : z;
                a[0] += v0.x; a[1] += v0.y; a[2] += v1.x; a[3] += v1.y; a[4] += v2.x; a[5] += v2.y;
                a[0] += w0.x; a[1] += w0.y; a[2] += w1.x; a[3] += w1.y; a[4] += w2.x; a[5] += w2.y;
            }
            for (int d = lpc >> 1; d > 0; d >>= 1)  // fixed order inside the group
#pragma unroll
                for (int k = 0; k < 6; ++k) a[k] += __shfl_down_sync(0xffffffffu, a[k], d, 32);
            if (li == 0 && valid) {
                double2* o = reinterpret_cast<double2*>(sW + (size_t)(c - c0) * 6);
                o[0] = make_double2(a[0], a[1]); o[1] = make_double2(a[2], a[3]); o[2] = make_double2(a[4], a[5]);
            }
        }
    }
    if (PEER) {
        __syncthreads();
        // the chunk's sums go (coalesced) into this rank's slot of every rank's buffer (push) or of the own buffer (pull)
        const size_t slot = xb_slot(seq, rank, nranks, cap) + (size_t)c0 * 6;
        if (push == 2) {
            // LL: (data, sequence) words straight into every rank's buffer; the readers below spin on the words themselves
            const unsigned long long f = (unsigned long long)(unsigned)seq << 32;
            for (int e = tid; e < n_el; e += PT_THREADS) {
                const unsigned long long bits = (unsigned long long)__double_as_longlong(sW[e]);
                for (int d = 0; d < nranks; ++d)
                    st_relaxed_sys_v2(xb_ll_slot(peers[d], seq, rank, nranks, cap) + (size_t)c0 * 6 + e, (bits & 0xffffffffull) | f, (bits >> 32) | f);
            }
            if (tid == 0) ok = 1;
        } else {
            for (int e = tid; e < n_el; e += PT_THREADS) {
                const double v = sW[e];
                if (push) {
                    for (int d = 0; d < nranks; ++d) peers[d][slot + e] = v;
                } else {
                    xbuf[slot + e] = v;
                }
            }
            __threadfence_system();  // this thread's part of the slot(s), before the cluster says "complete"
            cl.sync();
            if (R == 0 && tid < (push ? nranks : 1))
                st_release_sys(xb_flag(push ? peers[tid] : xbuf, seq, rank), seq);  // publish
            if (tid == 0) ok = 1;
            __syncthreads();
            if (tid < nranks) {  // one thread per rank polls that rank's flag (pull: the NVLink round trips overlap)
                const long long* flag = xb_flag(push ? xbuf : peers[tid], seq, tid);
                SpinGuard g;
                bool good = true;
                while (ld_acquire_sys(flag) < seq)
                    if (g.expired()) { good = false; break; }
                if (!good) { spin_failed(xstate, tid); ok = 0; }
            }
        }
    }
    __syncthreads();

    // ---- w = sum (+ the other ranks' sums, rank order) + lambda D u;  delta = w.u -------------------------
    double part = 0.0;
    for (int e = tid; e < n_el; e += PT_THREADS) {
        const size_t i = (size_t)c0 * 6 + e;
        double w = 0.0;
        if (PEER && push == 2) {
            const unsigned f = (unsigned)seq;
            for (int r0 = 0; r0 < nranks; r0 += 8) {  // eight loads in flight; a word that has not arrived yet is polled again
                ulonglong2 v[8];
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (r0 + j < nranks) v[j] = ld_relaxed_sys_v2(xb_ll_slot(xbuf, seq, r0 + j, nranks, cap) + i);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (r0 + j >= nranks) continue;
                    SpinGuard g;
                    while ((unsigned)(v[j].x >> 32) != f || (unsigned)(v[j].y >> 32) != f) {
                        if (g.expired()) { spin_failed(xstate, r0 + j); break; }
                        v[j] = ld_relaxed_sys_v2(xb_ll_slot(xbuf, seq, r0 + j, nranks, cap) + i);
                    }
                    w += __longlong_as_double((long long)((v[j].x & 0xffffffffull) | (v[j].y << 32)));
                }
            }
        } else if (PEER) {
            if (ok)
                for (int r0 = 0; r0 < nranks; r0 += 8) {  // eight loads in flight, then the adds in rank order
                    double v[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        v[j] = (r0 + j < nranks) ? __ldcv((push ? xbuf : peers[r0 + j]) + xb_slot(seq, r0 + j, nranks, cap) + i) : 0.0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) w += v[j];
                }
        } else {
            w = sW[e];
        }
        const double u = P.v_u[i];
        w += lam * P.cam_D[i] * u;
        sW[e] = w;
        part += w * u;
    }
    const double dpart = cta_sum_all(part, red);
    if (tid < CL) cl.map_shared_rank(part_d, tid)[R] = dpart;  // my partial into every CTA's table
    cl.sync();
    double delta = 0.0;
    for (int r = 0; r < CL; ++r) delta += part_d[r];
    const double beta = first ? 0.0 : gamma / gamma_prev;
    const double den = first ? delta : delta - beta * gamma / alpha_prev;
    const double alpha = (den > 0.0) ? gamma / den : 0.0;

    // ---- p = u + beta p,  s = w + beta s,  x += alpha p,  r -= alpha s ------------------------------------
    for (int e = tid; e < n_el; e += PT_THREADS) {
        const size_t i = (size_t)c0 * 6 + e;
        const double p = P.v_u[i] + beta * P.v_p[i];
        const double sv = sW[e] + beta * P.v_s[i];
        P.v_p[i] = p; P.v_s[i] = sv;
        P.v_x[i] += alpha * p;
        const double r = P.v_r[i] - alpha * sv;
        P.v_r[i] = r;
        sR[e] = r;
    }
    __syncthreads();
    double* sY = nullptr;  // coarse correction of the aggregates that meet this CTA's chunk, [aggregates][6]
    int a_first = 0;
    if (COARSE) {
        // every CTA's part of the new residual is in global memory (v_r); the cluster barrier makes it visible to all
        cl.sync();
        double* sRc = dsm + (size_t)chunk * 12;  // [nz] restricted residual
        sY = sRc + C.nz;
        coarse_restrict(C, P.v_r, P.n_cams, sRc);
        __syncthreads();
        a_first = c0 / C.agg;
        const int a_last = (c1 > c0) ? (c1 - 1) / C.agg : a_first - 1;
        const int rows = (a_last - a_first + 1) * 6;
        for (int rr = wid; rr < rows; rr += PT_THREADS / 32) {
            const double y = coarse_row(C, sRc, a_first * 6 + rr, lane);
            if (lane == 0) sY[rr] = y;
        }
        __syncthreads();
    }
    // ---- u = Minv r (6x6 block per camera) [+ coarse correction],  gamma = r.u -----------------------------
    part = 0.0;
    for (int e = tid; e < n_el; e += PT_THREADS) {
        const int cl_ = e / 6, k = e - 6 * cl_;
        const double* M = P.cam_Minv + ((size_t)(c0 + cl_) * 36 + 6 * k);
        const double* r = sR + 6 * cl_;
        double u = 0.0;
#pragma unroll
        for (int j = 0; j < 6; ++j) u += __ldg(M + j) * r[j];
        if (COARSE) u += sY[((c0 + cl_) / C.agg - a_first) * 6 + k];
        P.v_u[(size_t)c0 * 6 + e] = u;
        part += sR[e] * u;
    }
    const double gpart = cta_sum_all(part, red);
    if (tid == 0) cl.map_shared_rank(part_g, 0)[R] = gpart;
    cl.sync();
    if (R == 0 && tid == 0) {
        if (PEER) xstate[0] = (int)seq;  // read by every CTA at the start of the kernel, written after the last cluster barrier
        double g = 0.0;
        for (int r = 0; r < CL; ++r) g += part_g[r];
        st->gamma_prev = gamma;
        st->gamma = g; st->alpha = alpha; st->beta = beta;
        const int it = st->pcg_iter + 1;
        st->pcg_iter = it; st->pcg_total += 1;
        const bool done = !(alpha > 0.0) || !(g > tol * tol * st->gamma0) || it >= max_iters;
        if (done) st->pcg_done = 1;
        set_loop(pcg_loop, !done);
    }
}

// ---- coarse space: coloured probing, the inverse of the coarse matrix, the first preconditioned residual ---------------
//
// A0 = Z^T S Z couples two aggregates only if a point is seen from both.  adj[a * n_agg + b] = 1 marks such pairs (bytes,
// idempotent plain stores: no atomics, no order dependence).  The host colours the aggregates so that no aggregate is
// coupled to two of one colour (ba_solver.cu: coarse_colouring); one mat-vec then probes a whole colour at once.
__global__ void k_agg_adjacency(int n_pts, int n_cams, int agg, int n_agg, const int* __restrict__ pt_off,
                                const int2* __restrict__ pm_idx, unsigned char* __restrict__ adj) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pts) return;
    const int o0 = pt_off[p], o1 = pt_off[p + 1];
    for (int i = o0; i < o1; ++i) {
        const int ci = pm_idx[i].x;
        if (ci >= n_cams) continue;  // the fixed camera has no unknowns
        const int ai = ci / agg;
        for (int j = o0; j < o1; ++j) {
            const int cj = pm_idx[j].x;
            if (cj < n_cams) adj[(size_t)ai * n_agg + cj / agg] = 1;
        }
    }
}
// probe vector of (colour, component d): 1 in component d of every camera whose aggregate has that colour
__global__ void k_coarse_probe(int n_cams, int agg, const int* __restrict__ colour, int col, int d, double* __restrict__ probe) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (n_cams + 1) * 6) return;
    const int c = i / 6, e = i - 6 * c;
    probe[i] = (c < n_cams && e == d && colour[c / agg] == col) ? 1.0 : 0.0;
}
// w = S_local * probe(colour, d): the aggregate sums of w (camera order) are the entries A0[6 b + e][6 a + d] of the one
// aggregate a of that colour that b is coupled to (nb[b]; -1: none).
__global__ void k_coarse_scatter(int n_cams, int agg, int nz, int d, const int* __restrict__ nb, const double* __restrict__ w,
                                 double* __restrict__ A) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nz) return;
    const int b = t / 6, e = t - 6 * b;
    const int a = nb[b];
    if (a < 0) return;
    const int c0 = b * agg, c1 = min(c0 + agg, n_cams);
    double s = 0.0;
    for (int c = c0; c < c1; ++c) s += w[(size_t)c * 6 + e];
    A[(size_t)t * nz + 6 * a + d] = s;
}
// after the sum over the ranks: + lambda Z^T D Z (diagonal: the aggregates are disjoint) and symmetrise
// cam_U != nullptr (direct build: A holds the point part only): + Z^T U Z, the aggregate sums of the (replicated) camera blocks
__global__ void k_coarse_finish(int n_cams, int agg, int nz, const double* __restrict__ cam_D, const LMState* __restrict__ st,
                                const double* __restrict__ A, double* __restrict__ out, const double* __restrict__ cam_U = nullptr) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nz * nz) return;
    const int i = idx / nz, j = idx - i * nz;
    double v = 0.5 * (A[(size_t)i * nz + j] + A[(size_t)j * nz + i]);
    if (cam_U && i / 6 == j / 6) {
        const int a = i / 6, di = i - 6 * a, dj = j - 6 * a;
        const int c0 = a * agg, c1 = min(c0 + agg, n_cams);
        const int k = di <= dj ? sym6(di, dj) : sym6(dj, di);
        double s = 0.0;
        for (int c = c0; c < c1; ++c) s += cam_U[(size_t)c * 21 + k];
        v += s;
    }
    if (i == j) {
        const int a = i / 6, d = i - 6 * a;
        const int c0 = a * agg, c1 = min(c0 + agg, n_cams);
        double s = 0.0;
        for (int c = c0; c < c1; ++c) s += cam_D[(size_t)c * 6 + d];
        v += st->lambda * s;
    }
    out[idx] = v;
}

// ---- the coarse matrix without probing -----------------------------------------------------------------------------------
// A0 = Z^T S Z = Z^T U Z + lambda Z^T D Z - sum over the points of F_p^T (V_p + lambda D_p)^-1 F_p, with F_p = the point's E
// blocks added up per aggregate.  With local camera coordinates the camera block of an observation is Jc = Jp G_p,
// G_p = [-[X_p]x | I] (3 x 6), so E_cp^T = Jp^T W Jp G_p and the aggregate sum is F_{A,p} = W_{A,p} G_p, W_{A,p} = the sum of
// Jt^T Jt (Jacobian cache) over the point's observations from cameras of aggregate A: the point part of A0 needs nothing but
// the cache, the points and V^-1 - one pass over the point-major observation list instead of one mat-vec per probe vector
// (30 tile passes + 90 small kernels on the venice shape: 3.7 of the 51 ms of a solve).
//
// The blocks of all points meet in nz x nz sums.  To keep them bit-reproducible WITHOUT ordering the additions they are
// accumulated as 64-bit INTEGERS (integer addition is associative: any order, any mix of register, warp and L2 atomic adds
// gives the same bits): entry (i, j) is scaled by 2^58 / sqrt(d_i d_j), d = diag(Z^T U Z).  The Schur complement of a
// positive semi-definite matrix bounds every point's term, |t_ij| <= sqrt(u_ii u_jj) with u = the point's share of d, so by
// Cauchy-Schwarz the absolute sum over the points is at most sqrt(d_i d_j): the scaled sum stays below 2^58, and the
// rounding of a term is 2^-58 of that bound (18 k terms per entry on venice: 6e-14).
constexpr int CD_POOL = 64;        // (point, aggregate run) entries a warp holds in shared memory per batch
constexpr int CD_WARPS = 4;
constexpr double CD_SCALE = 536870912.0;  // 2^29 per factor: 2^58 per entry
__global__ void k_coarse_rscale(int n_cams, int agg, int nz, const double* __restrict__ cam_U, double* __restrict__ rscale) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nz) return;
    const int a = k / 6, i = k - 6 * a;
    const int c0 = a * agg, c1 = min(c0 + agg, n_cams);
    double d = 0.0;
    for (int c = c0; c < c1; ++c) d += cam_U[(size_t)c * 21 + sym6(i, i)];
    rscale[k] = (d > 0.0 && isfinite(d)) ? CD_SCALE / sqrt(d) : 0.0;
}
// A warp walks a contiguous range of points, 32 at a time: every lane turns the runs of equal aggregate in ITS point's track
// into entries F = W G (18 doubles) of the warp's pool, then the warp forms the 6 x 6 block of every ordered pair of entries
// of a point, lane = entry of the block (36 over two rounds).  Consecutive points mostly produce the same few aggregate pairs,
// so each lane keeps the running integer sums of four pairs in registers and sends one to L2 (RED.ADD.64) only when its pair
// is evicted.  overflow[0] = 1: a point with more than CD_POOL runs (the caller then probes instead).
__global__ void __launch_bounds__(CD_WARPS * 32) k_coarse_direct(Params P, int agg, int nz, const double* __restrict__ rscale,
                                                                  unsigned long long* __restrict__ acc, int* __restrict__ overflow) {
    __shared__ double sF[CD_WARPS][CD_POOL][18];
    __shared__ int sAgg[CD_WARPS][CD_POOL];
    __shared__ double sVi[CD_WARPS][32][6];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int b = P.st->cur;
    const int n_warps = gridDim.x * CD_WARPS, gw = blockIdx.x * CD_WARPS + w;
    const int per = ((P.n_pts + n_warps - 1) / n_warps + 31) & ~31;
    const int pa = min(gw * per, P.n_pts), pb = min(pa + per, P.n_pts);
    const double2* J = P.tc_J + (size_t)b * 3 * P.n_obs;
    // four cached pairs per lane: key (warp-uniform), two running sums (entries lane and lane + 32)
    int ckey[4] = {-1, -1, -1, -1};
    long long c0v[4] = {0, 0, 0, 0}, c1v[4] = {0, 0, 0, 0};
    int victim = 0;
    const int i0 = lane / 6, j0 = lane - 6 * i0;               // entry of round 0
    const int i1 = (lane + 32) / 6, j1 = (lane + 32) - 6 * i1;  // entry of round 1 (lanes 0-3)
    auto flush = [&](int k) {
        if (ckey[k] < 0) return;
        const int A = ckey[k] / (nz / 6), B = ckey[k] - A * (nz / 6);
        if (c0v[k]) atomicAdd(acc + (size_t)(A * 6 + i0) * nz + B * 6 + j0, (unsigned long long)c0v[k]);
        if (lane < 4 && c1v[k]) atomicAdd(acc + (size_t)(A * 6 + i1) * nz + B * 6 + j1, (unsigned long long)c1v[k]);
        ckey[k] = -1; c0v[k] = 0; c1v[k] = 0;
    };
    for (int p0 = pa; p0 < pb; p0 += 32) {
        const int p = p0 + lane;
        const bool valid = p < pb;
        const int o0 = valid ? P.pt_off[p] : 0, o1 = valid ? P.pt_off[p + 1] : 0;
        int runs = 0;
        {
            int last = -1;
            for (int j = o0; j < o1; ++j) {
                const int cam = P.pm_idx[j].x;
                if (cam >= P.n_cams) continue;  // the fixed camera has no unknowns
                const int A = cam / agg;
                if (A != last) { ++runs; last = A; }
            }
        }
        if (runs > CD_POOL) { *overflow = 1; runs = 0; }
        bool done = runs == 0;
        while (__any_sync(0xffffffffu, !done)) {
            const int need = done ? 0 : runs;
            int incl = need;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            const int first = incl - need;
            const bool take = !done && incl <= CD_POOL;  // a prefix of the lanes that are still waiting
            if (take) {
                double X[3], Vi[6];
                load_pt(P, b, p, X);
                load_sym3(P.pt_Vinv, p, Vi);
#pragma unroll
                for (int k = 0; k < 6; ++k) sVi[w][lane][k] = Vi[k];
                int e = first - 1, last = -1;
                double W[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
                auto emit = [&]() {  // F = W [-[X]x | I]: columns 0-2 = -W [X]x, columns 3-5 = W; column i scaled by rscale[6 A + i]
                    const double Wm[3][3] = {{W[0], W[1], W[2]}, {W[1], W[3], W[4]}, {W[2], W[4], W[5]}};
                    double* F = sF[w][e];
                    double rs[6];
#pragma unroll
                    for (int k = 0; k < 6; ++k) rs[k] = rscale[last * 6 + k];
#pragma unroll
                    for (int r = 0; r < 3; ++r) {
                        // row r of -W [X]x: (-[X]x) = [[0, X2, -X1], [-X2, 0, X0], [X1, -X0, 0]]
                        F[r * 6 + 0] = (-Wm[r][1] * X[2] + Wm[r][2] * X[1]) * rs[0];
                        F[r * 6 + 1] = (Wm[r][0] * X[2] - Wm[r][2] * X[0]) * rs[1];
                        F[r * 6 + 2] = (-Wm[r][0] * X[1] + Wm[r][1] * X[0]) * rs[2];
                        F[r * 6 + 3] = Wm[r][0] * rs[3]; F[r * 6 + 4] = Wm[r][1] * rs[4]; F[r * 6 + 5] = Wm[r][2] * rs[5];
                    }
                    sAgg[w][e] = last;
                };
                for (int j = o0; j < o1; ++j) {
                    const int cam = P.pm_idx[j].x;
                    if (cam >= P.n_cams) continue;
                    const int A = cam / agg;
                    if (A != last) {
                        if (last >= 0) emit();
                        ++e; last = A;
#pragma unroll
                        for (int k = 0; k < 6; ++k) W[k] = 0.0;
                    }
                    const int tc = P.pm_tc[j];
                    const double2 a0 = J[tc], a1 = J[(size_t)P.n_obs + tc], a2 = J[2 * (size_t)P.n_obs + tc];
                    const double jx[3] = {a0.x, a0.y, a1.x}, jy[3] = {a1.y, a2.x, a2.y};
                    W[0] += jx[0] * jx[0] + jy[0] * jy[0]; W[1] += jx[0] * jx[1] + jy[0] * jy[1]; W[2] += jx[0] * jx[2] + jy[0] * jy[2];
                    W[3] += jx[1] * jx[1] + jy[1] * jy[1]; W[4] += jx[1] * jx[2] + jy[1] * jy[2]; W[5] += jx[2] * jx[2] + jy[2] * jy[2];
                }
                if (last >= 0) emit();
            }
            __syncwarp();
            for (int q = 0; q < 32; ++q) {
                const int rq = __shfl_sync(0xffffffffu, take ? runs : 0, q);
                if (rq == 0) continue;
                const int fq = __shfl_sync(0xffffffffu, first, q);
                const double* Vi = sVi[w][q];
                const double v00 = Vi[0], v01 = Vi[1], v02 = Vi[2], v11 = Vi[3], v12 = Vi[4], v22 = Vi[5];
                // t(e, f)[i][j] = Fe[:, i] . V^-1 Ff[:, j] (already scaled).  Unordered pairs e <= f: the block of (f, e) is the
                // transpose.  Equal aggregates go to the diagonal key (the block, plus its transpose when e != f: a track that
                // returns to an aggregate); A_e < A_f to key (A_e, A_f); A_e > A_f (a track not sorted by camera) to key
                // (A_f, A_e) transposed.  k_coarse_from_acc mirrors the upper block triangle.
                auto tval = [&](const double* Fa, const double* Fb, int i, int j) {
                    const double f0 = Fa[i], f1 = Fa[6 + i], f2 = Fa[12 + i];
                    const double g0 = v00 * f0 + v01 * f1 + v02 * f2, g1 = v01 * f0 + v11 * f1 + v12 * f2, g2 = v02 * f0 + v12 * f1 + v22 * f2;
                    return g0 * Fb[j] + g1 * Fb[6 + j] + g2 * Fb[12 + j];
                };
                for (int e = fq; e < fq + rq; ++e) {
                    const double* Fe = sF[w][e];
                    const int Ae = sAgg[w][e];
                    for (int f = e; f < fq + rq; ++f) {
                        const double* Ff = sF[w][f];
                        const int Af = sAgg[w][f];
                        double t0, t1 = 0.0;
                        if (Ae <= Af) {  // warp-uniform
                            t0 = tval(Fe, Ff, i0, j0);
                            if (lane < 4) t1 = tval(Fe, Ff, i1, j1);
                            if (Ae == Af && e != f) {
                                t0 += tval(Fe, Ff, j0, i0);
                                if (lane < 4) t1 += tval(Fe, Ff, j1, i1);
                            }
                        } else {
                            t0 = tval(Fe, Ff, j0, i0);
                            if (lane < 4) t1 = tval(Fe, Ff, j1, i1);
                        }
                        const int key = min(Ae, Af) * (nz / 6) + max(Ae, Af);
                        const long long q0 = __double2ll_rn(t0);
                        const long long q1 = __double2ll_rn(t1);
                        int k = ckey[0] == key ? 0 : (ckey[1] == key ? 1 : (ckey[2] == key ? 2 : (ckey[3] == key ? 3 : -1)));
                        if (k < 0) {  // warp-uniform: evict round-robin
                            k = victim;
                            victim = (victim + 1) & 3;
#pragma unroll
                            for (int kk = 0; kk < 4; ++kk)
                                if (kk == k) { flush(kk); ckey[kk] = key; }
                        }
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk)
                            if (kk == k) { c0v[kk] += q0; c1v[kk] += q1; }
                    }
                }
            }
            __syncwarp();
            done |= take;
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) flush(k);
}
// point part of this rank's A0 as doubles (negative: it is subtracted), ready for the sum over the ranks
__global__ void k_coarse_from_acc(int nz, const double* __restrict__ rscale, const unsigned long long* __restrict__ acc,
                                  double* __restrict__ A, const int* __restrict__ overflow, double* __restrict__ overflow_out) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx == 0) overflow_out[0] = (double)overflow[0];  // as a double: it goes through the all-reduce (max) of a multi-GPU handle
    if (idx >= nz * nz) return;
    const int i = idx / nz, j = idx - i * nz;
    const double r = rscale[i] * rscale[j];
    // only the upper block triangle (aggregate of the row <= aggregate of the column) was accumulated
    const size_t src = (i / 6 <= j / 6) ? (size_t)idx : (size_t)j * nz + i;
    A[idx] = r > 0.0 ? -(double)(long long)acc[src] / r : 0.0;
}

// In-place inverse of the symmetric positive definite coarse matrix by Gauss-Jordan elimination without pivoting, on ONE
// thread-block cluster: CTA r keeps rows [r * rows_per, ...) in shared memory, the scaled pivot row of step k travels through
// distributed shared memory (two buffers alternate, so one cluster barrier per step suffices).  n^3 multiply-adds spread over
// the cluster, n barriers: ~0.5 ms for n = 336 against tens of milliseconds for a host factorisation plus two copies and a
// synchronisation.  flag[0] = 1 when a pivot is not positive (matrix not positive definite; the result is then unusable).
__host__ __device__ inline size_t spd_smem_bytes(int n, int cluster) {
    const int rows_per = (n + cluster - 1) / cluster;
    return ((size_t)rows_per * n + 3 * (size_t)n) * sizeof(double);
}
__global__ void __launch_bounds__(PT_THREADS, 1) k_spd_inverse(int n, const double* __restrict__ A, double* __restrict__ Ainv,
                                                                int* __restrict__ flag) {
    extern __shared__ __align__(16) double dsm[];
    __shared__ int s_bad;
    namespace cgx = cooperative_groups;
    cgx::cluster_group cl = cgx::this_cluster();
    const int R = (int)cl.block_rank(), CL = (int)cl.num_blocks();
    const int tid = threadIdx.x;
    const int rows_per = (n + CL - 1) / CL;
    const int r0 = min(R * rows_per, n), r1 = min(r0 + rows_per, n);
    const int nr = r1 - r0;
    double* M = dsm;                             // [rows_per][n] my rows
    double* prow = dsm + (size_t)rows_per * n;   // [2][n] pivot rows published by their owners
    double* lrow = prow + 2 * (size_t)n;         // [n] local copy of the current pivot row
    for (int l = tid; l < nr * n; l += PT_THREADS) M[l] = A[(size_t)r0 * n + l];
    if (tid == 0) s_bad = 0;
    __syncthreads();
    for (int k = 0; k < n; ++k) {
        const int owner = k / rows_per;
        double* mine = prow + (size_t)(k & 1) * n;
        if (R == owner) {
            double* row = M + (size_t)(k - r0) * n;
            const double piv = row[k];
            if (!(piv > 0.0) || !isfinite(piv)) { if (tid == 0) s_bad = 1; }
            const double ip = 1.0 / piv;
            __syncthreads();
            for (int j = tid; j < n; j += PT_THREADS) {
                const double v = (j == k) ? ip : row[j] * ip;
                row[j] = v;
                mine[j] = v;
            }
        }
        cl.sync();
        const double* pr = cl.map_shared_rank(prow, owner) + (size_t)(k & 1) * n;
        for (int j = tid; j < n; j += PT_THREADS) lrow[j] = pr[j];  // one remote read per element
        __syncthreads();
        // every row i != k of mine: f = a_ik; a_ij -= f * p_j (j != k); a_ik = -f * p_k
        for (int l = tid; l < nr * n; l += PT_THREADS) {
            const int i = l / n, j = l - i * n;
            if (r0 + i == k) continue;
            const double f = M[(size_t)i * n + k];
            // column k is read by every thread of the row in this step: it is rewritten after the barrier below
            if (j != k) M[l] -= f * lrow[j];
        }
        __syncthreads();
        {
            const double pk = lrow[k];
            for (int i = tid; i < nr; i += PT_THREADS)
                if (r0 + i != k) M[(size_t)i * n + k] = -M[(size_t)i * n + k] * pk;
        }
        __syncthreads();
    }
    cl.sync();  // nobody leaves while a peer may still read its pivot buffer
    for (int l = tid; l < nr * n; l += PT_THREADS) Ainv[(size_t)r0 * n + l] = M[l];
    if (tid == 0 && s_bad) atomicExch(flag, 1);
}
// The same elimination with 6 x 6 PIVOT BLOCKS (n = 6 x aggregates): one cluster barrier per aggregate instead of one per
// unknown - the scalar version spends its 1.1 ms (n = 336) on 336 barrier + remote-read round trips, not on arithmetic.
// Step K: the owner of block row K inverts its 6 x 6 pivot block (scalar Gauss-Jordan by one warp, the same pivots as the
// scalar elimination, so the positivity test is the same), replaces the block row by P^-1 [row] (P^-1 itself in block column
// K) and publishes the six rows; every CTA then subtracts F R from its other rows (F = their old block column K), block
// column K becoming -F P^-1.  Rows are dealt to the CTAs in multiples of 6.
__host__ __device__ inline int spd6_rows_per(int n, int cluster) { return 6 * ((n / 6 + cluster - 1) / cluster); }
__host__ __device__ inline size_t spd6_smem_bytes(int n, int cluster) {
    const int rp = spd6_rows_per(n, cluster);
    return ((size_t)rp * n + 18 * (size_t)n + (size_t)rp * 6 + 36) * sizeof(double);
}
__global__ void __launch_bounds__(PT_THREADS, 1) k_spd_inverse6(int n, const double* __restrict__ A, double* __restrict__ Ainv,
                                                                 int* __restrict__ flag) {
    extern __shared__ __align__(16) double dsm[];
    __shared__ int s_bad;
    namespace cgx = cooperative_groups;
    cgx::cluster_group cl = cgx::this_cluster();
    const int R = (int)cl.block_rank(), CL = (int)cl.num_blocks();
    const int tid = threadIdx.x, lane = tid & 31;
    const int rp = spd6_rows_per(n, CL);
    const int r0 = min(R * rp, n), r1 = min(r0 + rp, n);
    const int nr = r1 - r0;
    double* M = dsm;                          // [rp][n] my rows
    double* lrow = M + (size_t)rp * n;        // [2][6][n] the pivot rows of a step, PUSHED into every CTA by their owner (remote
                                              // stores before the cluster barrier: pulling them afterwards cost a distributed-
                                              // shared-memory load latency per step, 40 % of the kernel's stall samples)
    double* tmp = lrow + 12 * (size_t)n;      // [6][n] owner: the new block row before it replaces the old one
    double* fcol = tmp + 6 * (size_t)n;       // [rp][6] my rows' old block column K
    double* sP = fcol + (size_t)rp * 6;       // [36] the pivot block while it is inverted
    for (int l = tid; l < nr * n; l += PT_THREADS) M[l] = A[(size_t)r0 * n + l];
    if (tid == 0) s_bad = 0;
    __syncthreads();
    cl.sync();  // every CTA of the cluster is resident before the first remote store
    for (int K = 0; K < n / 6; ++K) {
        const int k0 = 6 * K, owner = k0 / rp;
        const size_t boff = (size_t)(K & 1) * 6 * n;  // a buffer is rewritten two steps later: everybody has passed a barrier since
        if (R == owner) {  // CTA-uniform
            const int lr = k0 - r0;
            if (tid < 36) sP[tid] = M[(size_t)(lr + tid / 6) * n + k0 + tid % 6];
            __syncthreads();
            if (tid < 32) {  // in-place inverse of the 6 x 6 block: lane e holds elements e and (lanes 0-3) e + 32
                for (int k = 0; k < 6; ++k) {
                    const double piv = sP[k * 6 + k];
                    if (lane == 0 && (!(piv > 0.0) || !isfinite(piv))) s_bad = 1;
                    const double ip = 1.0 / piv;
                    double nv[2] = {0.0, 0.0};
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        const int e = lane + 32 * hh;
                        if (e < 36) {
                            const int i = e / 6, j = e - 6 * i;
                            const double pj = (j == k) ? ip : sP[k * 6 + j] * ip;  // element j of the scaled pivot row
                            if (i == k) nv[hh] = pj;
                            else {
                                const double f = sP[i * 6 + k];
                                nv[hh] = (j == k) ? -f * ip : sP[e] - f * pj;
                            }
                        }
                    }
                    __syncwarp();
                    sP[lane] = nv[0];
                    if (lane < 4) sP[lane + 32] = nv[1];
                    __syncwarp();
                }
            }
            __syncthreads();
            for (int l = tid; l < 6 * n; l += PT_THREADS) {  // the new block row: P^-1 [old rows]; P^-1 in block column K
                const int r = l / n, j = l - r * n;
                double v = 0.0;
                if (j >= k0 && j < k0 + 6) v = sP[r * 6 + (j - k0)];
                else
#pragma unroll
                    for (int c = 0; c < 6; ++c) v += sP[r * 6 + c] * M[(size_t)(lr + c) * n + j];
                tmp[l] = v;
            }
            __syncthreads();  // the old rows have all been read
            for (int l = tid; l < 6 * n; l += PT_THREADS) {
                const double v = tmp[l];
                M[(size_t)lr * n + l] = v;
                for (int d = 0; d < CL; ++d) cl.map_shared_rank(lrow, d)[boff + l] = v;
            }
        }
        cl.sync();  // release / acquire: the owner's remote stores are visible
        const double* pr = lrow + boff;
        for (int l = tid; l < nr * 6; l += PT_THREADS) fcol[l] = M[(size_t)(l / 6) * n + k0 + (l % 6)];
        __syncthreads();
        for (int l = tid; l < nr * n; l += PT_THREADS) {
            const int i = l / n, j = l - i * n;
            if (r0 + i >= k0 && r0 + i < k0 + 6) continue;  // the pivot rows are done
            const double* f = fcol + 6 * i;
            double sacc = 0.0;
#pragma unroll
            for (int r = 0; r < 6; ++r) sacc += f[r] * pr[(size_t)r * n + j];
            M[l] = (j >= k0 && j < k0 + 6) ? -sacc : M[l] - sacc;
        }
        __syncthreads();
    }
    cl.sync();  // nobody leaves while a peer may still store into its pivot buffers
    for (int l = tid; l < nr * n; l += PT_THREADS) Ainv[(size_t)r0 * n + l] = M[l];
    if (tid == 0 && s_bad) atomicExch(flag, 1);
}
// After k_camera_finish_schur (u = Minv r, gamma = r.u): add the coarse correction to u and redo gamma.  One CTA.
__global__ void __launch_bounds__(PT_THREADS) k_coarse_first(Params P, Coarse C, unsigned long long pcg_loop) {
    extern __shared__ __align__(16) double sy[];  // [nz] correction, [nz] restricted residual
    __shared__ double red[33];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    double* s_rc = sy + C.nz;
    coarse_restrict(C, P.v_r, P.n_cams, s_rc);
    __syncthreads();
    for (int row = wid; row < C.nz; row += PT_THREADS / 32) {
        const double y = coarse_row(C, s_rc, row, lane);
        if (lane == 0) sy[row] = y;
    }
    __syncthreads();
    double part = 0.0;
    for (int i = tid; i < P.n_cams * 6; i += PT_THREADS) {
        const int c = i / 6, d = i - 6 * c;
        const double u = P.v_u[i] + sy[(c / C.agg) * 6 + d];
        P.v_u[i] = u;
        part += P.v_r[i] * u;
    }
    const double g = cta_sum_all(part, red);
    if (tid == 0) {
        LMState* st = P.st;
        st->gamma = g; st->gamma0 = g; st->gamma_prev = g;
        st->pcg_done = !(g > 0.0);
        set_loop(pcg_loop, g > 0.0);
    }
}

}  // namespace ba
