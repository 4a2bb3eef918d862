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
// Two slots alternate by sequence parity: a rank can only be one exchange ahead of the slowest rank
// (exchange n+1 needs everybody's flag n+1, which a rank sets after it has finished reading n), so slot
// (n & 1) is never rewritten while a peer still reads exchange n.
//
// Exchange buffer of a rank (doubles): [0] flag of slot 0, [1] flag of slot 1 (sequence numbers stored as
// doubles' bit patterns are avoided: flags are 64-bit integers in the same words), [16 ...) slot 0 payload,
// [16 + cap ...) slot 1 payload.
#pragma once
#include "ba_tiles.cuh"

namespace ba {

constexpr int XB_HDR = 16;                 // doubles reserved in front of the payloads (flags, padding to 128 B)
constexpr long long XB_SPIN_LIMIT = 1LL << 28;  // polls before a rank gives up on a peer (~ seconds) and flags an error

BA_D void st_release_sys(long long* p, long long v) {
    asm volatile("st.release.sys.global.s64 [%0], %1;\n" ::"l"(p), "l"(v) : "memory");
}
BA_D long long ld_acquire_sys(const long long* p) {
    long long v;
    asm volatile("ld.acquire.sys.global.s64 %0, [%1];\n" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// xstate: [0] sequence number of the last exchange this rank published, [1] ticket, [2] error flag
__global__ void __launch_bounds__(RS_WARPS * 32) k_rows_finish_xchg(Params P, double* __restrict__ xbuf, int cap,
                                                                    int* __restrict__ xstate) {
    __shared__ int last;
    if (P.st->pcg_done) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int c = blockIdx.x * RS_WARPS + wid;
    const long long seq = (long long)xstate[0] + 1;  // only the last CTA changes xstate[0], after everybody has read it
    double* slot = xbuf + XB_HDR + (size_t)(seq & 1) * cap;
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
            double2* o = reinterpret_cast<double2*>(slot + (size_t)c * 6);
            o[0] = make_double2(a[0], a[1]); o[1] = make_double2(a[2], a[3]); o[2] = make_double2(a[4], a[5]);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();  // this CTA's payload before its ticket
        last = (atomicAdd(xstate + 1, 1) == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence_system();
        xstate[1] = 0;
        xstate[0] = (int)seq;
        st_release_sys(reinterpret_cast<long long*>(xbuf) + (seq & 1), seq);  // publish
    }
}

// peers[r] = exchange buffer of rank r as mapped into this process (peers[own rank] = the local buffer)
__global__ void __launch_bounds__(VEC_THREADS) k_pcg_dots_peer(Params P, double* const* __restrict__ peers, int nranks,
                                                               int cap, int* __restrict__ xstate) {
    __shared__ double sm[VEC_THREADS / 32];
    __shared__ int ok;
    if (P.st->pcg_done) return;
    const long long seq = xstate[0];
    const size_t slot_off = XB_HDR + (size_t)(seq & 1) * cap;
    if (threadIdx.x == 0) {
        int good = 1;
        for (int r = 0; r < nranks && good; ++r) {
            const long long* flag = reinterpret_cast<const long long*>(peers[r]) + (seq & 1);
            long long spins = 0;
            while (ld_acquire_sys(flag) < seq) {
                if (++spins > XB_SPIN_LIMIT) { good = 0; break; }
            }
        }
        if (!good) atomicExch(xstate + 2, 1);
        ok = good;
    }
    __syncthreads();
    const int c = blockIdx.x * VEC_THREADS + threadIdx.x;
    double part = 0.0;
    if (c < P.n_cams && ok) {
        const double lam = P.st->lambda;
        double w[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        for (int r = 0; r < nranks; ++r) {  // rank order: identical bits on every rank
            const double2* q = reinterpret_cast<const double2*>(peers[r] + slot_off + (size_t)c * 6);
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

}  // namespace ba
