// ba_map.cuh - the device-resident map of successive local-BA windows.
//
// The reference flattens the window of every run() call from Python objects (_gather_local_data,
// /root/reference/src/bundle_adjuster.py:195-218, and the packing at :157-162).  Here the observation rows of every
// keyframe (keyframe id, map point id, keypoint) and the positions of the map points stay in HBM between calls; a call
// uploads what is new since the last one (the pipeline only appends: /root/reference/src/pipeline.py:172-173, 215-217,
// 277-278, 306-308) and these kernels build the flat problem of the window - the reference's row order and point
// numbering - without a host pass:
//   * rows of the window's keyframes whose map point is known, keyframe-major in window order and in list order inside
//     a keyframe (:210-216): one stable radix sort by window slot of the resident rows (append order = list order);
//   * point numbering = rank of the id among the window's ids (sorted(...) at :218): flags over the id range + scan;
//   * a (keyframe, map point) pair listed twice keeps both rows but the LAST keypoint (the dict at :216): the index of
//     the last row of every pair, by an integer atomicMax (order-independent), then one gather.
// Tables are dense in the ids (the reference numbers keyframes and map points from counters).
#pragma once
#include <cuda_runtime.h>

namespace ba {

// slot_of[kf_id[s]] = s for the window's keyframes (slot_of is -1 elsewhere)
__global__ void k_mw_slots(int n_kfs, const int* __restrict__ kf_id, int* __restrict__ slot_of) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n_kfs) slot_of[kf_id[s]] = s;
}

// key = window slot of the row's keyframe, or n_kfs when the row is not part of the window (other keyframe, or a map
// point that is not in the map: `if mp_id in gmap.map_points`, :213).  Marks the ids of the window's points.
__global__ void k_mw_keys(int n_rows, int n_kfs, int kf_cap, int mp_cap, const int* __restrict__ row_kf, const int* __restrict__ row_mp,
                          const int* __restrict__ slot_of, const unsigned char* __restrict__ mp_known, unsigned* __restrict__ key,
                          int* __restrict__ val, int* __restrict__ mp_flag, int* __restrict__ n_sel) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    const int kf = row_kf[r], mp = row_mp[r];
    const int s = (kf >= 0 && kf < kf_cap) ? slot_of[kf] : -1;
    const bool ok = s >= 0 && mp >= 0 && mp < mp_cap && mp_known[mp];
    key[r] = ok ? (unsigned)s : (unsigned)n_kfs;
    val[r] = r;
    if (ok) {
        mp_flag[mp] = 1;  // every writer stores the same value
        atomicAdd(n_sel, 1);
    }
}

// ids and positions of the window's points, in id order (local index = exclusive scan of the flags)
__global__ void k_mw_points(int mp_cap, const int* __restrict__ mp_flag, const int* __restrict__ mp_local, const double* __restrict__ mp_xyz,
                            int* __restrict__ ids_out, double* __restrict__ pts_out) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= mp_cap || !mp_flag[id]) return;
    const int p = mp_local[id];
    ids_out[p] = id;
    for (int k = 0; k < 3; ++k) pts_out[3 * (size_t)p + k] = mp_xyz[3 * (size_t)id + k];
}

// last[(slot, point)] = largest sorted row index of the pair
__global__ void k_mw_last(int n_obs, int n_pts, const unsigned* __restrict__ key_s, const int* __restrict__ order, const int* __restrict__ row_mp,
                          const int* __restrict__ mp_local, int* __restrict__ last) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    const int p = mp_local[row_mp[order[i]]];
    atomicMax(last + (size_t)key_s[i] * n_pts + p, i);
}

// the flat problem: slot 0 is the fixed keyframe (camera -1), slot s > 0 adjustable camera s - 1
__global__ void k_mw_fill(int n_obs, int n_pts, const unsigned* __restrict__ key_s, const int* __restrict__ order, const int* __restrict__ row_mp,
                          const double2* __restrict__ row_uv, const int* __restrict__ mp_local, const int* __restrict__ last,
                          int* __restrict__ obs_cam, int* __restrict__ obs_pt, double2* __restrict__ obs_uv) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    const int s = (int)key_s[i];
    const int p = mp_local[row_mp[order[i]]];
    obs_cam[i] = s - 1;
    obs_pt[i] = p;
    obs_uv[i] = row_uv[order[last[(size_t)s * n_pts + p]]];
}

// upsert of map point positions
__global__ void k_map_set_points(int n, const int* __restrict__ id, const double* __restrict__ xyz, double* __restrict__ mp_xyz,
                                 unsigned char* __restrict__ mp_known) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t m = (size_t)id[i];
    for (int k = 0; k < 3; ++k) mp_xyz[3 * m + k] = xyz[3 * (size_t)i + k];
    mp_known[m] = 1;
}
__global__ void k_map_forget_points(int n, int mp_cap, const int* __restrict__ id, unsigned char* __restrict__ mp_known) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && id[i] >= 0 && id[i] < mp_cap) mp_known[id[i]] = 0;
}
__global__ void k_map_fill_rows(int n, int kf, int* __restrict__ row_kf) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) row_kf[i] = kf;
}

// after an accepted solve: the solver's current points (internal order q, caller's index perm[q] = window-local index)
// go back into the resident table
__global__ void k_map_commit(int n_pts, int n_pad, const int* __restrict__ perm, const double* __restrict__ base, const int* __restrict__ win_ids,
                             double* __restrict__ mp_xyz) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_pts) return;
    const size_t m = (size_t)win_ids[perm[q]];
    const double2 a = reinterpret_cast<const double2*>(base)[q];
    mp_xyz[3 * m] = a.x;
    mp_xyz[3 * m + 1] = a.y;
    mp_xyz[3 * m + 2] = base[2 * (size_t)n_pad + q];
}

}  // namespace ba
