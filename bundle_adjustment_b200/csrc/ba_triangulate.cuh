// ba_triangulate.cuh - batched two-view linear triangulation and the cheirality test on the device
// (the producer of new map points, reference src/pipeline.py:315-336).
//
// The reference calls cv2.triangulatePoints(P1, P2, pts1, pts2) (OpenCV, third-party, unpinned; 4.13 here):
// per correspondence the 4x4 matrix
//        A = [ x1 P1[2] - P1[0] ;  y1 P1[2] - P1[1] ;  x2 P2[2] - P2[0] ;  y2 P2[2] - P2[1] ]
// and the right singular vector of its smallest singular value (homogeneous point, unit norm, sign
// arbitrary).  One thread per correspondence: one-sided (Hestenes) Jacobi SVD of A in registers - plane
// rotations make the columns of A orthogonal, V collects the rotations, the shortest column marks the
// singular vector.  No squaring of A, so the small singular value keeps its relative accuracy.
//
// The pipeline then forms  X = h[:3] / (h[3] + 1e-6)  and keeps points with z > 0 in both views
// (camera 2: R_rel X + t_rel).  k_triangulate does the same when asked to (`cheirality`).
#pragma once
#include "ba_kernels.cuh"

namespace ba {

// Smallest right singular vector of the row-major 4x4 matrix A -> h[4] (unit norm; sign: largest component positive is
// NOT enforced - callers divide by h[3]).
BA_D void smallest_singular_vector4(double A[16], double h[4]) {
    double V[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    for (int sweep = 0; sweep < 40; ++sweep) {
        bool changed = false;
#pragma unroll
        for (int p = 0; p < 3; ++p)
#pragma unroll
            for (int q = p + 1; q < 4; ++q) {
                double al = 0.0, be = 0.0, ga = 0.0;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    al += A[4 * i + p] * A[4 * i + p];
                    be += A[4 * i + q] * A[4 * i + q];
                    ga += A[4 * i + p] * A[4 * i + q];
                }
                if (fabs(ga) <= 1e-15 * sqrt(al * be) || ga == 0.0) continue;
                changed = true;
                const double zeta = (be - al) / (2.0 * ga);
                const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const double ap = A[4 * i + p], aq = A[4 * i + q];
                    A[4 * i + p] = c * ap - s * aq;
                    A[4 * i + q] = s * ap + c * aq;
                    const double vp = V[4 * i + p], vq = V[4 * i + q];
                    V[4 * i + p] = c * vp - s * vq;
                    V[4 * i + q] = s * vp + c * vq;
                }
            }
        if (!changed) break;
    }
    int best = 0;
    double nb = 0.0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        double n = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) n += A[4 * i + j] * A[4 * i + j];
        if (j == 0 || n < nb) { nb = n; best = j; }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {  // V[:, best] without dynamic register indexing
        double v = V[4 * i];
#pragma unroll
        for (int j = 1; j < 4; ++j) v = (best == j) ? V[4 * i + j] : v;
        h[i] = v;
    }
}

struct TriParams {
    double P1[12], P2[12];  // row-major 3x4 projection matrices
    double R[9], t[3];      // relative pose of camera 2 (cheirality test)
};

// pts1/pts2 [n][2];  hom [n][4] homogeneous solutions (may be null);  xyz [n][3] = hom[:3] / (hom[3] + 1e-6) and
// valid [n] = z > 0 in both views (both may be null)
__global__ void k_triangulate(TriParams T, int n, const double* __restrict__ pts1, const double* __restrict__ pts2,
                              double* __restrict__ hom, double* __restrict__ xyz, unsigned char* __restrict__ valid) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double x1 = pts1[2 * i], y1 = pts1[2 * i + 1], x2 = pts2[2 * i], y2 = pts2[2 * i + 1];
    double A[16], h[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        A[k] = x1 * T.P1[8 + k] - T.P1[k];
        A[4 + k] = y1 * T.P1[8 + k] - T.P1[4 + k];
        A[8 + k] = x2 * T.P2[8 + k] - T.P2[k];
        A[12 + k] = y2 * T.P2[8 + k] - T.P2[4 + k];
    }
    smallest_singular_vector4(A, h);
    if (hom) {
#pragma unroll
        for (int k = 0; k < 4; ++k) hom[4 * (size_t)i + k] = h[k];
    }
    if (xyz || valid) {
        const double w = h[3] + 1e-6;  // as the reference divides (src/pipeline.py:324)
        const double X = h[0] / w, Y = h[1] / w, Z = h[2] / w;
        if (xyz) { xyz[3 * (size_t)i] = X; xyz[3 * (size_t)i + 1] = Y; xyz[3 * (size_t)i + 2] = Z; }
        if (valid) {
            const double z2 = T.R[6] * X + T.R[7] * Y + T.R[8] * Z + T.t[2];
            valid[i] = (Z > 0.0 && z2 > 0.0) ? 1 : 0;
        }
    }
}

}  // namespace ba
