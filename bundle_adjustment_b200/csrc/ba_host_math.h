// ba_host_math.h - small dense linear algebra on the host (plain C++, no CUDA): the inverse of the coarse matrix of the
// two-level preconditioner (a few hundred unknowns, once per solve).  Tested on the CPU by tests/test_host_logic.py, which
// compiles this header with g++.
#pragma once
#include <cmath>
#include <cstddef>
#include <vector>

namespace ba {

// A (n x n, row-major, nearly symmetric: it is symmetrised first) -> Ainv = A^-1 through A = L L^T, A^-1 = L^-T L^-1.
// Returns false (Ainv unspecified) when A is not positive definite.  A is overwritten with L in its lower triangle.
inline bool spd_inverse_host(std::vector<double>& A, int n, std::vector<double>& Ainv) {
    const size_t N = (size_t)n;
    for (size_t i = 0; i < N; ++i)
        for (size_t j = 0; j < i; ++j) A[i * N + j] = A[j * N + i] = 0.5 * (A[i * N + j] + A[j * N + i]);
    for (size_t j = 0; j < N; ++j) {
        double d = A[j * N + j];
        for (size_t k = 0; k < j; ++k) d -= A[j * N + k] * A[j * N + k];
        if (!(d > 0.0) || !std::isfinite(d)) return false;
        const double l = std::sqrt(d);
        A[j * N + j] = l;
        for (size_t i = j + 1; i < N; ++i) {
            double v = A[i * N + j];
            for (size_t k = 0; k < j; ++k) v -= A[i * N + k] * A[j * N + k];
            A[i * N + j] = v / l;
        }
    }
    std::vector<double> Li(N * N, 0.0);  // L^-1 (lower triangular), column by column
    for (size_t c = 0; c < N; ++c) {
        Li[c * N + c] = 1.0 / A[c * N + c];
        for (size_t i = c + 1; i < N; ++i) {
            double v = 0.0;
            for (size_t k = c; k < i; ++k) v -= A[i * N + k] * Li[k * N + c];
            Li[i * N + c] = v / A[i * N + i];
        }
    }
    Ainv.assign(N * N, 0.0);
    for (size_t i = 0; i < N; ++i)
        for (size_t j = 0; j <= i; ++j) {
            double v = 0.0;
            for (size_t k = i; k < N; ++k) v += Li[k * N + i] * Li[k * N + j];
            Ainv[i * N + j] = Ainv[j * N + i] = v;
        }
    return true;
}

}  // namespace ba
