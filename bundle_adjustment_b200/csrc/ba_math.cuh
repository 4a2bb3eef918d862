// ba_math.cuh - per-observation / per-camera / per-point FP64 arithmetic of the B200 bundle adjuster.
//
// Everything here is small dense FP64 (2x3, 3x3, 6x6): no tensor cores, registers only.
// Camera model = the reference's (src/bundle_adjuster.py:67): x_cam = R(rvec) X + t,
// u = fx x/z + cx, v = fy y/z + cy; residual = observed - projected (src/bundle_adjuster.py:68-70).
#pragma once
#include <cuda_runtime.h>
#include <float.h>
#include <math.h>

namespace ba {

#define BA_HD __host__ __device__ __forceinline__
#define BA_D __device__ __forceinline__

// ---------------------------------------------------------------------------------------------
// SO(3)
// ---------------------------------------------------------------------------------------------

// rvec -> row-major R, the formula cv2.Rodrigues uses (identity below theta = DBL_EPSILON).
BA_HD void rodrigues(const double r[3], double R[9]) {
    const double th = sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
    if (th < DBL_EPSILON) {
        R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
        return;
    }
    const double c = cos(th), s = sin(th), c1 = 1.0 - c, it = 1.0 / th;
    const double kx = r[0] * it, ky = r[1] * it, kz = r[2] * it;
    R[0] = c + c1 * kx * kx;      R[1] = c1 * kx * ky - s * kz; R[2] = c1 * kx * kz + s * ky;
    R[3] = c1 * kx * ky + s * kz; R[4] = c + c1 * ky * ky;      R[5] = c1 * ky * kz - s * kx;
    R[6] = c1 * kx * kz - s * ky; R[7] = c1 * ky * kz + s * kx; R[8] = c + c1 * kz * kz;
}

// Right Jacobian of SO(3): R(r + dr) = R(r) Exp(Jr(r) dr) + O(|dr|^2).  Jr = I - B [r]x + C [r]x^2,
// B = (1 - cos th)/th^2, C = (th - sin th)/th^3 (series below th^2 = 1e-2).
BA_HD void right_jacobian(const double r[3], double J[9]) {
    const double t2 = r[0] * r[0] + r[1] * r[1] + r[2] * r[2];
    double B, C;
    if (t2 < 1e-2) {
        B = 0.5 - t2 * (1.0 / 24.0 - t2 * (1.0 / 720.0 - t2 * (1.0 / 40320.0)));
        C = 1.0 / 6.0 - t2 * (1.0 / 120.0 - t2 * (1.0 / 5040.0 - t2 * (1.0 / 362880.0)));
    } else {
        const double th = sqrt(t2), sh = sin(0.5 * th);
        B = 2.0 * sh * sh / t2;
        C = (th - sin(th)) / (t2 * th);
    }
    const double x = r[0], y = r[1], z = r[2];
    // [r]x^2 = r r^T - |r|^2 I
    J[0] = 1.0 + C * (x * x - t2); J[1] = B * z + C * x * y;      J[2] = -B * y + C * x * z;
    J[3] = -B * z + C * x * y;     J[4] = 1.0 + C * (y * y - t2); J[5] = B * x + C * y * z;
    J[6] = B * y + C * x * z;      J[7] = -B * x + C * y * z;     J[8] = 1.0 + C * (z * z - t2);
}

// Row-major R -> rvec through the unit quaternion (stable for every angle below pi).
BA_HD void log_so3(const double R[9], double r[3]) {
    double w, x, y, z;
    const double tr = R[0] + R[4] + R[8];
    if (tr > 0.0) {
        const double s = sqrt(tr + 1.0) * 2.0;
        w = 0.25 * s; x = (R[7] - R[5]) / s; y = (R[2] - R[6]) / s; z = (R[3] - R[1]) / s;
    } else if (R[0] > R[4] && R[0] > R[8]) {
        const double s = sqrt(1.0 + R[0] - R[4] - R[8]) * 2.0;
        w = (R[7] - R[5]) / s; x = 0.25 * s; y = (R[1] + R[3]) / s; z = (R[2] + R[6]) / s;
    } else if (R[4] > R[8]) {
        const double s = sqrt(1.0 + R[4] - R[0] - R[8]) * 2.0;
        w = (R[2] - R[6]) / s; x = (R[1] + R[3]) / s; y = 0.25 * s; z = (R[5] + R[7]) / s;
    } else {
        const double s = sqrt(1.0 + R[8] - R[0] - R[4]) * 2.0;
        w = (R[3] - R[1]) / s; x = (R[2] + R[6]) / s; y = (R[5] + R[7]) / s; z = 0.25 * s;
    }
    if (w < 0.0) { w = -w; x = -x; y = -y; z = -z; }
    const double n = sqrt(x * x + y * y + z * z);
    const double k = (n < 1e-12) ? 2.0 : 2.0 * atan2(n, w) / n;
    r[0] = k * x; r[1] = k * y; r[2] = k * z;
}

BA_HD void mat3_mul(const double A[9], const double B[9], double C[9]) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}

// ---------------------------------------------------------------------------------------------
// robust loss: Huber on each scalar residual (scipy least_squares loss='huber', f_scale = delta)
// ---------------------------------------------------------------------------------------------
enum { ROBUST_NONE = 0, ROBUST_IRLS = 1, ROBUST_TRIGGS = 2 };

// 1 / d in the hot loops (device): hardware seed (rcp.approx.ftz.f64, about 20 bits) and two Newton steps - five instructions
// without a branch, accurate to an ulp or two for normal d, against the library division's call with a slow path (three of
// those per observation were a third of the camera-major passes' FP64 instructions).  d = 0, inf, nan give nan (the division
// gives inf / 0 / nan): every caller treats non-finite results alike.
BA_HD double fast_rcp(double d) {
#ifdef __CUDA_ARCH__
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    r = fma(r, fma(-d, r, 1.0), r);
    r = fma(r, fma(-d, r, 1.0), r);
    return r;
#else
    return 1.0 / d;
#endif
}

// rho0 = delta^2 * rho((f/delta)^2); wg = rho' (gradient weight); wH = curvature weight.
BA_HD void robust_weights(double f, double delta, int mode, double& wH, double& wg, double& rho0) {
    const double a = fabs(f);
    if (mode == ROBUST_NONE || a <= delta) {
        wH = 1.0; wg = 1.0; rho0 = f * f;
    } else {
        wg = delta * fast_rcp(a);
        rho0 = delta * (2.0 * a - delta);
        wH = (mode == ROBUST_IRLS) ? wg : DBL_EPSILON;
    }
}

// ---------------------------------------------------------------------------------------------
// one observation
// ---------------------------------------------------------------------------------------------
struct Intr { double fx, fy, cx, cy; };

struct ObsLin {
    double rx, ry;        // residual (observed - projected)
    double jx[3], jy[3];  // rows of d r / d X  (= -d pi/d xc * R)
    double zc;            // depth in the camera
};

// Rt = R (9, row-major) followed by t (3).
BA_D void project_residual(const double* Rt, const double X[3], double u, double v, const Intr& K, double& rx,
                           double& ry) {
    const double xc = fma(Rt[0], X[0], fma(Rt[1], X[1], fma(Rt[2], X[2], Rt[9])));
    const double yc = fma(Rt[3], X[0], fma(Rt[4], X[1], fma(Rt[5], X[2], Rt[10])));
    const double zc = fma(Rt[6], X[0], fma(Rt[7], X[1], fma(Rt[8], X[2], Rt[11])));
    rx = u - (K.fx * xc / zc + K.cx);
    ry = v - (K.fy * yc / zc + K.cy);
}

BA_D void linearize_obs(const double* Rt, const double X[3], double u, double v, const Intr& K, ObsLin& o) {
    const double xc = fma(Rt[0], X[0], fma(Rt[1], X[1], fma(Rt[2], X[2], Rt[9])));
    const double yc = fma(Rt[3], X[0], fma(Rt[4], X[1], fma(Rt[5], X[2], Rt[10])));
    const double zc = fma(Rt[6], X[0], fma(Rt[7], X[1], fma(Rt[8], X[2], Rt[11])));
    const double iz = fast_rcp(zc);
    const double un = xc * iz, vn = yc * iz;
    o.zc = zc;
    o.rx = u - fma(K.fx, un, K.cx);
    o.ry = v - fma(K.fy, vn, K.cy);
    const double sx = -K.fx * iz, sy = -K.fy * iz;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        o.jx[k] = sx * fma(-un, Rt[6 + k], Rt[k]);
        o.jy[k] = sy * fma(-vn, Rt[6 + k], Rt[3 + k]);
    }
}

// rows of d r / d X only (what the Schur mat-vec needs: no measurement, no residual)
BA_D void point_jacobian(const double* Rt, const double X[3], const Intr& K, double jx[3], double jy[3]) {
    const double xc = fma(Rt[0], X[0], fma(Rt[1], X[1], fma(Rt[2], X[2], Rt[9])));
    const double yc = fma(Rt[3], X[0], fma(Rt[4], X[1], fma(Rt[5], X[2], Rt[10])));
    const double zc = fma(Rt[6], X[0], fma(Rt[7], X[1], fma(Rt[8], X[2], Rt[11])));
    const double iz = fast_rcp(zc);
    const double un = xc * iz, vn = yc * iz;
    const double sx = -K.fx * iz, sy = -K.fy * iz;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        jx[k] = sx * fma(-un, Rt[6 + k], Rt[k]);
        jy[k] = sy * fma(-vn, Rt[6 + k], Rt[3 + k]);
    }
}

BA_HD void cross3(const double a[3], const double b[3], double c[3]) {
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}

// index of (i,j), i <= j, in the row-major upper triangle of a symmetric NxN matrix
BA_HD constexpr int sym6(int i, int j) { return i * 6 - i * (i - 1) / 2 + (j - i); }
BA_HD constexpr int sym3(int i, int j) { return i * 3 - i * (i - 1) / 2 + (j - i); }

// symmetric 3x3 (6 upper entries: 00 01 02 11 12 22) times vector
BA_HD void sym3_mulv(const double S[6], const double x[3], double y[3]) {
    y[0] = S[0] * x[0] + S[1] * x[1] + S[2] * x[2];
    y[1] = S[1] * x[0] + S[3] * x[1] + S[4] * x[2];
    y[2] = S[2] * x[0] + S[4] * x[1] + S[5] * x[2];
}

// inverse of an SPD symmetric 3x3 through its LDL^T factors; returns false if a pivot is not positive.
BA_HD bool sym3_inverse(const double A[6], double Ai[6]) {
    const double d0 = A[0];
    if (!(d0 > 0.0)) return false;
    const double i0 = fast_rcp(d0);
    const double l10 = A[1] * i0, l20 = A[2] * i0;
    const double d1 = A[3] - l10 * A[1];
    if (!(d1 > 0.0)) return false;
    const double i1 = fast_rcp(d1);
    const double l21 = (A[4] - l20 * A[1]) * i1;
    const double d2 = A[5] - l20 * A[2] - l21 * l21 * d1;
    if (!(d2 > 0.0)) return false;
    const double i2 = fast_rcp(d2);
    // L^-1 = [[1,0,0],[-l10,1,0],[l10*l21-l20,-l21,1]]
    const double m10 = -l10, m20 = l10 * l21 - l20, m21 = -l21;
    Ai[0] = i0 + m10 * m10 * i1 + m20 * m20 * i2;
    Ai[1] = m10 * i1 + m20 * m21 * i2;
    Ai[2] = m20 * i2;
    Ai[3] = i1 + m21 * m21 * i2;
    Ai[4] = m21 * i2;
    Ai[5] = i2;
    return true;
}

// In-place inverse of an SPD 6x6 (full row-major storage) by Cholesky; returns false if not SPD.
BA_HD bool spd6_inverse(double A[36]) {
    double L[36];
#pragma unroll
    for (int i = 0; i < 36; ++i) L[i] = 0.0;
    for (int j = 0; j < 6; ++j) {
        double d = A[7 * j];
        for (int k = 0; k < j; ++k) d -= L[6 * j + k] * L[6 * j + k];
        if (!(d > 0.0)) return false;
        const double ljj = sqrt(d);
        L[7 * j] = ljj;
        for (int i = j + 1; i < 6; ++i) {
            double s = A[6 * i + j];
            for (int k = 0; k < j; ++k) s -= L[6 * i + k] * L[6 * j + k];
            L[6 * i + j] = s / ljj;
        }
    }
    // Li = L^-1 (lower triangular), then A^-1 = Li^T Li
    double Li[36];
#pragma unroll
    for (int i = 0; i < 36; ++i) Li[i] = 0.0;
    for (int j = 0; j < 6; ++j) {
        Li[7 * j] = 1.0 / L[7 * j];
        for (int i = j + 1; i < 6; ++i) {
            double s = 0.0;
            for (int k = j; k < i; ++k) s -= L[6 * i + k] * Li[6 * k + j];
            Li[6 * i + j] = s / L[7 * i];
        }
    }
    for (int i = 0; i < 6; ++i)
        for (int j = 0; j <= i; ++j) {
            double s = 0.0;
            for (int k = i; k < 6; ++k) s += Li[6 * k + i] * Li[6 * k + j];
            A[6 * i + j] = s;
            A[6 * j + i] = s;
        }
    return true;
}

}  // namespace ba
