// pose_math.cuh -- rotation-vector camera model and analytic reprojection Jacobians.
//
// Reference conventions: camera parameters are [rotvec rho(3), t(3)],
// R = Rotation.from_rotvec(rho) (NonlinearPnP.py:30, BundleAdjustment.py:87), P = K [R | t],
// projection u^ = a/d, v^ = b/d with d = w + 1e-8 (NonlinearPnP.py:39, BundleAdjustment.py:38),
// residual = observed - projected (NonlinearPnP.py:42, BundleAdjustment.py:107).
// The reference differentiates by finite differences inside scipy; we use the closed forms:
//   d(R X)/d rho = -R [X]x Jr(rho),  Jr = I - (1-cos th)/th^2 [rho]x + (th - sin th)/th^3 [rho]x^2
// (right Jacobian of SO(3)), with Taylor forms below th^2 < 1e-4 -- camera 1 of every
// reconstruction has rho == 0 exactly (Wrapper_dev.py:160-161).
#pragma once
#include "sfm_common.cuh"

namespace sfm {

struct Mat3 {
    double m[9];  // row-major
};

// a = sin(th)/th, b = (1-cos th)/th^2, c = (th - sin th)/th^3
SFM_HD void so3_coeffs(double th2, double &a, double &b, double &c) {
    if (th2 < 1e-4) {
        a = 1.0 - th2 * (1.0 / 6.0 - th2 * (1.0 / 120.0 - th2 / 5040.0));
        b = 0.5 - th2 * (1.0 / 24.0 - th2 * (1.0 / 720.0 - th2 / 40320.0));
        c = 1.0 / 6.0 - th2 * (1.0 / 120.0 - th2 * (1.0 / 5040.0 - th2 / 362880.0));
    } else {
        double th = sqrt(th2);
        double s = sin(th);
        a = s / th;
        // 1 - cos th = 2 sin^2(th/2): no cancellation
        double sh = sin(0.5 * th);
        b = 2.0 * sh * sh / th2;
        if (th2 < 0.5) {
            // th - sin th cancels badly for small th: keep the series (next term < 2e-14)
            c = 1.0 / 6.0 - th2 * (1.0 / 120.0 - th2 * (1.0 / 5040.0 - th2 * (1.0 / 362880.0 -
                th2 * (1.0 / 39916800.0 - th2 / 6227020800.0))));
        } else {
            c = (th - s) / (th2 * th);
        }
    }
}

// R = I + a [r]x + b [r]x^2 ;  Jr = I - b [r]x + c [r]x^2
SFM_HD void rotvec_to_R_Jr(const double *r, Mat3 &R, Mat3 *Jr) {
    double x = r[0], y = r[1], z = r[2];
    double th2 = x * x + y * y + z * z;
    double a, b, c;
    so3_coeffs(th2, a, b, c);
    // [r]x^2 = r r^T - th2 I
    double xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, yz = y * z;
    R.m[0] = 1.0 - b * (yy + zz); R.m[1] = b * xy - a * z;      R.m[2] = b * xz + a * y;
    R.m[3] = b * xy + a * z;      R.m[4] = 1.0 - b * (xx + zz); R.m[5] = b * yz - a * x;
    R.m[6] = b * xz - a * y;      R.m[7] = b * yz + a * x;      R.m[8] = 1.0 - b * (xx + yy);
    if (Jr) {
        Jr->m[0] = 1.0 - c * (yy + zz); Jr->m[1] = c * xy + b * z;      Jr->m[2] = c * xz - b * y;
        Jr->m[3] = c * xy - b * z;      Jr->m[4] = 1.0 - c * (xx + zz); Jr->m[5] = c * yz + b * x;
        Jr->m[6] = c * xz + b * y;      Jr->m[7] = c * yz - b * x;      Jr->m[8] = 1.0 - c * (xx + yy);
    }
}

SFM_HD void mat3_mul(const double *A, const double *B, double *C) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}

// Per-camera constants for one linearisation: M = K R, kt = K t, Jr.
struct CamLin {
    double M[9];
    double kt[3];
    double Jr[9];
};

SFM_HD void cam_linearise(const double *K, const double *cam6, CamLin &c) {
    Mat3 R, Jr;
    rotvec_to_R_Jr(cam6, R, &Jr);
    mat3_mul(K, R.m, c.M);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        c.kt[i] = K[3 * i] * cam6[3] + K[3 * i + 1] * cam6[4] + K[3 * i + 2] * cam6[5];
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) c.Jr[i] = Jr.m[i];
}

// Reprojection residual of one observation (ox, oy) of point X under a linearised camera,
// with the reference's +eps in the denominator. Returns r = obs - proj.
// If jc != nullptr: jc[2][6] = d r / d(rho, t), jp[2][3] = d r / d X.
SFM_HD void reproj_residual(const CamLin &c, const double *K, const double *X, double ox, double oy,
                            double eps, double &ru, double &rv, double (*jc)[6], double (*jp)[3]) {
    double a = c.M[0] * X[0] + c.M[1] * X[1] + c.M[2] * X[2] + c.kt[0];
    double b = c.M[3] * X[0] + c.M[4] * X[1] + c.M[5] * X[2] + c.kt[1];
    double w = c.M[6] * X[0] + c.M[7] * X[1] + c.M[8] * X[2] + c.kt[2];
    double d = w + eps;
    double inv = 1.0 / d;
    double pu = a * inv, pv = b * inv;
    ru = ox - pu;
    rv = oy - pv;
    if (!jp) return;
    // G = d r / d(Xc direction in K R coordinates) = -(1/d) [M0 - pu M2; M1 - pv M2]  (2x3) = dr/dX
    double G0[3], G1[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        G0[j] = -inv * (c.M[j] - pu * c.M[6 + j]);
        G1[j] = -inv * (c.M[3 + j] - pv * c.M[6 + j]);
        jp[0][j] = G0[j];
        jp[1][j] = G1[j];
    }
    if (!jc) return;
    // d r / d t = -(1/d) [K0 - pu K2; K1 - pv K2]
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        jc[0][3 + j] = -inv * (K[j] - pu * K[6 + j]);
        jc[1][3 + j] = -inv * (K[3 + j] - pv * K[6 + j]);
    }
    // d r / d rho = G (-[X]x) Jr ;  row g -> (g x X)... (-[X]x)^T g = X x g?  work it out row-wise:
    // (g^T (-[X]x))_j = -(g^T [X]x)_j = ([X]x g)_j  since [X]x^T = -[X]x  ->  h = X x g
    double h0[3] = {X[1] * G0[2] - X[2] * G0[1], X[2] * G0[0] - X[0] * G0[2], X[0] * G0[1] - X[1] * G0[0]};
    double h1[3] = {X[1] * G1[2] - X[2] * G1[1], X[2] * G1[0] - X[0] * G1[2], X[0] * G1[1] - X[1] * G1[0]};
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        jc[0][j] = h0[0] * c.Jr[j] + h0[1] * c.Jr[3 + j] + h0[2] * c.Jr[6 + j];
        jc[1][j] = h1[0] * c.Jr[j] + h1[1] * c.Jr[3 + j] + h1[2] * c.Jr[6 + j];
    }
}

}  // namespace sfm
