// ba_math.cuh -- per-observation / per-point arithmetic of the bundle-adjustment chain (K8-K12).
//
// Residual and parameter layout follow BundleAdjustment.py:43-110:
//   camera block [rotvec(3), t(3)], point block [X,Y,Z]; r = observed - projected with the
//   +1e-8 in the projective denominator (BundleAdjustment.py:38); observations point-major.
// The reference differentiates by finite differences inside scipy; the kernels use the analytic
// 2x6 / 2x3 blocks of pose_math.cuh (checked against central differences of the reference
// residual in tests/test_hostsim_lm.py).
#pragma once
#include "lm_small.cuh"
#include "pose_math.cuh"

namespace sfm {

constexpr double BA_EPS = 1e-8;          // BundleAdjustment.py:38
constexpr int CAMLIN_STRIDE = 22;        // doubles per linearised camera: M(9) kt(3) Jr(9) + 1 pad, so a
                                         // record is 16-byte aligned and read with 11 128-bit loads: the
                                         // per-lane gather of records from shared memory is bank-conflict
                                         // bound, and wide loads halve its wavefronts

SFM_HD void camlin_store(const CamLin &c, double *dst) {
#pragma unroll
    for (int i = 0; i < 9; ++i) dst[i] = c.M[i];
#pragma unroll
    for (int i = 0; i < 3; ++i) dst[9 + i] = c.kt[i];
#pragma unroll
    for (int i = 0; i < 9; ++i) dst[12 + i] = c.Jr[i];
    dst[21] = 0.0;
}

SFM_HD void camlin_load(const double *src, CamLin &c) {
#ifdef __CUDA_ARCH__
    double t[CAMLIN_STRIDE];
    const double2 *s2 = reinterpret_cast<const double2 *>(src);
#pragma unroll
    for (int i = 0; i < CAMLIN_STRIDE / 2; ++i) {
        const double2 v = s2[i];
        t[2 * i] = v.x;
        t[2 * i + 1] = v.y;
    }
#else
    const double *t = src;
#endif
#pragma unroll
    for (int i = 0; i < 9; ++i) c.M[i] = t[i];
#pragma unroll
    for (int i = 0; i < 3; ++i) c.kt[i] = t[9 + i];
#pragma unroll
    for (int i = 0; i < 9; ++i) c.Jr[i] = t[12 + i];
}

// M and kt only (the residual-only pass)
SFM_HD void camlin_load_proj(const double *src, CamLin &c) {
#ifdef __CUDA_ARCH__
    double t[12];
    const double2 *s2 = reinterpret_cast<const double2 *>(src);
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        const double2 v = s2[i];
        t[2 * i] = v.x;
        t[2 * i + 1] = v.y;
    }
#else
    const double *t = src;
#endif
#pragma unroll
    for (int i = 0; i < 9; ++i) c.M[i] = t[i];
#pragma unroll
    for (int i = 0; i < 3; ++i) c.kt[i] = t[9 + i];
}

// residual only (K8): no Jacobian work
SFM_HD void ba_residual_only(const CamLin &c, const double *X, double ox, double oy, double &ru, double &rv) {
    double a = c.M[0] * X[0] + c.M[1] * X[1] + c.M[2] * X[2] + c.kt[0];
    double b = c.M[3] * X[0] + c.M[4] * X[1] + c.M[5] * X[2] + c.kt[1];
    double w = c.M[6] * X[0] + c.M[7] * X[1] + c.M[8] * X[2] + c.kt[2];
    double inv = 1.0 / (w + BA_EPS);
    ru = ox - a * inv;
    rv = oy - b * inv;
}

// Inverse of a symmetric positive definite 3x3 given packed lower (00,10,11,20,21,22), through its
// Cholesky factor. Returns false when the matrix is not numerically positive definite.
SFM_HD bool spd3_inverse(const double *v, double *inv) {
    double l00 = v[0];
    if (!(l00 > 0.0)) return false;
    l00 = sqrt(l00);
    double i00 = 1.0 / l00;
    double l10 = v[1] * i00, l20 = v[3] * i00;
    double d1 = v[2] - l10 * l10;
    if (!(d1 > 0.0)) return false;
    double l11 = sqrt(d1);
    double i11 = 1.0 / l11;
    double l21 = (v[4] - l20 * l10) * i11;
    double d2 = v[5] - l20 * l20 - l21 * l21;
    if (!(d2 > 0.0)) return false;
    double l22 = sqrt(d2);
    double i22 = 1.0 / l22;
    // M = L^-1 (lower)
    double m10 = -l10 * i00 * i11;
    double m21 = -l21 * i11 * i22;
    double m20 = -(l20 * i00 + l21 * m10) * i22;
    // inv = M^T M
    inv[0] = i00 * i00 + m10 * m10 + m20 * m20;
    inv[1] = m10 * i11 + m20 * m21;
    inv[2] = i11 * i11 + m21 * m21;
    inv[3] = m20 * i22;
    inv[4] = m21 * i22;
    inv[5] = i22 * i22;
    return is_finite(inv[0]) && is_finite(inv[2]) && is_finite(inv[5]);
}

// y = S x for packed-lower symmetric 3x3
SFM_HD void sym3_mul(const double *s, const double *x, double *y) {
    y[0] = s[0] * x[0] + s[1] * x[1] + s[3] * x[2];
    y[1] = s[1] * x[0] + s[2] * x[1] + s[4] * x[2];
    y[2] = s[3] * x[0] + s[4] * x[1] + s[5] * x[2];
}

}  // namespace sfm
