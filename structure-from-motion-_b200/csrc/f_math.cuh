// f_math.cuh -- per-thread arithmetic of the fundamental-matrix path.
//
// Replaces (reference, "Phase 1/"):
//   EstimateFundamentalMatrix.py:30-78  -> f8pt_from_sample()
//   GetInliersRANSAC.py:67-81           -> f_is_inlier() / f_error_exact()
//
// Quirks kept on purpose: design rows encode n1^T F n2 while the
// de-normalisation is T2^T F T1 (Q1); +1e-8 on the Hartley scale and on both line norms
// (Q2); inlier test is strict '<' and NaN is never an inlier (Q3, Q10); the metric is the
// mean of the two point-to-epipolar-line distances, not Sampson (Q12).
#pragma once
#include "sfm_common.cuh"

namespace sfm {

// ---- one-sided (Hestenes) Jacobi on a 3x3 ------------------------------------------------
// Rotates columns p,q of G (3x3, column access through the three row arrays) and of V so
// that the two columns of G become orthogonal. Returns true when a rotation was applied.
SFM_HD bool jacobi_pair3(double &g0p, double &g1p, double &g2p, double &g0q, double &g1q,
                         double &g2q, double &v0p, double &v1p, double &v2p, double &v0q,
                         double &v1q, double &v2q) {
    double app = g0p * g0p + g1p * g1p + g2p * g2p;
    double aqq = g0q * g0q + g1q * g1q + g2q * g2q;
    double apq = g0p * g0q + g1p * g1q + g2p * g2q;
    // converged for this pair (also covers zero columns)
    if (fabs(apq) <= 1e-15 * sqrt(app * aqq) || apq == 0.0) return false;
    double zeta = (aqq - app) / (2.0 * apq);
    double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
    double c = 1.0 / sqrt(1.0 + t * t);
    double s = c * t;
    double a, b;
    a = g0p; b = g0q; g0p = c * a - s * b; g0q = s * a + c * b;
    a = g1p; b = g1q; g1p = c * a - s * b; g1q = s * a + c * b;
    a = g2p; b = g2q; g2p = c * a - s * b; g2q = s * a + c * b;
    a = v0p; b = v0q; v0p = c * a - s * b; v0q = s * a + c * b;
    a = v1p; b = v1q; v1p = c * a - s * b; v1q = s * a + c * b;
    a = v2p; b = v2q; v2p = c * a - s * b; v2q = s * a + c * b;
    return true;
}

// Rank-2 projection of a 3x3 M (row-major m[9]) : U diag(s0,s1,0) V^T, i.e. what
// EstimateFundamentalMatrix.py:70-72 does with numpy's SVD. In place.
SFM_HD void rank2_project3(double *m) {
    double g00 = m[0], g01 = m[1], g02 = m[2];
    double g10 = m[3], g11 = m[4], g12 = m[5];
    double g20 = m[6], g21 = m[7], g22 = m[8];
    double v00 = 1, v01 = 0, v02 = 0, v10 = 0, v11 = 1, v12 = 0, v20 = 0, v21 = 0, v22 = 1;
    for (int sweep = 0; sweep < 40; ++sweep) {
        bool r0 = jacobi_pair3(g00, g10, g20, g01, g11, g21, v00, v10, v20, v01, v11, v21);
        bool r1 = jacobi_pair3(g00, g10, g20, g02, g12, g22, v00, v10, v20, v02, v12, v22);
        bool r2 = jacobi_pair3(g01, g11, g21, g02, g12, g22, v01, v11, v21, v02, v12, v22);
        if (!(r0 | r1 | r2)) break;
    }
    // G = M V has orthogonal columns sigma_j u_j. Drop the smallest one:
    // M2 = sum over the two largest columns of g_j v_j^T.
    double n0 = g00 * g00 + g10 * g10 + g20 * g20;
    double n1 = g01 * g01 + g11 * g11 + g21 * g21;
    double n2 = g02 * g02 + g12 * g12 + g22 * g22;
    double w0 = 1.0, w1 = 1.0, w2 = 1.0;
    if (n0 <= n1 && n0 <= n2) w0 = 0.0;
    else if (n1 <= n0 && n1 <= n2) w1 = 0.0;
    else w2 = 0.0;
    g00 *= w0; g10 *= w0; g20 *= w0;
    g01 *= w1; g11 *= w1; g21 *= w1;
    g02 *= w2; g12 *= w2; g22 *= w2;
    m[0] = g00 * v00 + g01 * v01 + g02 * v02;
    m[1] = g00 * v10 + g01 * v11 + g02 * v12;
    m[2] = g00 * v20 + g01 * v21 + g02 * v22;
    m[3] = g10 * v00 + g11 * v01 + g12 * v02;
    m[4] = g10 * v10 + g11 * v11 + g12 * v12;
    m[5] = g10 * v20 + g11 * v21 + g12 * v22;
    m[6] = g20 * v00 + g21 * v01 + g22 * v02;
    m[7] = g20 * v10 + g21 * v11 + g22 * v12;
    m[8] = g20 * v20 + g21 * v21 + g22 * v22;
}

// ---- null vector of an 8x9 matrix ----------------------------------------------------------
// a[r][c] = row r (r = 0..7) of the design matrix. The reference takes Vt[-1] of the full
// SVD (EstimateFundamentalMatrix.py:65-67), which for an 8x9 matrix is the unit vector
// orthogonal to all eight rows. We get the same vector (up to the sign that the F[2][2]
// normalisation removes) as the last column of Q in the Householder QR of A^T (9x8): it is
// backward stable like the SVD, needs ~1.3 kflop instead of a full Jacobi SVD and every
// index below is a compile-time constant after unrolling, so A lives in registers.
// Rank-deficient samples (duplicated matches) have a >1-dimensional null space;
// LAPACK's pick there is rounding noise, ours is "orthogonal to the rows, from e9".
SFM_HD void null_vector_8x9(double (&a)[8][9], double (&q)[9]) {
    double beta[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        // column j of B = A^T is row j of A; entries j..8 are active
        double nrm2 = 0.0;
#pragma unroll
        for (int i = j; i < 9; ++i) nrm2 += a[j][i] * a[j][i];
        double nrm = sqrt(nrm2);
        double x0 = a[j][j];
        double alpha = (x0 >= 0.0) ? -nrm : nrm;
        double v0 = x0 - alpha;
        // v = (v0, a[j][j+1..8]);  v^T v = nrm2 - x0^2 + v0^2 = 2 nrm (nrm + |x0|)
        double vtv = 2.0 * nrm * (nrm + fabs(x0));
        double b = (vtv > 0.0) ? 2.0 / vtv : 0.0;
        beta[j] = b;
        a[j][j] = v0;
#pragma unroll
        for (int k = j + 1; k < 8; ++k) {
            double dot = 0.0;
#pragma unroll
            for (int i = j; i < 9; ++i) dot += a[j][i] * a[k][i];
            dot *= b;
#pragma unroll
            for (int i = j; i < 9; ++i) a[k][i] -= dot * a[j][i];
        }
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) q[i] = 0.0;
    q[8] = 1.0;
#pragma unroll
    for (int j = 7; j >= 0; --j) {
        double dot = 0.0;
#pragma unroll
        for (int i = j; i < 9; ++i) dot += a[j][i] * q[i];
        dot *= beta[j];
#pragma unroll
        for (int i = j; i < 9; ++i) q[i] -= dot * a[j][i];
    }
}

// ---- Hartley-normalised 8-point ------------------------------------------------------------
// px1[i],py1[i],px2[i],py2[i]: the eight sampled correspondences. F: row-major 3x3,
// divided by F[2][2] (EstimateFundamentalMatrix.py:78; a zero F[2][2] yields inf/nan exactly
// as numpy does, and such a hypothesis then scores zero inliers).
SFM_HD void f8pt_from_sample(const double (&px1)[8], const double (&py1)[8],
                             const double (&px2)[8], const double (&py2)[8], double *F) {
    double mx1 = 0, my1 = 0, mx2 = 0, my2 = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { mx1 += px1[i]; my1 += py1[i]; mx2 += px2[i]; my2 += py2[i]; }
    mx1 *= 0.125; my1 *= 0.125; mx2 *= 0.125; my2 *= 0.125;
    double d1 = 0, d2 = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        double ax = px1[i] - mx1, ay = py1[i] - my1;
        double bx = px2[i] - mx2, by = py2[i] - my2;
        d1 += sqrt(ax * ax + ay * ay);
        d2 += sqrt(bx * bx + by * by);
    }
    const double SQRT2 = 1.4142135623730951;
    double s1 = SQRT2 / (d1 * 0.125 + 1e-8);
    double s2 = SQRT2 / (d2 * 0.125 + 1e-8);
    double tx1 = -s1 * mx1, ty1 = -s1 * my1, tx2 = -s2 * mx2, ty2 = -s2 * my2;

    double a[8][9];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        // normalized = T @ [x, y, 1]  (EstimateFundamentalMatrix.py:54-55)
        double x1 = s1 * px1[i] + tx1, y1 = s1 * py1[i] + ty1;
        double x2 = s2 * px2[i] + tx2, y2 = s2 * py2[i] + ty2;
        a[i][0] = x1 * x2; a[i][1] = x1 * y2; a[i][2] = x1;
        a[i][3] = y1 * x2; a[i][4] = y1 * y2; a[i][5] = y1;
        a[i][6] = x2;      a[i][7] = y2;      a[i][8] = 1.0;
    }
    double f[9];
    null_vector_8x9(a, f);
    rank2_project3(f);
    // F = T2^T Fh T1   with T = [[s,0,tx],[0,s,ty],[0,0,1]]
    // (Fh T1): column ops; then T2^T: row ops.
    double g[9];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        g[3 * r + 0] = f[3 * r + 0] * s1;
        g[3 * r + 1] = f[3 * r + 1] * s1;
        g[3 * r + 2] = f[3 * r + 0] * tx1 + f[3 * r + 1] * ty1 + f[3 * r + 2];
    }
    double o[9];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        o[0 + c] = s2 * g[0 + c];
        o[3 + c] = s2 * g[3 + c];
        o[6 + c] = tx2 * g[0 + c] + ty2 * g[3 + c] + g[6 + c];
    }
    double inv = o[8];
#pragma unroll
    for (int i = 0; i < 9; ++i) F[i] = o[i] / inv;
}

// ---- inlier test -----------------------------------------------------------------------------
// err = (|e|/(n1+1e-8) + |e|/(n2+1e-8))/2 < thr   (GetInliersRANSAC.py:72-81), evaluated in
// the algebraically identical division-free form
//     |e| (n1+n2+2eps) < 2 thr (n1+eps)(n2+eps)
// (both denominators are positive, so the inequality direction is preserved). two_thr = 2*thr.
// NaN anywhere makes the comparison false, like numpy's.
struct FMat {
    double f0, f1, f2, f3, f4, f5, f6, f7, f8;
};

SFM_HD bool f_is_inlier(const FMat &F, double x1, double y1, double x2, double y2,
                        double two_thr) {
    double a = fma(F.f0, x1, fma(F.f1, y1, F.f2));
    double b = fma(F.f3, x1, fma(F.f4, y1, F.f5));
    double c = fma(F.f6, x1, fma(F.f7, y1, F.f8));
    double e = fma(x2, a, fma(y2, b, c));
    double p = fma(F.f0, x2, fma(F.f3, y2, F.f6));
    double q = fma(F.f1, x2, fma(F.f4, y2, F.f7));
    double n1 = fast_sqrt_pos(fma(a, a, fma(b, b, SFM_TINY)));
    double n2 = fast_sqrt_pos(fma(p, p, fma(q, q, SFM_TINY)));
    double u = n1 + n2;
    double lhs = fabs(e) * (u + 2e-8);
    double rhs = two_thr * fma(n1, n2, fma(1e-8, u, 1e-16));
    return lhs < rhs;
}

// ---- certified single-precision filter for the inlier test -------------------------------------
// The counting kernel evaluates d = |e| (n1+n2) - 2 thr n1 n2 in FP32 together with a forward
// error bound tol on |d - (lhs - rhs of f_is_inlier)|. When |d| > tol the sign of d IS the FP64
// decision and the test costs FP32 issue slots only; otherwise (a few tests in 10^4 on real
// data: |err - thr| below ~10 % of thr) the caller re-evaluates f_is_inlier in FP64. Counts are
// therefore identical to the all-FP64 kernel by construction.
//
// Error model (u = 2^-24, M1/M2 = max |coordinate| of image 1 / 2 over the staged rows):
//   |a_f - a| <= 4.01u Sa, Sa = (|F0|+|F1|) M1 + |F2|      (two input roundings + two FMAs)
//   |e_f - e| <= 7.1u Se,  Se = M2 (Sa+Sb) + Sc
//   |n_f - (n+eps)| <= 4.02u (Sa+Sb) + eps + 5.2u n_f       (sqrt.approx <= 2^-22 relative,
//                                                            3e-19 covers underflow / ftz)
// propagated through the two products; every constant below carries a factor >= 2 over the
// bound above, which also absorbs the rounding of the bound's own evaluation and the ~1e-12
// relative error of the FP64 form. Non-finite or out-of-range constants force the FP64 path.
struct FCert {
    float f0, f1, f2, f3, f4, f5, f6, f7, f8;
    float T;      // 2 thr
    float DeT;    // multiplies s = n1 + n2
    float Dn12;   // multiplies |e|
    float C0;     // constant term
    // certified early reject (no square roots): see f_cert_reject
    float nDe;    // -(bound on |e_f - e|)
    float Ea1;    // multiplies min(q1_f, q2_f)
    float Ea0;    // constant term (+inf: never reject early)
};

SFM_HD float up_f32(double v) {
    // float not below v (v >= 0): round to nearest then widen by 2^-20 relative
    float f = (float)(v * 1.000001);
    return f;
}

SFM_HD FCert f_cert_setup(const FMat &F, double M1, double M2, double two_thr) {
    const double u = 5.9604644775390625e-8;  // 2^-24
    FCert c;
    c.f0 = (float)F.f0; c.f1 = (float)F.f1; c.f2 = (float)F.f2;
    c.f3 = (float)F.f3; c.f4 = (float)F.f4; c.f5 = (float)F.f5;
    c.f6 = (float)F.f6; c.f7 = (float)F.f7; c.f8 = (float)F.f8;
    c.T = (float)two_thr;
    double Sa = (fabs(F.f0) + fabs(F.f1)) * M1 + fabs(F.f2);
    double Sb = (fabs(F.f3) + fabs(F.f4)) * M1 + fabs(F.f5);
    double Sc = (fabs(F.f6) + fabs(F.f7)) * M1 + fabs(F.f8);
    double Sp = (fabs(F.f0) + fabs(F.f3)) * M2 + fabs(F.f6);
    double Sq = (fabs(F.f1) + fabs(F.f4)) * M2 + fabs(F.f7);
    double Se = M2 * (Sa + Sb) + Sc;
    double De = 16.0 * u * Se;
    double Dn1 = 10.0 * u * (Sa + Sb) + 1e-8 + 3e-19;
    double Dn2 = 10.0 * u * (Sp + Sq) + 1e-8 + 3e-19;
    double T = fabs(two_thr);
    double DeT = De + 1.01 * T * fmax(Dn1, Dn2);
    double Dn12 = 1.01 * (Dn1 + Dn2);
    double C0 = 1.01 * (De * (Dn1 + Dn2) + T * Dn1 * Dn2) + 1e-36;
    // operands must stay in the range where FP32 neither overflows nor loses the bound to
    // underflow; anything else (including NaN) always takes the FP64 path
    bool ok = Se > 1e-12 && Se < 1e12 && M1 < 1e6 && M2 < 1e6 && T < 1e6 && two_thr >= 0.0;
    const float inf = 3.0e38f * 10.0f;
    c.DeT = ok ? up_f32(DeT) : inf;
    c.Dn12 = up_f32(Dn12);
    c.C0 = up_f32(C0);
    // ---- early reject. With N_i = n_i + 1e-8 the reference's test (|e|/N1 + |e|/N2)/2 < thr is
    // |e| (1/N1 + 1/N2) < T, and 1/N1 + 1/N2 > 1/min(N1, N2): a test with |e| >= T min(N1, N2) is an outlier --
    // no square root needed, compare squares. Bounds, with q_i = n_i^2 exact, Q_i = a_f^2 + b_f^2 of the FP32
    // intermediates, kap_1 = Sa + Sb >= n_1 (kap_2 likewise), kap = max, and 2 sqrt(x) d <= d (x / kap + kap):
    //   q_i <= Q_i (1 + 4.01u) + 4.01u kap_i^2 (1 + 4.01u)        (|a_f - a| <= 4.01u Sa, ...)
    //   Q_i <= q_i,f (1 + 2.1u)                                    (one FMUL + one FFMA rounding)
    //   min(N1,N2)^2 <= m (1 + eps / kap) + eps kap + eps^2,  m = min(q1, q2)
    // => min(N)^2 <= A min(q1_f, q2_f) + B,  A = 1 + 7u + eps/kap,  B = 4.02u kap^2 (1 + eps/kap) + eps kap + eps^2,
    // and |e| >= |e_f| - De. The factor 1.00002 on T^2 covers the FP32 roundings of t, t^2 and of the right-hand
    // side themselves (< 8u) and the ~1e-12 relative error of the FP64 form the decision is compared with.
    {
        const double eps = 1e-8, kap = fmax(fmax(Sa + Sb, Sp + Sq), 1e-300);
        const double A = 1.0 + 7.0 * u + eps / kap;
        // + 1e-40: q_i,f may be denormal or flushed (absolute rounding error of a few 2^-149)
        const double B = 4.02 * u * kap * kap * (1.0 + eps / kap) + eps * kap + eps * eps + 1e-40;
        const double T2 = T * T * 1.00002;
        c.nDe = -up_f32(De);
        c.Ea1 = up_f32(T2 * A);
        c.Ea0 = (ok && kap < 1e15) ? up_f32(T2 * B) + 1e-36f : inf;
    }
    return c;
}

// Certified outlier without square roots: true only if the FP64 decision f_is_inlier is certainly false. e_f and
// the two squared norms are the FP32 intermediates of f_cert_margin (same operations, same order).
SFM_HD bool f_cert_reject_from(const FCert &c, float e_abs, float q1, float q2) {
    const float t = e_abs + c.nDe;
    const float m = fminf(q1, q2);
    const float rhs = fmaf(c.Ea1, m, c.Ea0);
    return t > 0.0f && t * t > rhs;      // strict: an infinite right-hand side (range guard, overflow) never rejects
}

// First stage of the same certificate, from ONE of the two squared norms: fmaf is monotone in its operands and
// Ea1 > 0, so the right-hand side built from q1 is never below the one built from min(q1, q2) -- whatever this
// rejects, f_cert_reject_from rejects (tests/test_hostsim_f.py checks the implication). It needs neither F^T x2
// nor the second norm, which is what K2 saves on the ~93 % of its votes that pass here.
SFM_HD bool f_cert_reject_stage1_from(const FCert &c, float e_abs, float q1) {
    const float t = e_abs + c.nDe;
    const float rhs = fmaf(c.Ea1, q1, c.Ea0);
    return t > 0.0f && t * t > rhs;
}

SFM_HD bool f_cert_reject_stage1(const FCert &c, float x1, float y1, float x2, float y2) {
    float a = fmaf(c.f0, x1, fmaf(c.f1, y1, c.f2));
    float b = fmaf(c.f3, x1, fmaf(c.f4, y1, c.f5));
    float cc = fmaf(c.f6, x1, fmaf(c.f7, y1, c.f8));
    float e = fabsf(fmaf(x2, a, fmaf(y2, b, cc)));
    return f_cert_reject_stage1_from(c, e, fmaf(a, a, b * b));
}

SFM_HD bool f_cert_reject(const FCert &c, float x1, float y1, float x2, float y2) {
    float a = fmaf(c.f0, x1, fmaf(c.f1, y1, c.f2));
    float b = fmaf(c.f3, x1, fmaf(c.f4, y1, c.f5));
    float cc = fmaf(c.f6, x1, fmaf(c.f7, y1, c.f8));
    float e = fabsf(fmaf(x2, a, fmaf(y2, b, cc)));
    float p = fmaf(c.f0, x2, fmaf(c.f3, y2, c.f6));
    float q = fmaf(c.f1, x2, fmaf(c.f4, y2, c.f7));
    return f_cert_reject_from(c, e, fmaf(a, a, b * b), fmaf(p, p, q * q));
}

SFM_HD float sqrt_approx_f32(float s) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));
    return r;
#else
    return sqrtf(s);
#endif
}

// Returns d (negative = inlier) and sets tol; the decision is certified iff |d| > tol.
SFM_HD float f_cert_margin(const FCert &c, float x1, float y1, float x2, float y2, float &tol) {
    float a = fmaf(c.f0, x1, fmaf(c.f1, y1, c.f2));
    float b = fmaf(c.f3, x1, fmaf(c.f4, y1, c.f5));
    float cc = fmaf(c.f6, x1, fmaf(c.f7, y1, c.f8));
    float e = fabsf(fmaf(x2, a, fmaf(y2, b, cc)));
    float p = fmaf(c.f0, x2, fmaf(c.f3, y2, c.f6));
    float q = fmaf(c.f1, x2, fmaf(c.f4, y2, c.f7));
    float n1 = sqrt_approx_f32(fmaf(a, a, b * b));
    float n2 = sqrt_approx_f32(fmaf(p, p, q * q));
    float s = n1 + n2;
    float lhs = e * s;
    float d = fmaf(-(c.T * n1), n2, lhs);
    // rounding of the evaluation itself: <= 32u (lhs + rhs), and rhs = lhs - d <= lhs + |d|, so
    // |d| > 65u lhs + (rest) implies |d| > 32u (lhs + rhs) + (rest): rhs is never formed.
    tol = fmaf(3.874301910400391e-6f /* 65u */, lhs, c.C0);
    tol = fmaf(c.DeT, s, tol);
    tol = fmaf(e, c.Dn12, tol);
    return d;
}

// The reference's expression verbatim (for the reported per-point error of the winner).
SFM_HD double f_error_exact(const FMat &F, double x1, double y1, double x2, double y2) {
    double a = F.f0 * x1 + F.f1 * y1 + F.f2;
    double b = F.f3 * x1 + F.f4 * y1 + F.f5;
    double c = F.f6 * x1 + F.f7 * y1 + F.f8;
    double e = x2 * a + y2 * b + c;
    double p = F.f0 * x2 + F.f3 * y2 + F.f6;
    double q = F.f1 * x2 + F.f4 * y2 + F.f7;
    double dist1 = fabs(e) / (sqrt(a * a + b * b) + 1e-8);
    double dist2 = fabs(e) / (sqrt(p * p + q * q) + 1e-8);
    return (dist1 + dist2) / 2;
}

}  // namespace sfm
