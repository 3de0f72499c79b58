// h_math.cuh -- per-thread arithmetic of the homography RANSAC path.
//
// Replaces (reference, "Phase 1/GetHomographyInliers.py"):
//   find_homography :26-82 for a 4-point sample  -> h4pt_from_sample()
//   get_homography_inliers :139-149              -> h_is_inlier() / h_error_exact()
//
// Quirks kept: centroid + mean-distance normalisation with +1e-8 on the scale (:36-37), design rows
// [0,0,0,-x1,-y1,-1, y2 x1, y2 y1, y2] and [x1,y1,1,0,0,0, -x2 x1, -x2 y1, -x2] (:66-68), null vector
// = Vt[-1] of the 8x9 system, H = inv(T2) Hn T1 divided by H[2][2] (:77-80; a zero H[2][2] gives
// inf/nan like numpy and such a hypothesis scores zero), projection divided by (z + 1e-8) (:143),
// strict '<' on the Euclidean distance (:149), NaN never an inlier.
#pragma once
#include "f_math.cuh"

namespace sfm {

SFM_HD void h4pt_from_sample(const double (&px1)[4], const double (&py1)[4], const double (&px2)[4],
                             const double (&py2)[4], double *H) {
    double mx1 = 0, my1 = 0, mx2 = 0, my2 = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) { mx1 += px1[i]; my1 += py1[i]; mx2 += px2[i]; my2 += py2[i]; }
    mx1 *= 0.25; my1 *= 0.25; mx2 *= 0.25; my2 *= 0.25;
    double d1 = 0, d2 = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        double ax = px1[i] - mx1, ay = py1[i] - my1;
        double bx = px2[i] - mx2, by = py2[i] - my2;
        d1 += sqrt(ax * ax + ay * ay);
        d2 += sqrt(bx * bx + by * by);
    }
    const double SQRT2 = 1.4142135623730951;
    const double s1 = SQRT2 / (d1 * 0.25 + 1e-8), s2 = SQRT2 / (d2 * 0.25 + 1e-8);
    const double tx1 = -s1 * mx1, ty1 = -s1 * my1, tx2 = -s2 * mx2, ty2 = -s2 * my2;
    double a[8][9];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const double x1 = s1 * px1[i] + tx1, y1 = s1 * py1[i] + ty1;
        const double x2 = s2 * px2[i] + tx2, y2 = s2 * py2[i] + ty2;
        double *r0 = a[2 * i], *r1 = a[2 * i + 1];
        r0[0] = 0.0; r0[1] = 0.0; r0[2] = 0.0; r0[3] = -x1; r0[4] = -y1; r0[5] = -1.0;
        r0[6] = y2 * x1; r0[7] = y2 * y1; r0[8] = y2;
        r1[0] = x1; r1[1] = y1; r1[2] = 1.0; r1[3] = 0.0; r1[4] = 0.0; r1[5] = 0.0;
        r1[6] = -x2 * x1; r1[7] = -x2 * y1; r1[8] = -x2;
    }
    double f[9];
    null_vector_8x9(a, f);
    // (Hn T1): column operations, T = [[s,0,tx],[0,s,ty],[0,0,1]]
    double g[9];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        g[3 * r + 0] = f[3 * r + 0] * s1;
        g[3 * r + 1] = f[3 * r + 1] * s1;
        g[3 * r + 2] = f[3 * r + 0] * tx1 + f[3 * r + 1] * ty1 + f[3 * r + 2];
    }
    // inv(T2) = [[1/s2, 0, mx2], [0, 1/s2, my2], [0, 0, 1]]
    const double is2 = 1.0 / s2;
    double o[9];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        o[0 + c] = is2 * g[0 + c] + mx2 * g[6 + c];
        o[3 + c] = is2 * g[3 + c] + my2 * g[6 + c];
        o[6 + c] = g[6 + c];
    }
    const double inv = o[8];
#pragma unroll
    for (int i = 0; i < 9; ++i) H[i] = o[i] / inv;
}

struct HMat {
    double h0, h1, h2, h3, h4, h5, h6, h7, h8;
};

// |H x1 / (z + 1e-8) - x2| < thr in the division-free form (X - u w)^2 + (Y - v w)^2 < thr^2 w^2,
// w = z + 1e-8 (w^2 > 0 keeps the direction; w == 0 gives inf/nan in the reference and 'false' here
// because the left side is then >= 0 and the right side 0). thr2 = thr^2 for thr > 0, else -1.
SFM_HD bool h_is_inlier(const HMat &H, double x1, double y1, double x2, double y2, double thr2) {
    const double X = fma(H.h0, x1, fma(H.h1, y1, H.h2));
    const double Y = fma(H.h3, x1, fma(H.h4, y1, H.h5));
    const double w = fma(H.h6, x1, fma(H.h7, y1, H.h8)) + 1e-8;
    const double dx = fma(-x2, w, X), dy = fma(-y2, w, Y);
    return fma(dx, dx, dy * dy) < thr2 * (w * w);
}

// the reference's expression verbatim (per-point error of the winner)
SFM_HD double h_error_exact(const HMat &H, double x1, double y1, double x2, double y2) {
    const double X = H.h0 * x1 + H.h1 * y1 + H.h2;
    const double Y = H.h3 * x1 + H.h4 * y1 + H.h5;
    const double Z = H.h6 * x1 + H.h7 * y1 + H.h8;
    const double u = X / (Z + 1e-8) - x2, v = Y / (Z + 1e-8) - y2;
    return sqrt(u * u + v * v);
}

}  // namespace sfm
