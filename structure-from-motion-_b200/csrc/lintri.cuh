// lintri.cuh -- the per-point DLT of LinearTriangulation.py:57-92 (shared by lintri.cu and twoview.cu).
#pragma once
#include "sfm_common.cuh"

namespace sfm {

struct TriCams2 {
    double P1[12];
    double P2[12];
};

SFM_HD void dlt_triangulate(const TriCams2 &cams, double x1u, double x1v, double x2u, double x2v, double *X) {
    double g[4][4], v[4][4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        g[0][c] = x1v * cams.P1[8 + c] - cams.P1[4 + c];
        g[1][c] = cams.P1[c] - x1u * cams.P1[8 + c];
        g[2][c] = x2v * cams.P2[8 + c] - cams.P2[4 + c];
        g[3][c] = cams.P2[c] - x2u * cams.P2[8 + c];
#pragma unroll
        for (int r = 0; r < 4; ++r) v[r][c] = (r == c) ? 1.0 : 0.0;
    }
    for (int sweep = 0; sweep < 30; ++sweep) {
        bool rotated = false;
#pragma unroll
        for (int p = 0; p < 3; ++p) {
#pragma unroll
            for (int q = p + 1; q < 4; ++q) {
                double app = 0.0, aqq = 0.0, apq = 0.0;
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    app += g[r][p] * g[r][p];
                    aqq += g[r][q] * g[r][q];
                    apq += g[r][p] * g[r][q];
                }
                if (apq == 0.0 || fabs(apq) <= 1e-16 * sqrt(app * aqq)) continue;
                rotated = true;
                double zeta = (aqq - app) / (2.0 * apq);
                double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    double a = g[r][p], b = g[r][q];
                    g[r][p] = cs * a - sn * b;
                    g[r][q] = sn * a + cs * b;
                    a = v[r][p]; b = v[r][q];
                    v[r][p] = cs * a - sn * b;
                    v[r][q] = sn * a + cs * b;
                }
            }
        }
        if (!rotated) break;
    }
    double best = 1.7976931348623157e308;
    double h[4] = {0.0, 0.0, 0.0, 1.0};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        double nrm = g[0][c] * g[0][c] + g[1][c] * g[1][c] + g[2][c] * g[2][c] + g[3][c] * g[3][c];
        if (nrm < best) {           // NaN norms never win: a non-finite input yields the (0,0,0,1) default
            best = nrm;
#pragma unroll
            for (int r = 0; r < 4; ++r) h[r] = v[r][c];
        }
    }
    if (h[3] < 0.0) {
#pragma unroll
        for (int r = 0; r < 4; ++r) h[r] = -h[r];
    }
    if (fabs(h[3]) > 1e-8) {        // LinearTriangulation.py:88-89
        X[0] = h[0] / h[3]; X[1] = h[1] / h[3]; X[2] = h[2] / h[3];
    } else {                        // :90-92, point at infinity
        X[0] = h[0]; X[1] = h[1]; X[2] = h[2];
    }
}

}  // namespace sfm
