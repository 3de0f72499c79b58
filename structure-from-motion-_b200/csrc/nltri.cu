// nltri.cu -- K5: batched non-linear triangulation, one thread per point.
//
// Replaces the per-point scipy least_squares loop of NonLinearTriangulation.py:105-128 with
// the residual of Loss() (NonLinearTriangulation.py:5-50): r = [x1 - pi(P1 X~), x2 - pi(P2 X~)],
// pi = exact division (no +1e-8 here), a view whose |w| < 1e-8 contributes a zero
// residual. 3 parameters, 4 residuals, analytic 4x3 Jacobian, LM controller of lm_small.cuh.
// Memory per point: x1 16 + x2 16 + X0 24 in, X 24 (+ cost 8 + status 4) out -- FP64-ALU /
// latency bound, not HBM bound: reported as points/s.
#include "lm_small.cuh"

namespace sfm {

struct TriCams {
    double P1[12];
    double P2[12];
};

SFM_HD void tri_view_accumulate(const double *P, const double *X, double ox, double oy,
                                NormalEq<3> &ne) {
    double a = P[0] * X[0] + P[1] * X[1] + P[2] * X[2] + P[3];
    double b = P[4] * X[0] + P[5] * X[1] + P[6] * X[2] + P[7];
    double w = P[8] * X[0] + P[9] * X[1] + P[10] * X[2] + P[11];
    if (fabs(w) < 1e-8) return;   // NonLinearTriangulation.py:33-34,40-41: residual falls back to 0
    double inv = 1.0 / w;
    double pu = a * inv, pv = b * inv;
    double ru = ox - pu, rv = oy - pv;
    double ju[3], jv[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        ju[j] = -inv * (P[j] - pu * P[8 + j]);
        jv[j] = -inv * (P[4 + j] - pv * P[8 + j]);
    }
    ne.cost += 0.5 * (ru * ru + rv * rv);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        ne.g[r] += ju[r] * ru + jv[r] * rv;
#pragma unroll
        for (int c = 0; c <= r; ++c) ne.A[sym_idx(r, c)] += ju[r] * ju[c] + jv[r] * jv[c];
    }
}

SFM_HD void tri_eval(const TriCams &cams, const double *X, double x1u, double x1v, double x2u,
                     double x2v, NormalEq<3> &ne) {
#pragma unroll
    for (int i = 0; i < 6; ++i) ne.A[i] = 0.0;
    ne.g[0] = ne.g[1] = ne.g[2] = 0.0;
    ne.cost = 0.0;
    tri_view_accumulate(cams.P1, X, x1u, x1v, ne);
    tri_view_accumulate(cams.P2, X, x2u, x2v, ne);
}

// One complete LM solve. Returns the status; X is updated in place, cost/iters reported.
SFM_HD int tri_solve_point(const TriCams &cams, double *X, double x1u, double x1v, double x2u,
                           double x2v, int max_iter, double tol, double *cost_out, int *iters_out) {
    NormalEq<3> ne, net;
    LMState<3> s;
    s.x[0] = X[0]; s.x[1] = X[1]; s.x[2] = X[2];
    tri_eval(cams, s.x, x1u, x1v, x2u, x2v, ne);
    lm_init<3>(s, ne);
    while (s.status == LM_RUNNING) {
        if (s.need_outer) {
            lm_begin_outer<3>(s, ne, tol);
            if (s.status != LM_RUNNING) break;
        }
        double p[3], xt[3];
        lm_propose<3>(s, ne, p, xt);
        tri_eval(cams, xt, x1u, x1v, x2u, x2v, net);
        lm_judge<3>(s, ne, net, p, xt, tol, tol, max_iter);
    }
    if (s.status != LM_FAILED && is_finite(s.x[0]) && is_finite(s.x[1]) && is_finite(s.x[2])) {
        X[0] = s.x[0]; X[1] = s.x[1]; X[2] = s.x[2];
    }   // else: keep the initial estimate, like the reference's except branch (:126-128)
    if (cost_out) *cost_out = ne.cost;
    if (iters_out) *iters_out = s.nfev;
    return s.status;
}

#if defined(__CUDACC__)
__global__ void __launch_bounds__(128)
nltri_kernel(TriCams cams, const double *__restrict__ x1, const double *__restrict__ x2,
             const double *__restrict__ X0, int n, int max_iter, double tol,
             double *__restrict__ Xout, double *__restrict__ cost, int32_t *__restrict__ status) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double2 a = __ldg(reinterpret_cast<const double2 *>(x1) + i);
    double2 b = __ldg(reinterpret_cast<const double2 *>(x2) + i);
    double X[3] = {X0[3 * (size_t)i], X0[3 * (size_t)i + 1], X0[3 * (size_t)i + 2]};
    double c;
    int it;
    int st = tri_solve_point(cams, X, a.x, a.y, b.x, b.y, max_iter, tol, &c, &it);
    Xout[3 * (size_t)i] = X[0];
    Xout[3 * (size_t)i + 1] = X[1];
    Xout[3 * (size_t)i + 2] = X[2];
    if (cost) cost[i] = c;
    if (status) status[i] = (st << 16) | (it & 0xFFFF);
}
#endif

}  // namespace sfm

#if defined(__CUDACC__)
using namespace sfm;

extern "C" int sfm_nltri(const double *h_P1, const double *h_P2, const double *d_x1,
                         const double *d_x2, const double *d_X0, int32_t n, int32_t max_iter,
                         double tol, double *d_X, double *d_cost, int32_t *d_status, void *stream) {
    if (!h_P1 || !h_P2 || !d_x1 || !d_x2 || !d_X0 || !d_X || n < 0 || max_iter < 0) return SFM_E_INVALID;
    if (n == 0) return SFM_OK;
    TriCams cams;
    for (int i = 0; i < 12; ++i) { cams.P1[i] = h_P1[i]; cams.P2[i] = h_P2[i]; }
    nltri_kernel<<<ceil_div(n, 128), 128, 0, as_stream(stream)>>>(cams, d_x1, d_x2, d_X0, n, max_iter,
                                                                  tol, d_X, d_cost, d_status);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}
#endif
