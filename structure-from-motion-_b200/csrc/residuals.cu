// residuals.cu -- the residual functions the reference exports next to its solvers, as kernels of their own:
//   Loss(X, x1, x2, P1, P2)            NonLinearTriangulation.py:5-50   -> sfm_tri_residuals
//   NonLinearPnPLoss(X0, X, x, K)      NonlinearPnP.py:5-44             -> sfm_pnp_residuals
// K5 / K7 evaluate the same expressions inside their LM loops (nltri.cu tri_view_accumulate, pose_math.cuh
// reproj_residual); these entry points expose them for the drop-in names `Loss` / `NonLinearPnPLoss` and pin them
// against the reference's own residual vectors (tests/golden: loss0 in nltri.npz, loss0_* in pnp.npz).
#include "pose_math.cuh"

namespace sfm {

struct TriP { double P1[12], P2[12]; };

// one view of Loss(): exact division, a view with |w| < 1e-8 contributes a zero residual (:33-34, :40-41)
__device__ __forceinline__ void tri_view_residual(const double *P, const double *X, double ox, double oy, double &ru,
                                                  double &rv) {
    const double a = P[0] * X[0] + P[1] * X[1] + P[2] * X[2] + P[3];
    const double b = P[4] * X[0] + P[5] * X[1] + P[6] * X[2] + P[7];
    const double w = P[8] * X[0] + P[9] * X[1] + P[10] * X[2] + P[11];
    if (fabs(w) < 1e-8) { ru = 0.0; rv = 0.0; return; }
    const double inv = 1.0 / w;
    ru = ox - a * inv;
    rv = oy - b * inv;
}

__global__ void tri_residuals_kernel(TriP p, const double2 *__restrict__ x1, const double2 *__restrict__ x2,
                                     const double *__restrict__ X, int n, double *__restrict__ r) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double Xi[3] = {X[3 * (size_t)i], X[3 * (size_t)i + 1], X[3 * (size_t)i + 2]};
    const double2 a = x1[i], b = x2[i];
    double r0, r1, r2, r3;
    tri_view_residual(p.P1, Xi, a.x, a.y, r0, r1);
    tri_view_residual(p.P2, Xi, b.x, b.y, r2, r3);
    reinterpret_cast<double2 *>(r)[2 * (size_t)i] = make_double2(r0, r1);
    reinterpret_cast<double2 *>(r)[2 * (size_t)i + 1] = make_double2(r2, r3);
}

struct PnpK { double K[9], rt[6]; };

__global__ void pnp_residuals_kernel(PnpK p, const double *__restrict__ X, const double2 *__restrict__ x, int n,
                                     double2 *__restrict__ r) {
    __shared__ CamLin c;
    if (threadIdx.x == 0) cam_linearise(p.K, p.rt, c);
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double Xi[3] = {X[3 * (size_t)i], X[3 * (size_t)i + 1], X[3 * (size_t)i + 2]};
    const double2 o = x[i];
    double ru, rv;
    reproj_residual(c, p.K, Xi, o.x, o.y, 1e-8, ru, rv, nullptr, nullptr);   // +1e-8 in the denominator (:39)
    r[i] = make_double2(ru, rv);                                      // interleaved [du, dv] (:42)
}

}  // namespace sfm

using namespace sfm;

extern "C" int sfm_tri_residuals(const double *h_P1, const double *h_P2, const double *d_x1, const double *d_x2,
                                 const double *d_X, int32_t n, double *d_r, void *stream) {
    if (!h_P1 || !h_P2 || !d_x1 || !d_x2 || !d_X || !d_r || n < 0) return SFM_E_INVALID;
    if (n == 0) return SFM_OK;
    TriP p;
    for (int i = 0; i < 12; ++i) { p.P1[i] = h_P1[i]; p.P2[i] = h_P2[i]; }
    tri_residuals_kernel<<<ceil_div(n, 256), 256, 0, as_stream(stream)>>>(
        p, reinterpret_cast<const double2 *>(d_x1), reinterpret_cast<const double2 *>(d_x2), d_X, n, d_r);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_pnp_residuals(const double *h_K, const double *h_rt, const double *d_X, const double *d_x, int32_t n,
                                 double *d_r, void *stream) {
    if (!h_K || !h_rt || !d_X || !d_x || !d_r || n < 0) return SFM_E_INVALID;
    if (n == 0) return SFM_OK;
    PnpK p;
    for (int i = 0; i < 9; ++i) p.K[i] = h_K[i];
    for (int i = 0; i < 6; ++i) p.rt[i] = h_rt[i];
    pnp_residuals_kernel<<<ceil_div(n, 256), 256, 0, as_stream(stream)>>>(p, d_X, reinterpret_cast<const double2 *>(d_x), n,
                                                                          reinterpret_cast<double2 *>(d_r));
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}
