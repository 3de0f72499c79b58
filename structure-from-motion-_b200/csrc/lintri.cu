// lintri.cu -- N1: batched linear (DLT) triangulation, one thread per point.
//
// Replaces the per-point numpy loop of LinearTriangulation.py:57-90: the 4x4 system
//   [y1 P1[2] - P1[1];  P1[0] - x1 P1[2];  y2 P2[2] - P2[1];  P2[0] - x2 P2[2]] X~ = 0   (:76-81)
// is solved by the right singular vector of its smallest singular value (Vt[-1], :84-85), then
// de-homogenised when |X~[3]| > 1e-8, otherwise the first three components are returned as they
// are (:88-92). The reference calls LAPACK's dgesdd; here a one-sided (Hestenes) Jacobi SVD runs
// in registers: columns of A are rotated until mutually orthogonal, the column of least norm
// marks the singular vector in V. Jacobi is at least as accurate as the bidiagonalisation route,
// and the vector's sign -- the only LAPACK artefact -- cancels in the division (the |X~[3]| <= 1e-8
// branch, a point at infinity, keeps our sign convention: X~[3] >= 0).
// It produces x0 for every NonLinearTriangulation call (Wrapper_dev.py:179,243,252,280).
#include "sfm_common.cuh"
#include "lintri.cuh"

namespace sfm {

#if defined(__CUDACC__)
__global__ void __launch_bounds__(128)
lintri_kernel(TriCams2 cams, const double *__restrict__ x1, const double *__restrict__ x2, int n,
              double *__restrict__ Xout) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double2 a = __ldg(reinterpret_cast<const double2 *>(x1) + i);
    double2 b = __ldg(reinterpret_cast<const double2 *>(x2) + i);
    double X[3];
    dlt_triangulate(cams, a.x, a.y, b.x, b.y, X);
    Xout[3 * (size_t)i] = X[0];
    Xout[3 * (size_t)i + 1] = X[1];
    Xout[3 * (size_t)i + 2] = X[2];
}
#endif

}  // namespace sfm

#if defined(__CUDACC__)
using namespace sfm;

extern "C" int sfm_lintri(const double *h_P1, const double *h_P2, const double *d_x1, const double *d_x2,
                          int32_t n, double *d_X, void *stream) {
    if (!h_P1 || !h_P2 || !d_x1 || !d_x2 || !d_X || n < 0) return SFM_E_INVALID;
    if (n == 0) return SFM_OK;
    TriCams2 cams;
    for (int i = 0; i < 12; ++i) { cams.P1[i] = h_P1[i]; cams.P2[i] = h_P2[i]; }
    lintri_kernel<<<ceil_div(n, 128), 128, 0, as_stream(stream)>>>(cams, d_x1, d_x2, n, d_X);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}
#endif
