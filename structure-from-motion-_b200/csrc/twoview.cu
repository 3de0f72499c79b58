// twoview.cu -- SURVEY.md 8(f) N4: the two-view pose front end on the device, so a pair stays in HBM
// between F-RANSAC and triangulation.
//
// Replaces (reference, "Phase 1/"):
//   EssentialMatrixFromFundamentalMatrix.py:19   E = K^T F K
//   ExtractCameraPose.py:24-81                   SVD of E, det fix, R1 = U W V^T, R2 = U W^T V^T,
//                                                t = U[:,2]/|U[:,2]|, C = -R^T (+-t)  -> 4 configurations
//   LinearTriangulation.py:44-92                 DLT of every correspondence under each configuration
//   DisambiguateCameraPose.py:53-82              cheirality vote: z > 0 in both cameras, strict '>' =
//                                                first maximum, starting from -1
//
// Order of the four configurations: the reference inherits it from LAPACK's sign choice for the
// singular vectors (flipping a (u_i, v_i) pair permutes the list, the set is invariant). Here the
// convention is fixed (see svd3_desc); parity tests compare the configurations as a set and the
// selected pose, points and vote exactly.
#include "sfm_common.cuh"
#include "f_math.cuh"
#include "lintri.cuh"

namespace sfm {

// M (row-major) = U diag(s) V^T, s descending, by one-sided Jacobi on the columns of M V.
// Signs: every column of V is oriented so that its largest-magnitude entry is positive; u_j follows
// from M v_j = s_j u_j; a (numerically) zero singular value gets u_2 = u_0 x u_1.
SFM_HD void svd3_desc(const double *m, double *U, double *s, double *V) {
    double g[3][3], v[3][3];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) { g[r][c] = m[3 * r + c]; v[r][c] = r == c ? 1.0 : 0.0; }
    for (int sweep = 0; sweep < 60; ++sweep) {
        bool r0 = jacobi_pair3(g[0][0], g[1][0], g[2][0], g[0][1], g[1][1], g[2][1], v[0][0], v[1][0], v[2][0], v[0][1], v[1][1], v[2][1]);
        bool r1 = jacobi_pair3(g[0][0], g[1][0], g[2][0], g[0][2], g[1][2], g[2][2], v[0][0], v[1][0], v[2][0], v[0][2], v[1][2], v[2][2]);
        bool r2 = jacobi_pair3(g[0][1], g[1][1], g[2][1], g[0][2], g[1][2], g[2][2], v[0][1], v[1][1], v[2][1], v[0][2], v[1][2], v[2][2]);
        if (!(r0 | r1 | r2)) break;
    }
    double n[3];
    int ord[3] = {0, 1, 2};
    for (int c = 0; c < 3; ++c) n[c] = sqrt(g[0][c] * g[0][c] + g[1][c] * g[1][c] + g[2][c] * g[2][c]);
    for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2 - a; ++b)
            if (n[ord[b]] < n[ord[b + 1]]) { int t = ord[b]; ord[b] = ord[b + 1]; ord[b + 1] = t; }
    for (int j = 0; j < 3; ++j) {
        const int c = ord[j];
        s[j] = n[c];
        int big = 0;
        for (int r = 1; r < 3; ++r) if (fabs(v[r][c]) > fabs(v[big][c])) big = r;
        const double sg = v[big][c] < 0.0 ? -1.0 : 1.0;
        for (int r = 0; r < 3; ++r) {
            V[3 * r + j] = sg * v[r][c];
            U[3 * r + j] = n[c] > 0.0 ? sg * g[r][c] / n[c] : 0.0;
        }
    }
    if (!(s[2] > 1e-12 * s[0])) {       // rank 2 (an essential matrix): complete the basis
        U[2] = U[3] * U[7] - U[6] * U[4];
        U[5] = U[6] * U[1] - U[0] * U[7];
        U[8] = U[0] * U[4] - U[3] * U[1];
        double nn = sqrt(U[2] * U[2] + U[5] * U[5] + U[8] * U[8]);
        if (nn > 0.0) { U[2] /= nn; U[5] /= nn; U[8] /= nn; }
    }
}

SFM_HD double det3(const double *a) {
    return a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6]);
}

// cfg [4][12] = (C[3], R[9] row-major) in the reference's order (R1,t) (R1,-t) (R2,t) (R2,-t);
// P [5][12]: P[0] = K [I | 0], P[1+i] = K R_i [I | -C_i] (LinearTriangulation.py:44-50)
SFM_HD void two_view_configs(const double *F, const double *K, double *cfg, double *P) {
    double FK[9], E[9];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) FK[3 * r + c] = F[3 * r] * K[c] + F[3 * r + 1] * K[3 + c] + F[3 * r + 2] * K[6 + c];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) E[3 * r + c] = K[r] * FK[c] + K[3 + r] * FK[3 + c] + K[6 + r] * FK[6 + c];
    double U[9], s[3], V[9];
    svd3_desc(E, U, s, V);
    if (det3(U) * det3(V) < 0.0) { U[2] = -U[2]; U[5] = -U[5]; U[8] = -U[8]; }      // ExtractCameraPose.py:30-31
    // U W V^T = -u0 v1^T + u1 v0^T + u2 v2^T ; U W^T V^T = u0 v1^T - u1 v0^T + u2 v2^T
    double R[2][9];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
            const double a = U[3 * r] * V[3 * c + 1], b = U[3 * r + 1] * V[3 * c], d = U[3 * r + 2] * V[3 * c + 2];
            R[0][3 * r + c] = -a + b + d;
            R[1][3 * r + c] = a - b + d;
        }
    double t[3] = {U[2], U[5], U[8]};
    const double tn = sqrt(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
    for (int i = 0; i < 3; ++i) t[i] /= tn;
    for (int i = 0; i < 12; ++i) P[i] = 0.0;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) P[4 * r + c] = K[3 * r + c];
    for (int i = 0; i < 4; ++i) {
        const double *Ri = R[i >> 1];
        const double sg = (i & 1) ? -1.0 : 1.0;
        double *o = cfg + 12 * i, *Pi = P + 12 * (i + 1);
        for (int c = 0; c < 3; ++c) o[c] = -(Ri[c] * t[0] + Ri[3 + c] * t[1] + Ri[6 + c] * t[2]) * sg;   // C = -R^T (+-t)
        for (int e = 0; e < 9; ++e) o[3 + e] = Ri[e];
        double Rt[12];                                     // [R | -R C]
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) Rt[4 * r + c] = Ri[3 * r + c];
            Rt[4 * r + 3] = -(Ri[3 * r] * o[0] + Ri[3 * r + 1] * o[1] + Ri[3 * r + 2] * o[2]);
        }
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 4; ++c) Pi[4 * r + c] = K[3 * r] * Rt[c] + K[3 * r + 1] * Rt[4 + c] + K[3 * r + 2] * Rt[8 + c];
    }
}

SFM_HD bool in_front_of_both(const double *cfg, const double *X) {
    const double *C = cfg, *R = cfg + 3;
    const double z2 = R[6] * (X[0] - C[0]) + R[7] * (X[1] - C[1]) + R[8] * (X[2] - C[2]);
    return X[2] > 0.0 && z2 > 0.0;                         // DisambiguateCameraPose.py:66-76
}

#if defined(__CUDACC__)
struct K3x3 { double k[9]; };

__global__ void two_view_setup_kernel(const double *__restrict__ F, K3x3 K, double *__restrict__ cfg,
                                      double *__restrict__ P, int32_t *__restrict__ counts) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double f[9];
        for (int i = 0; i < 9; ++i) f[i] = F[i];
        two_view_configs(f, K.k, cfg, P);
        for (int i = 0; i < 4; ++i) counts[i] = 0;
    }
}

// thread per correspondence: DLT under each of the four configurations + cheirality vote
__global__ void __launch_bounds__(128)
two_view_tri_kernel(const double *__restrict__ cfg, const double *__restrict__ P, const double *__restrict__ x1,
                    const double *__restrict__ x2, int n, double *__restrict__ Xall, int32_t *__restrict__ counts) {
    __shared__ double sP[60], sC[48];
    if (threadIdx.x < 60) sP[threadIdx.x] = P[threadIdx.x];
    if (threadIdx.x >= 64 && threadIdx.x < 112) sC[threadIdx.x - 64] = cfg[threadIdx.x - 64];
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = i < n;
    double2 a = make_double2(0.0, 0.0), b = a;
    if (active) {
        a = __ldg(reinterpret_cast<const double2 *>(x1) + i);
        b = __ldg(reinterpret_cast<const double2 *>(x2) + i);
    }
    TriCams2 cams;
#pragma unroll
    for (int e = 0; e < 12; ++e) cams.P1[e] = sP[e];
    for (int c = 0; c < 4; ++c) {
        bool front = false;
        if (active) {
#pragma unroll
            for (int e = 0; e < 12; ++e) cams.P2[e] = sP[12 * (c + 1) + e];
            double X[3];
            dlt_triangulate(cams, a.x, a.y, b.x, b.y, X);
            double *o = Xall + ((size_t)c * n + i) * 3;
            o[0] = X[0]; o[1] = X[1]; o[2] = X[2];
            front = in_front_of_both(sC + 12 * c, X);
        }
        const unsigned m = __ballot_sync(0xffffffffu, front);
        if ((threadIdx.x & 31) == 0 && m) atomicAdd(&counts[c], __popc(m));
    }
}

// cheirality vote alone, for points triangulated elsewhere (DisambiguateCameraPose's signature)
__global__ void __launch_bounds__(128)
cheirality_kernel(const double *__restrict__ cfg, const double *__restrict__ Xall, int n,
                  int32_t *__restrict__ counts) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, c = blockIdx.y;
    bool front = false;
    if (i < n) {
        const double *x = Xall + ((size_t)c * n + i) * 3;
        double X[3] = {x[0], x[1], x[2]}, cf[12];
        for (int e = 0; e < 12; ++e) cf[e] = cfg[12 * c + e];
        front = in_front_of_both(cf, X);
    }
    const unsigned m = __ballot_sync(0xffffffffu, front);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(&counts[c], __popc(m));
}

// first maximum with the reference's strict '>' from -1; best = (index, count); pose = its (C, R)
__global__ void two_view_select_kernel(const int32_t *__restrict__ counts, const double *__restrict__ cfg,
                                       int32_t *__restrict__ best, double *__restrict__ pose) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        int bi = 0, bc = -1;
        for (int i = 0; i < 4; ++i)
            if (counts[i] > bc) { bc = counts[i]; bi = i; }
        best[0] = bi;
        best[1] = bc;
        for (int e = 0; e < 12; ++e) pose[e] = cfg[12 * bi + e];
    }
}
#endif

}  // namespace sfm

#if defined(__CUDACC__)
using namespace sfm;

extern "C" int sfm_two_view_configs(const double *d_F, const double *h_K, double *d_cfg48, double *d_P60,
                                    int32_t *d_counts4, void *stream) {
    if (!d_F || !h_K || !d_cfg48 || !d_P60 || !d_counts4) return SFM_E_INVALID;
    K3x3 K;
    for (int i = 0; i < 9; ++i) K.k[i] = h_K[i];
    two_view_setup_kernel<<<1, 32, 0, as_stream(stream)>>>(d_F, K, d_cfg48, d_P60, d_counts4);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_cheirality_count(const double *d_cfg48, const double *d_Xall, int32_t n, int32_t *d_counts4,
                                    void *stream) {
    if (!d_cfg48 || !d_counts4 || n < 0 || (n > 0 && !d_Xall)) return SFM_E_INVALID;
    cudaStream_t st = as_stream(stream);
    SFM_CUDA_TRY(cudaMemsetAsync(d_counts4, 0, 4 * sizeof(int32_t), st));
    if (n == 0) return SFM_OK;
    cheirality_kernel<<<dim3(ceil_div(n, 128), 4), 128, 0, st>>>(d_cfg48, d_Xall, n, d_counts4);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_two_view_pose(const double *d_F, const double *h_K, const double *d_x1, const double *d_x2,
                                 int32_t n, double *d_cfg48, double *d_P60, double *d_Xall, int32_t *d_counts4,
                                 int32_t *d_best2, double *d_pose12, void *stream) {
    if (!d_F || !h_K || !d_cfg48 || !d_P60 || !d_counts4 || !d_best2 || !d_pose12 || n < 0) return SFM_E_INVALID;
    if (n > 0 && (!d_x1 || !d_x2 || !d_Xall)) return SFM_E_INVALID;
    cudaStream_t st = as_stream(stream);
    int rc = sfm_two_view_configs(d_F, h_K, d_cfg48, d_P60, d_counts4, stream);
    if (rc != SFM_OK) return rc;
    if (n > 0) {
        two_view_tri_kernel<<<ceil_div(n, 128), 128, 0, st>>>(d_cfg48, d_P60, d_x1, d_x2, n, d_Xall, d_counts4);
        SFM_LAUNCH_CHECK();
    }
    two_view_select_kernel<<<1, 32, 0, st>>>(d_counts4, d_cfg48, d_best2, d_pose12);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}
#endif
