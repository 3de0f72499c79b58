// linpnp.cu -- K6a: LinearPnP on the device (LinearPnP.py:3-96), the hypothesis generator of PnPRANSAC
// (PnPRANSAC.py:49-55) and its all-points fallback (:82-87).
//
//   linear_pnp4_kernel     thread per hypothesis: gather the 4 sampled correspondences, 8 x 12 DLT matrix in
//                          registers / local memory, numpy's Vt[-1] by replaying LAPACK's bidiagonalisation
//                          (linpnp_math.cuh rule (1)), QR clean-up with LAPACK's sign convention, pose + K [R | -R C].
//                          FP64 latency bound; 1,000 hypotheses of a registration are one small launch, a batch of
//                          registrations one launch over registrations x hypotheses.
//   linear_pnp_npt_kernel  one 1024-thread block for N > 4 correspondences: 2N x 12 design matrix column-major in a
//                          global workspace, one-sided Jacobi on its twelve columns (A^T A is never formed), the
//                          column of least norm = Vt[-1] up to sign -- and the sign does not matter: det(M) < 0
//                          flips it (LinearPnP.py:65-67).
#include "linpnp_math.cuh"

namespace sfm {

constexpr int LP_THREADS = 128;
constexpr int LPN_THREADS = 1024;

struct LinPnpK { double K[9], Kinv[9]; };

// d_sample [h,4] indices into the correspondences of registration `reg` (offsets[reg] .. offsets[reg+1]);
// offsets == nullptr: one registration of n correspondences.
__global__ void __launch_bounds__(LP_THREADS)
linear_pnp4_kernel(LinPnpK k, const double *__restrict__ X, const double *__restrict__ x, const int32_t *__restrict__ offsets,
                   int n, const int32_t *__restrict__ sample, int h_per_reg, int h_total, int mute_reflected,
                   double *__restrict__ C,
                   double *__restrict__ R, double *__restrict__ P, unsigned char *__restrict__ flag) {
    const int hyp = blockIdx.x * LP_THREADS + threadIdx.x;
    if (hyp >= h_total) return;
    const int reg = hyp / h_per_reg;
    const int base = offsets ? offsets[reg] : 0;
    const int cnt = offsets ? offsets[reg + 1] - base : n;
    double X4[12], x4[8];
    bool ok = true;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int s = sample[4 * (size_t)hyp + i];
        ok = ok && s >= 0 && s < cnt;
        const size_t q = (size_t)base + (ok ? s : 0);
        X4[3 * i] = X[3 * q]; X4[3 * i + 1] = X[3 * q + 1]; X4[3 * i + 2] = X[3 * q + 2];
        x4[2 * i] = x[2 * q]; x4[2 * i + 1] = x[2 * q + 1];
    }
    double c[3], r[9], p[12];
    unsigned char f = linear_pnp4(k.K, k.Kinv, X4, x4, c, r, p);
    if ((f & LINPNP_FLAG_REFLECTED) && mute_reflected) {
        // parity policy: in the reference this hypothesis carries a pose drawn from rounding noise, which all but
        // never collects a consensus; here it keeps its documented pose in C / R but does not vote (NaN projection
        // matrix -> K6 counts zero inliers, and a count of zero never wins: PnPRANSAC.py:72)
        const double nan = __longlong_as_double(0x7ff8000000000000LL);
#pragma unroll
        for (int i = 0; i < 12; ++i) p[i] = nan;
    }
    if (!ok) {                                  // an index outside the registration: no pose
        f = LINPNP_FLAG_NONFINITE;
        const double nan = __longlong_as_double(0x7ff8000000000000LL);
#pragma unroll
        for (int i = 0; i < 12; ++i) p[i] = nan;
#pragma unroll
        for (int i = 0; i < 9; ++i) r[i] = nan;
        c[0] = c[1] = c[2] = nan;
    }
    if (C) { C[3 * (size_t)hyp] = c[0]; C[3 * (size_t)hyp + 1] = c[1]; C[3 * (size_t)hyp + 2] = c[2]; }
    if (R) {
#pragma unroll
        for (int i = 0; i < 9; ++i) R[9 * (size_t)hyp + i] = r[i];
    }
#pragma unroll
    for (int i = 0; i < 12; ++i) P[12 * (size_t)hyp + i] = p[i];
    if (flag) flag[hyp] = f;
}

// ---- N-point: block-wide one-sided Jacobi on twelve columns -------------------------------------------
template <int K>
__device__ __forceinline__ void lpn_block_sum(double (&v)[K], double *sred /* [32*K] */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; ++k)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    __syncthreads();
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < K; ++k) sred[warp * K + k] = v[k];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < K; ++k) {
        double s = 0.0;
        for (int w = 0; w < LPN_THREADS / 32; ++w) s += sred[w * K + k];     // fixed order: identical in every thread
        v[k] = s;
    }
}

__global__ void __launch_bounds__(LPN_THREADS)
linear_pnp_npt_kernel(LinPnpK k, const double *__restrict__ X, const double *__restrict__ x, int n,
                      double *__restrict__ A /* [12][2n] column-major */, double *__restrict__ CR, double *__restrict__ P,
                      unsigned char *__restrict__ flag) {
    __shared__ double sred[32 * 3];
    __shared__ double V[12][12];
    const int tid = threadIdx.x, m = 2 * n;
    if (m < 12) {
        // fewer equations than unknowns (n = 4, 5): Vt[-1] is LAPACK's pick inside the null space (linpnp_math.cuh (1))
        if (tid == 0) {
            for (int i = 0; i < n; ++i) linpnp_rows(k.Kinv, X + 3 * (size_t)i, x + 2 * (size_t)i, A + 24 * i, A + 24 * i + 12);
            double p[12], cc[3], r[9], pp[12];
            gebd2_null_vector_rt(A, m, p);
            const unsigned char f = linpnp_finish(p, k.K, cc, r, pp);
            CR[0] = cc[0]; CR[1] = cc[1]; CR[2] = cc[2];
            for (int j = 0; j < 9; ++j) CR[3 + j] = r[j];
            if (P) for (int j = 0; j < 12; ++j) P[j] = pp[j];
            if (flag) flag[0] = f;
        }
        return;
    }
    for (int i = tid; i < n; i += LPN_THREADS) {
        double r0[12], r1[12];
        linpnp_rows(k.Kinv, X + 3 * (size_t)i, x + 2 * (size_t)i, r0, r1);
#pragma unroll
        for (int c = 0; c < 12; ++c) {
            A[(size_t)c * m + 2 * i] = r0[c];
            A[(size_t)c * m + 2 * i + 1] = r1[c];
        }
    }
    for (int e = tid; e < 144; e += LPN_THREADS) V[e / 12][e % 12] = (e / 12 == e % 12) ? 1.0 : 0.0;
    __syncthreads();
    for (int sweep = 0; sweep < 60; ++sweep) {
        int rotated = 0;
        for (int p = 0; p < 11; ++p)
            for (int q = p + 1; q < 12; ++q) {
                double *cp = A + (size_t)p * m, *cq = A + (size_t)q * m;
                double acc[3] = {0, 0, 0};
                for (int i = tid; i < m; i += LPN_THREADS) {
                    const double a = cp[i], b = cq[i];
                    acc[0] += a * a; acc[1] += b * b; acc[2] += a * b;
                }
                lpn_block_sum<3>(acc, sred);
                const double app = acc[0], aqq = acc[1], apq = acc[2];
                if (apq == 0.0 || fabs(apq) <= 1e-15 * sqrt(app * aqq)) continue;      // block-uniform
                rotated = 1;
                const double zeta = (aqq - app) / (2.0 * apq);
                const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
                for (int i = tid; i < m; i += LPN_THREADS) {
                    const double a = cp[i], b = cq[i];
                    cp[i] = c * a - s * b;
                    cq[i] = s * a + c * b;
                }
                if (tid < 12) {
                    const double a = V[tid][p], b = V[tid][q];
                    V[tid][p] = c * a - s * b;
                    V[tid][q] = s * a + c * b;
                }
            }
        if (!rotated) break;
    }
    __syncthreads();
    // the column of least norm carries the smallest singular value
    double best = 0.0;
    int jmin = 0;
    for (int c = 0; c < 12; ++c) {
        double acc[1] = {0.0};
        for (int i = tid; i < m; i += LPN_THREADS) { const double a = A[(size_t)c * m + i]; acc[0] += a * a; }
        lpn_block_sum<1>(acc, sred);
        if (c == 0 || acc[0] < best) { best = acc[0]; jmin = c; }
    }
    if (tid == 0) {
        double p[12], cc[3], r[9], pp[12];
        for (int j = 0; j < 12; ++j) p[j] = V[j][jmin];
        const unsigned char f = linpnp_finish(p, k.K, cc, r, pp);
        CR[0] = cc[0]; CR[1] = cc[1]; CR[2] = cc[2];
        for (int j = 0; j < 9; ++j) CR[3 + j] = r[j];
        if (P) for (int j = 0; j < 12; ++j) P[j] = pp[j];
        if (flag) flag[0] = f;
    }
}

static bool make_k(const double *h_K, const double *h_Kinv, LinPnpK &k) {
    for (int i = 0; i < 9; ++i) k.K[i] = h_K[i];
    if (h_Kinv) {
        for (int i = 0; i < 9; ++i) k.Kinv[i] = h_Kinv[i];
        return true;
    }
    const double *a = h_K;                      // adjugate / determinant
    const double det = a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6]);
    if (!(det != 0.0)) return false;
    const double id = 1.0 / det;
    k.Kinv[0] = (a[4] * a[8] - a[5] * a[7]) * id; k.Kinv[1] = (a[2] * a[7] - a[1] * a[8]) * id; k.Kinv[2] = (a[1] * a[5] - a[2] * a[4]) * id;
    k.Kinv[3] = (a[5] * a[6] - a[3] * a[8]) * id; k.Kinv[4] = (a[0] * a[8] - a[2] * a[6]) * id; k.Kinv[5] = (a[2] * a[3] - a[0] * a[5]) * id;
    k.Kinv[6] = (a[3] * a[7] - a[4] * a[6]) * id; k.Kinv[7] = (a[1] * a[6] - a[0] * a[7]) * id; k.Kinv[8] = (a[0] * a[4] - a[1] * a[3]) * id;
    return true;
}

}  // namespace sfm

using namespace sfm;

extern "C" int sfm_linear_pnp4_batch(const double *d_X, const double *d_x, const int32_t *d_offsets, int32_t n_regs,
                                     int32_t n, const int32_t *d_sample, int32_t h_per_reg, const double *h_K,
                                     const double *h_Kinv, int32_t policy, double *d_C, double *d_R, double *d_P,
                                     uint8_t *d_flag, void *stream) {
    if (!d_X || !d_x || !d_sample || !h_K || !d_P || n_regs < 1 || h_per_reg < 1) return SFM_E_INVALID;
    if (!d_offsets && (n_regs != 1 || n < 4)) return SFM_E_INVALID;
    LinPnpK k;
    if (!make_k(h_K, h_Kinv, k)) return SFM_E_INVALID;
    const long long total = (long long)n_regs * h_per_reg;
    if (total > 0x7fffffffLL) return SFM_E_INVALID;
    linear_pnp4_kernel<<<(unsigned)((total + LP_THREADS - 1) / LP_THREADS), LP_THREADS, 0, as_stream(stream)>>>(
        k, d_X, d_x, d_offsets, n, d_sample, h_per_reg, (int)total, policy & 1, d_C, d_R, d_P, d_flag);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" size_t sfm_linear_pnp_npt_workspace_bytes(int32_t n) {
    return n > 0 ? sizeof(double) * 24 * (size_t)n : 0;
}

extern "C" int sfm_linear_pnp_npt(const double *d_X, const double *d_x, int32_t n, const double *h_K,
                                  const double *h_Kinv, double *d_CR, double *d_P, uint8_t *d_flag, void *d_ws,
                                  size_t ws_bytes, void *stream) {
    if (!d_X || !d_x || !h_K || !d_CR || !d_ws || n < 4) return SFM_E_INVALID;
    if (ws_bytes < sfm_linear_pnp_npt_workspace_bytes(n)) return SFM_E_WORKSPACE;
    LinPnpK k;
    if (!make_k(h_K, h_Kinv, k)) return SFM_E_INVALID;
    linear_pnp_npt_kernel<<<1, LPN_THREADS, 0, as_stream(stream)>>>(k, d_X, d_x, n, reinterpret_cast<double *>(d_ws), d_CR,
                                                                   d_P, d_flag);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}
