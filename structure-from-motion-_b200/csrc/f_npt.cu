// f_npt.cu -- EstimateFundamentalMatrix for an arbitrary number of correspondences n >= 8
// (EstimateFundamentalMatrix.py:3-83 called outside RANSAC, e.g. on all inliers).
//
// Cold path (one call per image pair at most), so one 1024-thread block does everything:
// Hartley normalisation by block reductions, the n x 9 design matrix in a column-major
// global workspace (coalesced; L2 resident), and a one-sided (Hestenes) Jacobi SVD directly
// on its columns -- A^T A is never formed, so the smallest right singular vector keeps
// LAPACK-class accuracy (hard part 2). The column that converges to the
// smallest norm gives Vt[-1].
#include "f_math.cuh"
#include <mutex>
#include <stdio.h>

namespace sfm {

constexpr int NPT_THREADS = 1024;

template <int K>
__device__ __forceinline__ void block_sum(double (&v)[K], double *sred /* [32*K] */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    }
    __syncthreads();  // protect sred from the previous use
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < K; ++k) sred[warp * K + k] = v[k];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < K; ++k) {
        double s = 0.0;
        // fixed order over the 32 warps: every thread gets the identical sum
        for (int w = 0; w < NPT_THREADS / 32; ++w) s += sred[w * K + k];
        v[k] = s;
    }
}

// One-sided (Hestenes) Jacobi on the nine columns of A [9][n] (column-major, global workspace),
// accumulating the rotations in V (shared, identity on entry). Returns the index of the column of
// least norm -- V[:, jmin] is the right singular vector of the smallest singular value (Vt[-1]).
// Every thread of the block gets the same value.
__device__ int block_jacobi_null9(double *__restrict__ A, int n, double (*V)[9], double *sred) {
    const int tid = threadIdx.x;
    for (int sweep = 0; sweep < 60; ++sweep) {
        int rotated = 0;
        for (int p = 0; p < 8; ++p) {
            for (int q = p + 1; q < 9; ++q) {
                double *cp = A + (size_t)p * n, *cq = A + (size_t)q * n;
                double acc[3] = {0, 0, 0};
                for (int i = tid; i < n; i += NPT_THREADS) {
                    double a = cp[i], b = cq[i];
                    acc[0] += a * a; acc[1] += b * b; acc[2] += a * b;
                }
                block_sum<3>(acc, sred);
                const double app = acc[0], aqq = acc[1], apq = acc[2];
                // identical values in every thread -> uniform branch
                if (apq == 0.0 || fabs(apq) <= 1e-15 * sqrt(app * aqq)) continue;
                rotated = 1;
                double zeta = (aqq - app) / (2.0 * apq);
                double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
                for (int i = tid; i < n; i += NPT_THREADS) {
                    double a = cp[i], b = cq[i];
                    cp[i] = c * a - s * b;
                    cq[i] = s * a + c * b;
                }
                if (tid < 9) {
                    double a = V[tid][p], b = V[tid][q];
                    V[tid][p] = c * a - s * b;
                    V[tid][q] = s * a + c * b;
                }
                // the next pair's reads of A/V are ordered by the barriers inside block_sum
            }
        }
        if (!rotated) break;
    }
    __syncthreads();
    // column norms -> smallest singular value's column
    double nrm[9];
    {
        double acc9[9];
#pragma unroll
        for (int c = 0; c < 9; ++c) acc9[c] = 0.0;
        for (int i = tid; i < n; i += NPT_THREADS) {
#pragma unroll
            for (int c = 0; c < 9; ++c) { double a = A[(size_t)c * n + i]; acc9[c] += a * a; }
        }
        // three rounds of 3 to reuse the 4-wide scratch
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            double t3[3] = {acc9[3 * r], acc9[3 * r + 1], acc9[3 * r + 2]};
            block_sum<3>(t3, sred);
            nrm[3 * r] = t3[0]; nrm[3 * r + 1] = t3[1]; nrm[3 * r + 2] = t3[2];
        }
    }
    int jmin = 0;
    for (int c = 1; c < 9; ++c) if (nrm[c] < nrm[jmin]) jmin = c;
    return jmin;
}

__global__ void __launch_bounds__(NPT_THREADS)
fundamental_npt_kernel(const double *__restrict__ x1, const double *__restrict__ x2, int n,
                       double *__restrict__ A /* [9][n] */, double *__restrict__ Fout) {
    __shared__ double sred[32 * 4];
    __shared__ double V[9][9];
    const int tid = threadIdx.x;

    double m[4] = {0, 0, 0, 0};
    for (int i = tid; i < n; i += NPT_THREADS) {
        m[0] += x1[2 * i]; m[1] += x1[2 * i + 1]; m[2] += x2[2 * i]; m[3] += x2[2 * i + 1];
    }
    block_sum<4>(m, sred);
    const double inv_n = 1.0 / (double)n;
    const double mx1 = m[0] * inv_n, my1 = m[1] * inv_n, mx2 = m[2] * inv_n, my2 = m[3] * inv_n;
    double d[2] = {0, 0};
    for (int i = tid; i < n; i += NPT_THREADS) {
        double ax = x1[2 * i] - mx1, ay = x1[2 * i + 1] - my1;
        double bx = x2[2 * i] - mx2, by = x2[2 * i + 1] - my2;
        d[0] += sqrt(ax * ax + ay * ay);
        d[1] += sqrt(bx * bx + by * by);
    }
    block_sum<2>(d, sred);
    const double SQRT2 = 1.4142135623730951;
    const double s1 = SQRT2 / (d[0] * inv_n + 1e-8), s2 = SQRT2 / (d[1] * inv_n + 1e-8);
    const double tx1 = -s1 * mx1, ty1 = -s1 * my1, tx2 = -s2 * mx2, ty2 = -s2 * my2;

    for (int i = tid; i < n; i += NPT_THREADS) {
        double u1 = s1 * x1[2 * i] + tx1, v1 = s1 * x1[2 * i + 1] + ty1;
        double u2 = s2 * x2[2 * i] + tx2, v2 = s2 * x2[2 * i + 1] + ty2;
        A[0 * (size_t)n + i] = u1 * u2; A[1 * (size_t)n + i] = u1 * v2; A[2 * (size_t)n + i] = u1;
        A[3 * (size_t)n + i] = v1 * u2; A[4 * (size_t)n + i] = v1 * v2; A[5 * (size_t)n + i] = v1;
        A[6 * (size_t)n + i] = u2;      A[7 * (size_t)n + i] = v2;      A[8 * (size_t)n + i] = 1.0;
    }
    if (tid < 81) V[tid / 9][tid % 9] = (tid / 9 == tid % 9) ? 1.0 : 0.0;
    __syncthreads();

    const int jmin = block_jacobi_null9(A, n, V, sred);
    if (tid == 0) {
        double f[9];
        for (int r = 0; r < 9; ++r) f[r] = V[r][jmin];
        rank2_project3(f);
        double g[9], o[9];
        for (int r = 0; r < 3; ++r) {
            g[3 * r + 0] = f[3 * r + 0] * s1;
            g[3 * r + 1] = f[3 * r + 1] * s1;
            g[3 * r + 2] = f[3 * r + 0] * tx1 + f[3 * r + 1] * ty1 + f[3 * r + 2];
        }
        for (int c = 0; c < 3; ++c) {
            o[0 + c] = s2 * g[0 + c];
            o[3 + c] = s2 * g[3 + c];
            o[6 + c] = tx2 * g[0 + c] + ty2 * g[3 + c] + g[6 + c];
        }
        double inv = o[8];
        for (int i = 0; i < 9; ++i) Fout[i] = o[i] / inv;
    }
}

// find_homography (GetHomographyInliers.py:4-82) for n >= 4 correspondences: same block-wide scheme,
// design matrix [9][2n].
__global__ void __launch_bounds__(NPT_THREADS)
homography_npt_kernel(const double *__restrict__ x1, const double *__restrict__ x2, int n,
                      double *__restrict__ A /* [9][2n] */, double *__restrict__ Hout) {
    __shared__ double sred[32 * 4];
    __shared__ double V[9][9];
    const int tid = threadIdx.x;
    double m[4] = {0, 0, 0, 0};
    for (int i = tid; i < n; i += NPT_THREADS) {
        m[0] += x1[2 * i]; m[1] += x1[2 * i + 1]; m[2] += x2[2 * i]; m[3] += x2[2 * i + 1];
    }
    block_sum<4>(m, sred);
    const double inv_n = 1.0 / (double)n;
    const double mx1 = m[0] * inv_n, my1 = m[1] * inv_n, mx2 = m[2] * inv_n, my2 = m[3] * inv_n;
    double d[2] = {0, 0};
    for (int i = tid; i < n; i += NPT_THREADS) {
        double ax = x1[2 * i] - mx1, ay = x1[2 * i + 1] - my1;
        double bx = x2[2 * i] - mx2, by = x2[2 * i + 1] - my2;
        d[0] += sqrt(ax * ax + ay * ay);
        d[1] += sqrt(bx * bx + by * by);
    }
    block_sum<2>(d, sred);
    const double SQRT2 = 1.4142135623730951;
    const double s1 = SQRT2 / (d[0] * inv_n + 1e-8), s2 = SQRT2 / (d[1] * inv_n + 1e-8);
    const double tx1 = -s1 * mx1, ty1 = -s1 * my1, tx2 = -s2 * mx2, ty2 = -s2 * my2;
    const size_t m2 = 2 * (size_t)n;
    for (int i = tid; i < n; i += NPT_THREADS) {
        const double u1 = s1 * x1[2 * i] + tx1, v1 = s1 * x1[2 * i + 1] + ty1;
        const double u2 = s2 * x2[2 * i] + tx2, v2 = s2 * x2[2 * i + 1] + ty2;
        const size_t r0 = 2 * (size_t)i, r1 = r0 + 1;
        A[0 * m2 + r0] = 0.0; A[1 * m2 + r0] = 0.0; A[2 * m2 + r0] = 0.0;
        A[3 * m2 + r0] = -u1; A[4 * m2 + r0] = -v1; A[5 * m2 + r0] = -1.0;
        A[6 * m2 + r0] = v2 * u1; A[7 * m2 + r0] = v2 * v1; A[8 * m2 + r0] = v2;
        A[0 * m2 + r1] = u1; A[1 * m2 + r1] = v1; A[2 * m2 + r1] = 1.0;
        A[3 * m2 + r1] = 0.0; A[4 * m2 + r1] = 0.0; A[5 * m2 + r1] = 0.0;
        A[6 * m2 + r1] = -u2 * u1; A[7 * m2 + r1] = -u2 * v1; A[8 * m2 + r1] = -u2;
    }
    if (tid < 81) V[tid / 9][tid % 9] = (tid / 9 == tid % 9) ? 1.0 : 0.0;
    __syncthreads();
    const int jmin = block_jacobi_null9(A, (int)m2, V, sred);
    if (tid == 0) {
        double f[9], g[9], o[9];
        for (int r = 0; r < 9; ++r) f[r] = V[r][jmin];
        for (int r = 0; r < 3; ++r) {
            g[3 * r + 0] = f[3 * r + 0] * s1;
            g[3 * r + 1] = f[3 * r + 1] * s1;
            g[3 * r + 2] = f[3 * r + 0] * tx1 + f[3 * r + 1] * ty1 + f[3 * r + 2];
        }
        const double is2 = 1.0 / s2;
        for (int c = 0; c < 3; ++c) {
            o[0 + c] = is2 * g[0 + c] + mx2 * g[6 + c];
            o[3 + c] = is2 * g[3 + c] + my2 * g[6 + c];
            o[6 + c] = g[6 + c];
        }
        const double inv = o[8];
        for (int i = 0; i < 9; ++i) Hout[i] = o[i] / inv;
    }
}

}  // namespace sfm

using namespace sfm;

extern "C" size_t sfm_fundamental_npt_workspace_bytes(int32_t n) {
    return n > 0 ? sizeof(double) * 9 * (size_t)n : 0;
}

extern "C" int sfm_fundamental_npt(const double *d_x1, const double *d_x2, int32_t n, double *d_F,
                                   void *d_ws, size_t ws_bytes, void *stream) {
    if (!d_x1 || !d_x2 || !d_F || !d_ws || n < 8) return SFM_E_INVALID;
    if (ws_bytes < sfm_fundamental_npt_workspace_bytes(n)) return SFM_E_WORKSPACE;
    fundamental_npt_kernel<<<1, NPT_THREADS, 0, as_stream(stream)>>>(
        d_x1, d_x2, n, reinterpret_cast<double *>(d_ws), d_F);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

// ---- allocation-free variant (no counterpart in the reference) ----
// Callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc) have no
// workspace to hand over, so the 9 x n design matrix of fundamental_npt_kernel lives in a file-scope device array
// of fixed size: 9 x 16384 doubles = 1.18 MB, held by every GPU that loads the library until the process exits.
// The array's NAME is a host-side stand-in in host code, not a device address: the pointer the kernel gets comes
// from cudaGetSymbolAddress, looked up once per device with that device current. All calls on a device share the
// array (it keeps the last call's contents; the kernel writes every element it later reads, so nothing has to be
// cleared), hence two calls that overlap on the GPU race.
namespace sfm {
void record_error_text(const char *msg);

__device__ __align__(256) static double g_npt_static_scratch[9 * (size_t)SFM_FUNDAMENTAL_NPT_STATIC_MAX_N];

constexpr int NPT_MAX_DEVICES = 64;
static std::mutex g_npt_mu;
static double *g_npt_scratch_ptr[NPT_MAX_DEVICES] = {};

// device address of the scratch array on the CURRENT device (set up on first use); nullptr on failure
static double *npt_static_scratch_ptr() {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { record_cuda_error(e, "cudaGetDevice(sfm_fundamental_npt_static)"); return nullptr; }
    if (dev < 0 || dev >= NPT_MAX_DEVICES) { record_error_text("sfm_fundamental_npt_static: device ordinal out of range"); return nullptr; }
    std::lock_guard<std::mutex> g(g_npt_mu);
    if (!g_npt_scratch_ptr[dev]) {
        double *ptr = nullptr;
        e = cudaGetSymbolAddress((void **)&ptr, g_npt_static_scratch);
        if (e != cudaSuccess || !ptr) {
            record_cuda_error(e, "cudaGetSymbolAddress(g_npt_static_scratch)");
            return nullptr;
        }
        g_npt_scratch_ptr[dev] = ptr;
    }
    return g_npt_scratch_ptr[dev];
}
}  // namespace sfm

extern "C" int sfm_fundamental_npt_static(const double *d_x1, const double *d_x2, int32_t n, double *d_F,
                                          void *stream) {
    if (!d_x1 || !d_x2 || !d_F || n < 8) return SFM_E_INVALID;
    if (n > SFM_FUNDAMENTAL_NPT_STATIC_MAX_N) {
        char msg[160];
        snprintf(msg, sizeof(msg), "sfm_fundamental_npt_static: n = %d exceeds the static scratch maximum of %d "
                 "correspondences; use sfm_fundamental_npt with a caller workspace", (int)n,
                 (int)SFM_FUNDAMENTAL_NPT_STATIC_MAX_N);
        record_error_text(msg);
        return SFM_E_WORKSPACE;
    }
    double *scratch = npt_static_scratch_ptr();
    if (!scratch) return SFM_E_CUDA;
    fundamental_npt_kernel<<<1, NPT_THREADS, 0, as_stream(stream)>>>(d_x1, d_x2, n, scratch, d_F);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" size_t sfm_homography_npt_workspace_bytes(int32_t n) {
    return n > 0 ? sizeof(double) * 18 * (size_t)n : 0;
}

extern "C" int sfm_homography_npt(const double *d_x1, const double *d_x2, int32_t n, double *d_H, void *d_ws,
                                  size_t ws_bytes, void *stream) {
    if (!d_x1 || !d_x2 || !d_H || !d_ws || n < 4) return SFM_E_INVALID;
    if (ws_bytes < sfm_homography_npt_workspace_bytes(n)) return SFM_E_WORKSPACE;
    homography_npt_kernel<<<1, NPT_THREADS, 0, as_stream(stream)>>>(d_x1, d_x2, n, reinterpret_cast<double *>(d_ws), d_H);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}
