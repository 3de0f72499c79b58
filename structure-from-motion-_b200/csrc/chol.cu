// chol.cu -- K11: blocked right-looking Cholesky factorisation and solve of the reduced camera
// system S dc = g_red, as ONE cooperative persistent kernel (grid-wide barriers between panel
// steps instead of ~3 launches per panel).
//
// In the reference this solve is hidden inside MINPACK (dense pivoted QR of the full Jacobian,
// BundleAdjustment.py:205-212 -> scipy least_squares(method='lm') -> lmder/lmpar/qrsolv); here it
// is the dense SPD solve of the Schur complement.
//
// Layout: A is row-major with leading dimension ld; only the lower triangle of its leading
// npad x npad part is read or written (npad = multiple of NB; rows/cols >= n are identity
// padding supplied by the caller). The right-hand side rides along as block-row nb (rows
// npad .. npad+NB-1, one meaningful row): it goes through the panel triangular solves and
// the trailing updates like any other block-row, which leaves y = L^-1 g in it for free; the
// backward substitution L^T x = y then runs block by block from the bottom.
//
// Per panel step k (two grid barriers):
//   phase 1  every CTA factors the NB x NB diagonal block in shared memory (redundantly: cheaper
//            than a barrier), inverts the factor, and solves its share of the block-rows below:
//            A_ik <- A_ik L_kk^-T as a dense product with the inverse;
//   phase 2  trailing update A_ij -= L_ik L_jk^T over the lower-triangle tiles, round-robin.
#include "sfm_common.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace sfm {

constexpr int CH_NB = 48;          // panel width: a multiple of 6 (camera blocks) and of 8
constexpr int CH_LDS = CH_NB + 1;  // padded shared-memory row (conflict-free column access)
constexpr int CH_THREADS = 256;

struct CholSmem {
    double L[CH_NB * CH_LDS];      // diagonal block / left operand
    double B[CH_NB * CH_LDS];      // inverse of the diagonal factor / right operand
    double C[CH_NB * CH_LDS];      // tile being solved or updated
    double x[CH_NB];
    int fail;
};

// load an NB x NB block (rows r0.., cols c0..) of the row-major matrix into padded smem
__device__ __forceinline__ void load_block(double *s, const double *A, size_t ld, int r0, int c0) {
    for (int e = threadIdx.x; e < CH_NB * CH_NB; e += CH_THREADS) {
        int r = e / CH_NB, c = e - r * CH_NB;
        s[r * CH_LDS + c] = A[(size_t)(r0 + r) * ld + c0 + c];
    }
}
__device__ __forceinline__ void store_block(const double *s, double *A, size_t ld, int r0, int c0) {
    for (int e = threadIdx.x; e < CH_NB * CH_NB; e += CH_THREADS) {
        int r = e / CH_NB, c = e - r * CH_NB;
        A[(size_t)(r0 + r) * ld + c0 + c] = s[r * CH_LDS + c];
    }
}

// In-place Cholesky of the NB x NB block in sm.L (lower), then sm.B = L^-1 (lower).
// All threads of the CTA take part. Sets sm.fail on a non-positive pivot.
__device__ void factor_diag_block(CholSmem &sm) {
    double *L = sm.L;
    for (int j = 0; j < CH_NB; ++j) {
        __syncthreads();
        double d = L[j * CH_LDS + j];
        bool bad = !(d > 0.0) || !(d < 1.7e308);
        double piv = bad ? 1.0 : sqrt(d);
        __syncthreads();
        if (threadIdx.x == 0) {
            L[j * CH_LDS + j] = piv;
            if (bad) sm.fail = 1;
        }
        double inv = 1.0 / piv;
        for (int i = j + 1 + threadIdx.x; i < CH_NB; i += CH_THREADS) L[i * CH_LDS + j] *= inv;
        __syncthreads();
        // trailing rank-1 update of the lower triangle: (i, m) with j < m <= i
        const int rem = CH_NB - 1 - j;
        for (int e = threadIdx.x; e < rem * rem; e += CH_THREADS) {
            int a = e / rem, b = e - a * rem;
            if (b <= a) {
                int i = j + 1 + a, m = j + 1 + b;
                L[i * CH_LDS + m] -= L[i * CH_LDS + j] * L[m * CH_LDS + j];
            }
        }
    }
    __syncthreads();
    // zero the strict upper part so that products with the block need no masking
    for (int e = threadIdx.x; e < CH_NB * CH_NB; e += CH_THREADS) {
        int r = e / CH_NB, c = e - r * CH_NB;
        if (c > r) L[r * CH_LDS + c] = 0.0;
        sm.B[r * CH_LDS + c] = 0.0;
    }
    __syncthreads();
    // B = L^-1: column c by forward substitution, one thread per column
    if (threadIdx.x < CH_NB) {
        const int c = threadIdx.x;
        sm.B[c * CH_LDS + c] = 1.0 / L[c * CH_LDS + c];
        for (int i = c + 1; i < CH_NB; ++i) {
            double s = 0.0;
            for (int m = c; m < i; ++m) s += L[i * CH_LDS + m] * sm.B[m * CH_LDS + c];
            sm.B[i * CH_LDS + c] = -s / L[i * CH_LDS + i];
        }
    }
    __syncthreads();
}

// C(48x48) = or -= X * Y^T with X, Y, C in padded smem; thread (ty, tx) of a 16 x 16 layout owns
// the 3 x 3 outputs (ty + 16a, tx + 16b). MODE 0: C = X Y^T; MODE 1: C -= X Y^T.
template <int MODE>
__device__ __forceinline__ void block_xyT(const double *X, const double *Y, double *C) {
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    double acc[3][3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) acc[a][b] = 0.0;
#pragma unroll 4
    for (int m = 0; m < CH_NB; ++m) {
        double xv[3], yv[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) xv[a] = X[(ty + 16 * a) * CH_LDS + m];
#pragma unroll
        for (int b = 0; b < 3; ++b) yv[b] = Y[(tx + 16 * b) * CH_LDS + m];
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int b = 0; b < 3; ++b) acc[a][b] = fma(xv[a], yv[b], acc[a][b]);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            double *c = &C[(ty + 16 * a) * CH_LDS + tx + 16 * b];
            if (MODE == 0) *c = acc[a][b]; else *c -= acc[a][b];
        }
}

// nb = npad / NB diagonal blocks; block-row nb is the right-hand side. x_out [npad].
// Linv_ws [nb, NB*NB]: inverses of the diagonal factors (for the backward substitution).
__global__ void __launch_bounds__(CH_THREADS)
chol_solve_kernel(double *__restrict__ A, size_t ld, int nb, double *__restrict__ Linv_ws,
                  double *__restrict__ x_out, int *__restrict__ fail_flag) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    CholSmem &sm = *reinterpret_cast<CholSmem *>(smem_raw);
    cg::grid_group grid = cg::this_grid();
    const int G = gridDim.x, cta = blockIdx.x;
    if (threadIdx.x == 0) sm.fail = 0;
    __syncthreads();

    for (int k = 0; k < nb; ++k) {
        // ---- phase 1: diagonal factor (redundant per CTA) + panel solve ----
        const int rows_below = nb - k;            // block-rows k+1 .. nb (nb = RHS)
        if (cta < rows_below || cta == 0) {
            load_block(sm.L, A, ld, k * CH_NB, k * CH_NB);
            factor_diag_block(sm);
            if (cta == 0) {
                store_block(sm.L, A, ld, k * CH_NB, k * CH_NB);
                for (int e = threadIdx.x; e < CH_NB * CH_NB; e += CH_THREADS) {
                    int r = e / CH_NB, c = e - r * CH_NB;
                    Linv_ws[(size_t)k * CH_NB * CH_NB + e] = sm.B[r * CH_LDS + c];
                }
                if (threadIdx.x == 0 && sm.fail) *fail_flag = 1;
            }
            for (int i = k + 1 + cta; i <= nb; i += G) {
                __syncthreads();
                load_block(sm.C, A, ld, i * CH_NB, k * CH_NB);
                __syncthreads();
                // X = A_ik L^-T :  X[r][c] = sum_m A_ik[r][m] Linv[c][m]; result into sm.L? no:
                // sm.L holds the factor (unused from here on in this step) -> reuse as output
                block_xyT<0>(sm.C, sm.B, sm.L);
                __syncthreads();
                store_block(sm.L, A, ld, i * CH_NB, k * CH_NB);
            }
        }
        grid.sync();
        // ---- phase 2: trailing update over tiles (i, j), k < j <= i <= nb, j < nb ----
        const int m = nb - 1 - k;                 // remaining diagonal blocks
        if (m > 0) {
            // tiles: for j' in [0, m): i' in [j', m]  (i' == m is the RHS block-row)
            const int ntiles = m * (m + 1) / 2 + m;
            for (int t = cta; t < ntiles; t += G) {
                int ip, jp;
                if (t < m * (m + 1) / 2) {
                    // lower triangle of an m x m block matrix, row-major enumeration
                    ip = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
                    while ((ip + 1) * (ip + 2) / 2 <= t) ++ip;
                    while (ip * (ip + 1) / 2 > t) --ip;
                    jp = t - ip * (ip + 1) / 2;
                } else {
                    ip = m;
                    jp = t - m * (m + 1) / 2;
                }
                const int i = k + 1 + ip, j = k + 1 + jp;
                __syncthreads();
                load_block(sm.L, A, ld, i * CH_NB, k * CH_NB);
                load_block(sm.B, A, ld, j * CH_NB, k * CH_NB);
                load_block(sm.C, A, ld, i * CH_NB, j * CH_NB);
                __syncthreads();
                block_xyT<1>(sm.L, sm.B, sm.C);
                __syncthreads();
                store_block(sm.C, A, ld, i * CH_NB, j * CH_NB);
            }
        }
        grid.sync();
    }

    // ---- backward substitution L^T x = y, y = first row of the RHS block-row ----
    const double *yrow = A + (size_t)nb * CH_NB * ld;
    for (int e = cta * CH_THREADS + threadIdx.x; e < nb * CH_NB; e += G * CH_THREADS) x_out[e] = yrow[e];
    grid.sync();
    for (int k = nb - 1; k >= 0; --k) {
        // every CTA: x_k = Linv_kk^T y_k  (48 x 48 mat-vec from global Linv)
        __syncthreads();
        if (threadIdx.x < CH_NB) sm.x[threadIdx.x] = x_out[k * CH_NB + threadIdx.x];
        __syncthreads();
        double xk = 0.0;
        if (threadIdx.x < CH_NB) {
            const double *Li = Linv_ws + (size_t)k * CH_NB * CH_NB;
            const int c = threadIdx.x;
            for (int r = c; r < CH_NB; ++r) xk += Li[r * CH_NB + c] * sm.x[r];
        }
        __syncthreads();
        if (threadIdx.x < CH_NB) sm.x[threadIdx.x] = xk;
        __syncthreads();
        // y_b -= L_kb^T x_k for b < k, round-robin over CTAs; CTA (k % G) also commits x_k
        for (int b = cta; b < k; b += G) {
            if (threadIdx.x < CH_NB) {
                const int c = threadIdx.x;
                double s = 0.0;
                for (int r = 0; r < CH_NB; ++r)
                    s += A[(size_t)(k * CH_NB + r) * ld + b * CH_NB + c] * sm.x[r];
                x_out[b * CH_NB + c] -= s;
            }
        }
        grid.sync();
        if (cta == 0 && threadIdx.x < CH_NB) x_out[k * CH_NB + threadIdx.x] = sm.x[threadIdx.x];
    }
}

static int g_chol_grid = 0;

int chol_grid_size() {
    if (g_chol_grid) return g_chol_grid;
    int per_sm = 0;
    if (cudaFuncSetAttribute(chol_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)sizeof(CholSmem)) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_solve_kernel, CH_THREADS,
                                                      sizeof(CholSmem)) != cudaSuccess || per_sm < 1) {
        (void)cudaGetLastError();
        return 0;
    }
    g_chol_grid = sm_count() * (per_sm > 2 ? 2 : per_sm);
    return g_chol_grid;
}

// Enqueue the factor+solve. A: (npad + NB) x ld; d_Linv: nb*NB*NB doubles; d_x: npad doubles;
// d_fail: int (zeroed by the caller).
int chol_solve_launch(double *A, size_t ld, int npad, double *d_Linv, double *d_x, int *d_fail,
                      cudaStream_t st) {
    int G = chol_grid_size();
    if (G <= 0) return SFM_E_CUDA;
    int nb = npad / CH_NB;
    // no point in more CTAs than the widest phase can use
    int need = nb * (nb + 1) / 2 + nb;
    if (need < 1) need = 1;
    if (G > need) G = need;
    void *args[] = {&A, &ld, &nb, &d_Linv, &d_x, &d_fail};
    cudaError_t e = cudaLaunchCooperativeKernel((void *)chol_solve_kernel, dim3(G), dim3(CH_THREADS), args,
                                                sizeof(CholSmem), st);
    count_launch();
    if (e != cudaSuccess) {
        record_cuda_error(e, "cudaLaunchCooperativeKernel(chol_solve_kernel)");
        return SFM_E_CUDA;
    }
    return SFM_OK;
}

// ---- standalone entry (tests): pack a dense SPD system into the padded layout ----------------
__global__ void chol_pack_kernel(const double *__restrict__ S, const double *__restrict__ g, int n, int npad,
                                 size_t ld, double *__restrict__ A) {
    size_t total = (size_t)(npad + CH_NB) * ld;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        int r = (int)(e / ld), c = (int)(e - (size_t)r * ld);
        double v = 0.0;
        if (r < n && c < n) v = S[(size_t)r * n + c];
        else if (r < npad && r == c) v = 1.0;
        else if (r == npad && c < n) v = g[c];
        A[e] = v;
    }
}

}  // namespace sfm

using namespace sfm;

extern "C" size_t sfm_chol_solve_workspace_bytes(int32_t n) {
    if (n < 1) return 0;
    size_t npad = (size_t)ceil_div(n, CH_NB) * CH_NB;
    size_t bytes = (npad + CH_NB) * npad * sizeof(double);              // A
    bytes += (npad / CH_NB) * (size_t)CH_NB * CH_NB * sizeof(double);    // Linv
    bytes += npad * sizeof(double) + 256;                                // x, flag
    return bytes;
}

extern "C" int sfm_chol_solve(const double *d_S, const double *d_g, int32_t n, double *d_x,
                              int32_t *d_fail, void *d_ws, size_t ws_bytes, void *stream) {
    if (!d_S || !d_g || !d_x || !d_fail || !d_ws || n < 1) return SFM_E_INVALID;
    if (ws_bytes < sfm_chol_solve_workspace_bytes(n)) return SFM_E_WORKSPACE;
    cudaStream_t st = as_stream(stream);
    int npad = ceil_div(n, CH_NB) * CH_NB;
    size_t ld = (size_t)npad;
    double *A = reinterpret_cast<double *>(d_ws);
    double *Linv = A + (size_t)(npad + CH_NB) * ld;
    double *x = Linv + (size_t)(npad / CH_NB) * CH_NB * CH_NB;
    int *flag = reinterpret_cast<int *>(x + npad);
    SFM_CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(int), st));
    chol_pack_kernel<<<sm_count() * 4, 256, 0, st>>>(d_S, d_g, n, npad, ld, A);
    SFM_LAUNCH_CHECK();
    int rc = chol_solve_launch(A, ld, npad, Linv, x, flag, st);
    if (rc != SFM_OK) return rc;
    SFM_CUDA_TRY(cudaMemcpyAsync(d_x, x, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, st));
    SFM_CUDA_TRY(cudaMemcpyAsync(d_fail, flag, sizeof(int), cudaMemcpyDeviceToDevice, st));
    return SFM_OK;
}
