// pnp.cu -- K6 (PnP-RANSAC consensus scoring) and K7 (batched non-linear PnP).
//
//   K6 pnp_score_kernel  PnPRANSAC.py:59-69: reprojection test of every hypothesis pose against
//                        every 3D-2D correspondence, strict '<', +1e-8 in the denominator.
//   K7 nlpnp_kernel      NonlinearPnP.py:5-44,97-116: 6-parameter (rotvec, t) LM refinement,
//                        one 128-thread block per registration, residual of NonLinearPnPLoss.
#include "lm_small.cuh"
#include "pose_math.cuh"

namespace sfm {

// ---- K7 arithmetic ------------------------------------------------------------------------------
SFM_HD void pnp_zero(NormalEq<6> &ne) {
#pragma unroll
    for (int i = 0; i < 21; ++i) ne.A[i] = 0.0;
#pragma unroll
    for (int i = 0; i < 6; ++i) ne.g[i] = 0.0;
    ne.cost = 0.0;
}

SFM_HD void pnp_accumulate_point(const CamLin &c, const double *K, const double *X, double ox,
                                 double oy, NormalEq<6> &ne) {
    double ru, rv, jc[2][6], jp[2][3];
    reproj_residual(c, K, X, ox, oy, 1e-8, ru, rv, jc, jp);
    ne.cost += 0.5 * (ru * ru + rv * rv);
#pragma unroll
    for (int r = 0; r < 6; ++r) {
        ne.g[r] += jc[0][r] * ru + jc[1][r] * rv;
#pragma unroll
        for (int q = 0; q <= r; ++q) ne.A[sym_idx(r, q)] += jc[0][r] * jc[0][q] + jc[1][r] * jc[1][q];
    }
}

// serial evaluation (host simulation and tiny problems)
SFM_HD void pnp_eval_serial(const double *K, const double *rt, const double *X, const double *x,
                            int n, NormalEq<6> &ne) {
    CamLin c;
    cam_linearise(K, rt, c);
    pnp_zero(ne);
    for (int i = 0; i < n; ++i) pnp_accumulate_point(c, K, X + 3 * i, x[2 * i], x[2 * i + 1], ne);
}

SFM_HD int pnp_solve_serial(const double *K, double *rt, const double *X, const double *x, int n,
                            int max_iter, double tol, double *cost_out, int *iters_out) {
    NormalEq<6> ne, net;
    LMState<6> s;
    for (int j = 0; j < 6; ++j) s.x[j] = rt[j];
    pnp_eval_serial(K, s.x, X, x, n, ne);
    lm_init<6>(s, ne);
    while (s.status == LM_RUNNING) {
        if (s.need_outer) {
            lm_begin_outer<6>(s, ne, tol);
            if (s.status != LM_RUNNING) break;
        }
        double p[6], xt[6];
        lm_propose<6>(s, ne, p, xt);
        pnp_eval_serial(K, xt, X, x, n, net);
        lm_judge<6>(s, ne, net, p, xt, tol, tol, max_iter);
    }
    bool ok = s.status != LM_FAILED;
    for (int j = 0; j < 6; ++j) ok = ok && is_finite(s.x[j]);
    if (ok) for (int j = 0; j < 6; ++j) rt[j] = s.x[j];
    if (cost_out) *cost_out = ne.cost;
    if (iters_out) *iters_out = s.nfev;
    return ok ? s.status : LM_FAILED;
}

#if defined(__CUDACC__)
constexpr int PNP_THREADS = 128;

// block-wide sum of the 28 numbers of a NormalEq<6>; every thread receives the identical total
__device__ __forceinline__ void pnp_block_reduce(NormalEq<6> &ne, double *sred /*[4][28]*/) {
    double v[28];
#pragma unroll
    for (int i = 0; i < 21; ++i) v[i] = ne.A[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) v[21 + i] = ne.g[i];
    v[27] = ne.cost;
#pragma unroll
    for (int i = 0; i < 28; ++i) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < 28; ++i) sred[warp * 28 + i] = v[i];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 28; ++i) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < PNP_THREADS / 32; ++w) s += sred[w * 28 + i];
        v[i] = s;
    }
#pragma unroll
    for (int i = 0; i < 21; ++i) ne.A[i] = v[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) ne.g[i] = v[21 + i];
    ne.cost = v[27];
}

struct KMat { double k[9]; };

__device__ __forceinline__ void pnp_eval_block(const KMat &K, const double *rt, const double *X,
                                               const double *x, int n, NormalEq<6> &ne, double *sred) {
    CamLin c;
    cam_linearise(K.k, rt, c);
    pnp_zero(ne);
    for (int i = threadIdx.x; i < n; i += PNP_THREADS) {
        double Xi[3] = {X[3 * (size_t)i], X[3 * (size_t)i + 1], X[3 * (size_t)i + 2]};
        double2 o = __ldg(reinterpret_cast<const double2 *>(x) + i);
        pnp_accumulate_point(c, K.k, Xi, o.x, o.y, ne);
    }
    pnp_block_reduce(ne, sred);
}

__global__ void __launch_bounds__(PNP_THREADS)
nlpnp_kernel(KMat K, const double *__restrict__ Xall, const double *__restrict__ xall,
             const int32_t *__restrict__ offsets, const double *__restrict__ rt0, int max_iter,
             double tol, double *__restrict__ rt_out, double *__restrict__ cost_out,
             int32_t *__restrict__ status_out) {
    __shared__ double sred[(PNP_THREADS / 32) * 28];
    const int prob = blockIdx.x;
    const int beg = offsets[prob], n = offsets[prob + 1] - beg;
    const double *X = Xall + 3 * (size_t)beg;
    const double *x = xall + 2 * (size_t)beg;   // 16-byte aligned: beg * 16 bytes
    NormalEq<6> ne, net;
    LMState<6> s;
#pragma unroll
    for (int j = 0; j < 6; ++j) s.x[j] = rt0[6 * (size_t)prob + j];
    pnp_eval_block(K, s.x, X, x, n, ne, sred);
    lm_init<6>(s, ne);
    if (n < 4) s.status = LM_FAILED;            // NonlinearPnP.py:82-84: return the input pose
    while (s.status == LM_RUNNING) {
        if (s.need_outer) {
            lm_begin_outer<6>(s, ne, tol);
            if (s.status != LM_RUNNING) break;
        }
        double p[6], xt[6];
        lm_propose<6>(s, ne, p, xt);
        pnp_eval_block(K, xt, X, x, n, net, sred);
        lm_judge<6>(s, ne, net, p, xt, tol, tol, max_iter);
    }
    if (threadIdx.x == 0) {
        bool ok = s.status != LM_FAILED;
#pragma unroll
        for (int j = 0; j < 6; ++j) ok = ok && is_finite(s.x[j]);
#pragma unroll
        for (int j = 0; j < 6; ++j) rt_out[6 * (size_t)prob + j] = ok ? s.x[j] : rt0[6 * (size_t)prob + j];
        if (cost_out) cost_out[prob] = ne.cost;
        if (status_out) status_out[prob] = ((ok ? s.status : (int)LM_FAILED) << 16) | (s.nfev & 0xFFFF);
    }
}

// ---- K6 ------------------------------------------------------------------------------------------
// thread = hypothesis (its 3x4 P in registers); a block stages a chunk of correspondences
// (X,Y,Z,u,v = 40 B) in shared memory and reads them back as broadcasts.
//   err = sqrt((u - a/d)^2 + (v - b/d)^2) < thr ,  d = w + 1e-8           (PnPRANSAC.py:62-68)
//   <=>  (u d - a)^2 + (v d - b)^2 < thr^2 d^2   and thr > 0               (division free)
constexpr int K6_THREADS = 256;
constexpr int K6_CHUNK = 512;

__global__ void __launch_bounds__(K6_THREADS)
pnp_score_kernel(const double *__restrict__ Pall, const double *__restrict__ X,
                 const double *__restrict__ x, int n, int h, double thr, int32_t *__restrict__ counts,
                 const int32_t *__restrict__ offsets = nullptr) {
    __shared__ double sX[K6_CHUNK * 3];
    __shared__ double sx[K6_CHUNK * 2];
    if (offsets) {
        // batched form (sfm_pnp_score_count_batch): blockIdx.z = registration, its rows, hypotheses and counts
        const int b = blockIdx.z, o = offsets[b];
        n = offsets[b + 1] - o;
        X += 3 * (size_t)o;
        x += 2 * (size_t)o;
        Pall += (size_t)b * h * 12;
        counts += (size_t)b * h;
    }
    const int p0 = blockIdx.y * K6_CHUNK;
    if (p0 >= n) return;                                   // block-uniform: before any barrier
    const int np = min(K6_CHUNK, n - p0);
    for (int i = threadIdx.x; i < np * 3; i += K6_THREADS) sX[i] = X[3 * (size_t)p0 + i];
    for (int i = threadIdx.x; i < np * 2; i += K6_THREADS) sx[i] = x[2 * (size_t)p0 + i];
    const int hyp = blockIdx.x * K6_THREADS + threadIdx.x;
    double P[12];
    {
        const double *p = Pall + (size_t)min(hyp, h - 1) * 12;
#pragma unroll
        for (int i = 0; i < 12; ++i) P[i] = __ldg(p + i);
    }
    __syncthreads();
    const double thr2 = thr > 0.0 ? thr * thr : -1.0;   // thr <= 0 (or NaN): nothing is an inlier
    int cnt = 0;
#pragma unroll 4
    for (int i = 0; i < np; ++i) {
        double X0 = sX[3 * i], X1 = sX[3 * i + 1], X2 = sX[3 * i + 2];
        double a = fma(P[0], X0, fma(P[1], X1, fma(P[2], X2, P[3])));
        double b = fma(P[4], X0, fma(P[5], X1, fma(P[6], X2, P[7])));
        double w = fma(P[8], X0, fma(P[9], X1, fma(P[10], X2, P[11])));
        double d = w + 1e-8;
        double eu = fma(sx[2 * i], d, -a);
        double ev = fma(sx[2 * i + 1], d, -b);
        double lhs = fma(eu, eu, ev * ev);
        double rhs = thr2 * (d * d);
        cnt += (lhs < rhs) ? 1 : 0;
    }
    if (hyp < h && cnt) atomicAdd(&counts[hyp], cnt);
}
#endif

}  // namespace sfm

#if defined(__CUDACC__)
using namespace sfm;

extern "C" int sfm_pnp_score_count(const double *d_P, const double *d_X, const double *d_x, int32_t n,
                                   int32_t h, double threshold, int32_t *d_counts, void *stream) {
    if (!d_P || !d_X || !d_x || !d_counts || n < 0 || h < 0) return SFM_E_INVALID;
    if (h == 0) return SFM_OK;
    cudaStream_t st = as_stream(stream);
    SFM_CUDA_TRY(cudaMemsetAsync(d_counts, 0, sizeof(int32_t) * (size_t)h, st));
    if (n == 0) return SFM_OK;
    dim3 grid(ceil_div(h, K6_THREADS), ceil_div(n, K6_CHUNK));
    pnp_score_kernel<<<grid, K6_THREADS, 0, st>>>(d_P, d_X, d_x, n, h, threshold, d_counts);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_pnp_score_count_batch(const double *d_P, const double *d_X, const double *d_x,
                                         const int32_t *d_offsets, int32_t n_regs, int32_t max_n, int32_t h_per_reg,
                                         double threshold, int32_t *d_counts, void *stream) {
    if (!d_P || !d_X || !d_x || !d_offsets || !d_counts || n_regs < 0 || max_n < 0 || h_per_reg < 0) return SFM_E_INVALID;
    if (n_regs == 0 || h_per_reg == 0) return SFM_OK;
    if (n_regs > 65535) return SFM_E_INVALID;             // gridDim.z
    cudaStream_t st = as_stream(stream);
    SFM_CUDA_TRY(cudaMemsetAsync(d_counts, 0, sizeof(int32_t) * (size_t)n_regs * h_per_reg, st));
    if (max_n == 0) return SFM_OK;
    dim3 grid(ceil_div(h_per_reg, K6_THREADS), ceil_div(max_n, K6_CHUNK), n_regs);
    pnp_score_kernel<<<grid, K6_THREADS, 0, st>>>(d_P, d_X, d_x, 0, h_per_reg, threshold, d_counts, d_offsets);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_nlpnp_batch(const double *h_K, const double *d_X, const double *d_x,
                               const int32_t *d_offsets, int32_t n_problems, const double *d_rt0,
                               int32_t max_iter, double tol, double *d_rt, double *d_cost,
                               int32_t *d_status, void *stream) {
    if (!h_K || !d_X || !d_x || !d_offsets || !d_rt0 || !d_rt || n_problems < 0 || max_iter < 0)
        return SFM_E_INVALID;
    if (n_problems == 0) return SFM_OK;
    KMat K;
    for (int i = 0; i < 9; ++i) K.k[i] = h_K[i];
    nlpnp_kernel<<<n_problems, PNP_THREADS, 0, as_stream(stream)>>>(K, d_X, d_x, d_offsets, d_rt0, max_iter,
                                                                  tol, d_rt, d_cost, d_status);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}
#endif
