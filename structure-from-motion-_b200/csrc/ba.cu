// ba.cu -- K8..K12: Schur-complement Levenberg-Marquardt bundle adjustment on the GPU.
//
// Replaces perform_bundle_adjustment's scipy least_squares(method='lm') call
// (BundleAdjustment.py:205-212) and the residual it minimises (bundle_adjustment_residuals,
// BundleAdjustment.py:43-110; project_points :8-40). Parameter layout, residual sign/order and
// the +1e-8 denominator are the reference's (SURVEY.md 8.1-Q2,Q5); observations arrive sorted by
// point (the reference's own order, BundleAdjustment.py:164-169).
//
// Kernel chain per LM linearisation:
//   ba_cam_linearise   per camera: M = K R, K t, right Jacobian of SO(3)        (21 doubles)
//   K9p point pass     thread per observation: residual, analytic 2x6 / 2x3 Jacobians;
//                      W = Jc^T Jp streamed out as 18 coalesced planes; V = sum Jp^T Jp and
//                      g_p = sum Jp^T r by a segmented warp reduction over the point-sorted
//                      observations; 0.5|r|^2 into per-CTA partials (deterministic)
//   K9c camera pass    the same residual re-evaluated in camera-major order (index built once
//                      per problem): U = sum Jc^T Jc, g_c = sum Jc^T r per camera, block
//                      reductions, no atomics
// and per trial step (damping lambda):
//   K10 schur          warp per point: (V + lambda Dp^2)^-1 in registers, Y = W V^-1,
//                      S -= Y W^T (lower block triangle) and g_red -= Y g_p
//   K11 chol_solve     blocked Cholesky + solve of S dc = g_red (chol.cu)
//   K12 update         dp = V^-1 (g_p - W^T dc); trial parameters; predicted-reduction terms
//   K8  cost pass      0.5 |r|^2 at the trial point
// The LM controller (host, a few scalars per trial) uses MINPACK's scaling (D = running max of
// the Jacobian column norms, scipy x_scale='jac'), MINPACK's gain-ratio acceptance and its
// ftol / xtol / gtol stopping rules, with a direct damping update instead of lmpar's secular
// equation: the problem has a 7-dof gauge freedom (no camera is fixed, BundleAdjustment.py:180-199),
// so J is rank deficient and lmpar's trajectory is rounding noise; both converge to the same
// minimum (DESIGN.md "BA").
#include "ba_math.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <algorithm>
#include <mutex>
#include <vector>

namespace sfm {

int chol_solve_launch(double *A, size_t ld, int npad, double *d_Linv, double *d_x, int *d_fail,
                      cudaStream_t st, unsigned long long *d_prof);
// K11s (chol_dag.cu): structure-aware factorisation, planned once per problem
struct CholDagPlan;
CholDagPlan *chol_dag_create(const uint8_t *h_pattern, int nb, int order_mode, cudaStream_t st, int *rc);
int chol_dag_launch(CholDagPlan *p, const double *A, size_t ld, int npad, double *d_x, int *d_fail, cudaStream_t st);
void chol_dag_destroy(CholDagPlan *p, cudaStream_t st);
int chol_dag_zero_system(CholDagPlan *p, double *A, size_t ld, cudaStream_t st);
constexpr int BA_NB = 48;                 // == CH_NB in chol.cu
constexpr int BA_CHUNK = 2048;            // observations per camera-pass CTA
constexpr int K9P_WARP_SCRATCH = 9 * 33 + 18;   // doubles of shared memory per warp of K9p (T[9][33] + 34 ints, padded)
constexpr size_t BA_SMEM_MAX = 232448 - 1024;      // opt-in dynamic shared memory per CTA minus the static part
constexpr size_t BA_SMEM_CAM_LIMIT = BA_SMEM_MAX - 16 * sizeof(double) * K9P_WARP_SCRATCH;   // table + 512-thread scratch

struct KParam { double k[9]; };

// ------------------------------------------------------------------------------ small helpers
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// block-wide deterministic sum; result valid in thread 0. sbuf: >= 32 doubles.
__device__ __forceinline__ double block_sum(double v, double *sbuf) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sbuf[threadIdx.x >> 5] = v;
    __syncthreads();
    double s = 0.0;
    if (threadIdx.x == 0)
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += sbuf[w];
    return s;
}

// Segmented sum over lanes with equal (sorted, contiguous) key: afterwards the first lane of
// every run holds the run's total.
template <int N>
__device__ __forceinline__ void seg_reduce(double (&v)[N], int key, int lane) {
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        int k2 = __shfl_down_sync(0xffffffffu, key, off);
        bool take = (lane + off < 32) && (k2 == key);
#pragma unroll
        for (int i = 0; i < N; ++i) {
            double t = __shfl_down_sync(0xffffffffu, v[i], off);
            if (take) v[i] += t;
        }
    }
}

__device__ __forceinline__ void atomic_max_pos(unsigned long long *addr, double v) {
    // non-negative doubles order like their bit patterns
    if (v > 0.0) atomicMax(addr, (unsigned long long)__double_as_longlong(v));
}

// ------------------------------------------------------------------------------ index building
__global__ void ba_iota_kernel(int32_t *idx, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) idx[i] = i;
}

// CSR pointer from a sorted key array: ptr[k] = first position with key >= k, ptr[nkeys] = n.
// Also validates range and order (flag |= 1 on violation).
__global__ void ba_csr_from_sorted_kernel(const int32_t *__restrict__ keys, int n, int nkeys,
                                          int32_t *__restrict__ ptr, int *__restrict__ flag) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    int prev = i == 0 ? -1 : keys[i - 1];
    int cur = i == n ? nkeys : keys[i];
    if (i < n && (cur < 0 || cur >= nkeys)) { atomicOr(flag, 1); return; }
    if (i > 0 && i < n && cur < prev) { atomicOr(flag, 1); return; }
    if (prev < 0) prev = -1;
    for (int k = prev + 1; k <= cur && k <= nkeys; ++k) ptr[k] = i;
}

__global__ void ba_range_check_kernel(const int32_t *__restrict__ keys, int n, int nkeys, int *__restrict__ flag) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && (keys[i] < 0 || keys[i] >= nkeys)) atomicOr(flag, 1);
}

__global__ void ba_gather_cam_major_kernel(const int32_t *__restrict__ perm, const double2 *__restrict__ obs,
                                           const int32_t *__restrict__ pt_idx, int n,
                                           double2 *__restrict__ obs_c, int32_t *__restrict__ pt_c) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int o = perm[i];
    obs_c[i] = obs[o];
    pt_c[i] = pt_idx[o];
}

// ------------------------------------------------------------------------------ linearisation
__global__ void ba_cam_linearise_kernel(KParam K, const double *__restrict__ cam, int nc,
                                        double *__restrict__ camlin) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nc) return;
    CamLin cl;
    cam_linearise(K.k, cam + 6 * (size_t)c, cl);
    camlin_store(cl, camlin + (size_t)CAMLIN_STRIDE * c);
}

// K9p. Persistent grid; thread per observation (grid-stride in whole warps).
template <bool SMEM_CAMS, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
ba_point_pass_kernel(KParam K, const double *__restrict__ camlin, int nc, const double *__restrict__ pts,
                     const double2 *__restrict__ obs, const int32_t *__restrict__ cam_idx,
                     const int32_t *__restrict__ pt_idx, const int32_t *__restrict__ pt_ptr, int nobs,
                     double *__restrict__ W, double *__restrict__ V, double *__restrict__ gp,
                     double *__restrict__ costpart) {
    extern __shared__ __align__(16) double scam[];
    __shared__ double sred[32];
    const double *ctab = camlin;
    if (SMEM_CAMS) {
        for (int i = threadIdx.x; i < nc * CAMLIN_STRIDE; i += THREADS) scam[i] = camlin[i];
        __syncthreads();
        ctab = scam;
    }
    const int lane = threadIdx.x & 31;
    // per-warp scratch of the segmented sum, behind the camera table: T[9][33] and the segment starts
    double *sT = scam + (SMEM_CAMS ? (size_t)nc * CAMLIN_STRIDE : 0) + (size_t)(threadIdx.x >> 5) * K9P_WARP_SCRATCH;
    int *sSeg = reinterpret_cast<int *>(sT + 9 * 33);
    double cost = 0.0;
    const int stride = gridDim.x * THREADS;
    const int nround = (nobs + 31) & ~31;
    // software pipeline: the pass is latency bound (one wide CTA per SM when the camera table fills
    // shared memory), so the index / observation loads run two iterations ahead and the point gather,
    // whose address they produce, one iteration ahead of the arithmetic that consumes them
    int o = blockIdx.x * THREADS + threadIdx.x;
    int c_cur = 0, p_cur = -1, c_nx = 0, p_nx = -1;
    double2 ob_cur = make_double2(0.0, 0.0), ob_nx = ob_cur;
    double X[3] = {0.0, 0.0, 0.0};
    if (o < nobs) { c_cur = cam_idx[o]; p_cur = pt_idx[o]; ob_cur = obs[o]; }
    if (o + stride < nobs) { c_nx = cam_idx[o + stride]; p_nx = pt_idx[o + stride]; ob_nx = obs[o + stride]; }
    if (p_cur >= 0) { X[0] = pts[3 * (size_t)p_cur]; X[1] = pts[3 * (size_t)p_cur + 1]; X[2] = pts[3 * (size_t)p_cur + 2]; }
    for (; o < nround; o += stride) {
        const bool active = o < nobs;
        const long long o2 = (long long)o + 2LL * stride;
        int c_n2 = 0, p_n2 = -1;
        double2 ob_n2 = make_double2(0.0, 0.0);
        if (o2 < nobs) { c_n2 = cam_idx[o2]; p_n2 = pt_idx[o2]; ob_n2 = obs[o2]; }
        double Xn[3] = {0.0, 0.0, 0.0};
        if (p_nx >= 0) { Xn[0] = pts[3 * (size_t)p_nx]; Xn[1] = pts[3 * (size_t)p_nx + 1]; Xn[2] = pts[3 * (size_t)p_nx + 2]; }
        const int p = active ? p_cur : -1;
        double v[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) v[i] = 0.0;
        if (active) {
            const int c = c_cur;
            const double2 ob = ob_cur;
            CamLin cl;
            camlin_load(ctab + (size_t)CAMLIN_STRIDE * c, cl);
            double ru, rv, jc[2][6], jp[2][3];
            reproj_residual(cl, K.k, X, ob.x, ob.y, BA_EPS, ru, rv, jc, jp);
            cost += 0.5 * (ru * ru + rv * rv);
#pragma unroll
            for (int r = 0; r < 6; ++r)
#pragma unroll
                for (int a = 0; a < 3; ++a)
                    W[(size_t)(r * 3 + a) * nobs + o] = jc[0][r] * jp[0][a] + jc[1][r] * jp[1][a];
            v[0] = jp[0][0] * jp[0][0] + jp[1][0] * jp[1][0];
            v[1] = jp[0][1] * jp[0][0] + jp[1][1] * jp[1][0];
            v[2] = jp[0][1] * jp[0][1] + jp[1][1] * jp[1][1];
            v[3] = jp[0][2] * jp[0][0] + jp[1][2] * jp[1][0];
            v[4] = jp[0][2] * jp[0][1] + jp[1][2] * jp[1][1];
            v[5] = jp[0][2] * jp[0][2] + jp[1][2] * jp[1][2];
            v[6] = jp[0][0] * ru + jp[1][0] * rv;
            v[7] = jp[0][1] * ru + jp[1][1] * rv;
            v[8] = jp[0][2] * ru + jp[1][2] * rv;
        }
        c_cur = c_nx; p_cur = p_nx; ob_cur = ob_nx;
        c_nx = c_n2; p_nx = p_n2; ob_nx = ob_n2;
        X[0] = Xn[0]; X[1] = Xn[1]; X[2] = Xn[2];
        // Segmented sum of the nine per-observation values over the runs of equal point index in
        // this warp. A shuffle scan moves 9 doubles through 5 levels (~200 SHFL per warp and pass:
        // the MIO queue, not HBM, bounded the kernel); here every lane parks its values in shared
        // memory and lane (s, k) of 3 x 9 sums quantity k of segment s, three segments per round.
        const int prev = __shfl_up_sync(0xffffffffu, p, 1);
        const bool head = (lane == 0) || (prev != p);
        const unsigned heads = __ballot_sync(0xffffffffu, head);
        const int nseg = __popc(heads);
        __syncwarp();                              // the previous pass has been read
#pragma unroll
        for (int k = 0; k < 9; ++k) sT[k * 33 + lane] = v[k];
        if (head) sSeg[__popc(heads & ((1u << lane) - 1u))] = lane;
        if (lane == 0) sSeg[nseg] = 32;
        __syncwarp();
        const int sl = lane / 9, k9 = lane - 9 * sl;
        for (int s0 = 0; s0 < nseg; s0 += 3) {
            const int sg = s0 + sl;
            const bool mine = lane < 27 && sg < nseg;
            const int j0 = mine ? sSeg[sg] : 0, j1 = mine ? sSeg[sg + 1] : 0;
            const int ps = __shfl_sync(0xffffffffu, p, j0);
            if (mine && ps >= 0) {
                double acc = 0.0;
                for (int j = j0; j < j1; ++j) acc += sT[k9 * 33 + j];
                const int seg_beg = pt_ptr[ps], seg_end = pt_ptr[ps + 1];
                const int ow = o - lane;           // first observation of this warp's pass
                const bool complete = (ow + j0 == seg_beg) && (seg_end <= ow + 32);
                double *dst = k9 < 6 ? V + 6 * (size_t)ps + k9 : gp + 3 * (size_t)ps + (k9 - 6);
                if (complete) *dst = acc;
                else atomicAdd(dst, acc);
            }
        }
    }
    double s = block_sum(cost, sred);
    if (threadIdx.x == 0) costpart[blockIdx.x] = s;
}

// K8. Same traversal as K9p, residual only. r_out (may be null): [nobs,2] interleaved residuals.
// The pass moves 48 B per observation and computes almost nothing, so it is written around the
// memory pipeline: shared memory holds only the projective part of each camera record (M, K t:
// 12 of the 22 doubles, padded to 14: 112 KB at 1,000 cameras), the index/observation loads of the first two
// rounds are requested BEFORE the table is staged (the staging would otherwise run with HBM idle),
// and inside the loop the indices run two rounds and the point gather one round ahead of the
// arithmetic, as in K9p.
constexpr int K8_CAM_DOUBLES = 12;
// shared-memory record stride: 7 x 16 B. An odd number of 16-byte words spreads the per-lane
// record gather (LDS.128 at 16 B x (7 c + j)) over all eight 4-bank groups; at the natural
// 6-word stride 6c mod 8 only reaches the even groups and the gather takes twice the wavefronts.
constexpr int K8_CAM_STRIDE = 14;
template <bool SMEM_CAMS, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
ba_cost_kernel(const double *__restrict__ camlin, int nc, const double *__restrict__ pts,
               const double2 *__restrict__ obs, const int32_t *__restrict__ cam_idx,
               const int32_t *__restrict__ pt_idx, int nobs, double *__restrict__ costpart,
               double2 *__restrict__ r_out) {
    extern __shared__ __align__(16) double scam[];
    __shared__ double sred[32];
    const int stride = gridDim.x * THREADS;
    int o = blockIdx.x * THREADS + threadIdx.x;
    int c_cur = 0, p_cur = 0, c_nx = 0, p_nx = 0;
    double2 ob_cur = make_double2(0.0, 0.0), ob_nx = ob_cur;
    const bool v0 = o < nobs, v1 = o + stride < nobs;
    if (v0) { c_cur = cam_idx[o]; p_cur = pt_idx[o]; ob_cur = obs[o]; }
    if (v1) { c_nx = cam_idx[o + stride]; p_nx = pt_idx[o + stride]; ob_nx = obs[o + stride]; }
    const double *ctab = camlin;
    constexpr int CS = SMEM_CAMS ? K8_CAM_STRIDE : CAMLIN_STRIDE;
    if (SMEM_CAMS) {
        // 6 of the 11 16-byte words of every record
        const double2 *src = reinterpret_cast<const double2 *>(camlin);
        double2 *dst = reinterpret_cast<double2 *>(scam);
        for (int i = threadIdx.x; i < nc * (K8_CAM_DOUBLES / 2); i += THREADS) {
            const int c = i / (K8_CAM_DOUBLES / 2), j = i - c * (K8_CAM_DOUBLES / 2);
            dst[c * (K8_CAM_STRIDE / 2) + j] = __ldg(src + c * (CAMLIN_STRIDE / 2) + j);
        }
        ctab = scam;
    }
    double X[3] = {0.0, 0.0, 0.0};
    if (v0) { X[0] = pts[3 * (size_t)p_cur]; X[1] = pts[3 * (size_t)p_cur + 1]; X[2] = pts[3 * (size_t)p_cur + 2]; }
    if (SMEM_CAMS) __syncthreads();
    double cost = 0.0;
    for (; o < nobs; o += stride) {
        const int o2 = o + 2 * stride;
        int c_n2 = 0, p_n2 = 0;
        double2 ob_n2 = make_double2(0.0, 0.0);
        if (o2 < nobs) { c_n2 = cam_idx[o2]; p_n2 = pt_idx[o2]; ob_n2 = obs[o2]; }
        double Xn[3] = {0.0, 0.0, 0.0};
        if (o + stride < nobs) {
            Xn[0] = pts[3 * (size_t)p_nx]; Xn[1] = pts[3 * (size_t)p_nx + 1]; Xn[2] = pts[3 * (size_t)p_nx + 2];
        }
        CamLin cl;
        camlin_load_proj(ctab + (size_t)CS * c_cur, cl);
        double ru, rv;
        ba_residual_only(cl, X, ob_cur.x, ob_cur.y, ru, rv);
        cost += 0.5 * (ru * ru + rv * rv);
        if (r_out) r_out[o] = make_double2(ru, rv);
        c_cur = c_nx; c_nx = c_n2; p_nx = p_n2;
        ob_cur = ob_nx; ob_nx = ob_n2;
        X[0] = Xn[0]; X[1] = Xn[1]; X[2] = Xn[2];
    }
    double s = block_sum(cost, sred);
    if (threadIdx.x == 0) costpart[blockIdx.x] = s;
}

// K9c. One CTA per (camera, chunk of its observations) in camera-major order.
constexpr int K9C_THREADS = 128;
__global__ void __launch_bounds__(K9C_THREADS, 3)
ba_camera_pass_kernel(KParam K, const double *__restrict__ camlin, const double *__restrict__ pts,
                      const double2 *__restrict__ obs_c, const int32_t *__restrict__ pt_c,
                      const int32_t *__restrict__ chunk_cam, const int32_t *__restrict__ chunk_beg,
                      const int32_t *__restrict__ chunk_end, double *__restrict__ Upart) {
    __shared__ double sred[K9C_THREADS / 32][27];
    const int ch = blockIdx.x;
    const int c = chunk_cam[ch], beg = chunk_beg[ch], end = chunk_end[ch];
    CamLin cl;
    camlin_load(camlin + (size_t)CAMLIN_STRIDE * c, cl);
    double acc[27];
#pragma unroll
    for (int i = 0; i < 27; ++i) acc[i] = 0.0;
    // software pipeline (as in K9p): index / observation two iterations ahead, point gather one ahead
    int o = beg + threadIdx.x;
    int p_nx = -1;
    double2 ob_cur = make_double2(0.0, 0.0), ob_nx = ob_cur;
    double X[3] = {0.0, 0.0, 0.0};
    if (o < end) {
        const int p0 = pt_c[o];
        ob_cur = obs_c[o];
        X[0] = pts[3 * (size_t)p0]; X[1] = pts[3 * (size_t)p0 + 1]; X[2] = pts[3 * (size_t)p0 + 2];
    }
    if (o + K9C_THREADS < end) { p_nx = pt_c[o + K9C_THREADS]; ob_nx = obs_c[o + K9C_THREADS]; }
    for (; o < end; o += K9C_THREADS) {
        int p_n2 = -1;
        double2 ob_n2 = make_double2(0.0, 0.0);
        if (o + 2 * K9C_THREADS < end) { p_n2 = pt_c[o + 2 * K9C_THREADS]; ob_n2 = obs_c[o + 2 * K9C_THREADS]; }
        double Xn[3] = {0.0, 0.0, 0.0};
        if (p_nx >= 0) { Xn[0] = pts[3 * (size_t)p_nx]; Xn[1] = pts[3 * (size_t)p_nx + 1]; Xn[2] = pts[3 * (size_t)p_nx + 2]; }
        const double2 ob = ob_cur;
        double ru, rv, jc[2][6], jp[2][3];
        reproj_residual(cl, K.k, X, ob.x, ob.y, BA_EPS, ru, rv, jc, jp);
        ob_cur = ob_nx; ob_nx = ob_n2; p_nx = p_n2;
        X[0] = Xn[0]; X[1] = Xn[1]; X[2] = Xn[2];
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            acc[21 + r] += jc[0][r] * ru + jc[1][r] * rv;
#pragma unroll
            for (int q = 0; q <= r; ++q) acc[sym_idx(r, q)] += jc[0][r] * jc[0][q] + jc[1][r] * jc[1][q];
        }
    }
#pragma unroll
    for (int i = 0; i < 27; ++i) acc[i] = warp_sum(acc[i]);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < 27; ++i) sred[warp][i] = acc[i];
    }
    __syncthreads();
    if (threadIdx.x < 27) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < K9C_THREADS / 32; ++w) s += sred[w][threadIdx.x];
        Upart[(size_t)ch * 27 + threadIdx.x] = s;
    }
}

// Sum the chunk partials of every camera in chunk order -> U [nc,21], g_c [nc,6] (this rank's share).
// Warp per camera, lane = one of the 27 partial-sum slots (U 21 | g_c 6), chunks added in order (the same
// fixed summation order as ever). Single GPU (Dc != nullptr): U and g_c are final here, so the MINPACK
// scaling of ba_cam_scale_kernel is done in the same launch from the values the lanes already hold.
__global__ void __launch_bounds__(256)
ba_cam_finalize_kernel(const double *__restrict__ Upart, const int32_t *__restrict__ chunk_ptr, int nc,
                       double *__restrict__ U, double *__restrict__ gc, int first, double *__restrict__ Dc,
                       unsigned long long *__restrict__ gmax) {
    const int c = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (c >= nc) return;                                   // warp-uniform
    double v = 0.0;
    if (lane < 27)
        for (int ch = chunk_ptr[c]; ch < chunk_ptr[c + 1]; ++ch) v += Upart[(size_t)ch * 27 + lane];
    if (lane < 21) U[(size_t)c * 21 + lane] = v;
    else if (lane < 27) gc[(size_t)c * 6 + (lane - 21)] = v;
    if (!Dc) return;
    const int j = lane < 6 ? lane : 0;
    const double ujj = __shfl_sync(0xffffffffu, v, sym_idx(j, j)), gj = __shfl_sync(0xffffffffu, v, 21 + j);
    double gm = 0.0;
    if (lane < 6) {
        const double cn = sqrt(ujj);
        const double d = first ? (cn == 0.0 ? 1.0 : cn) : fmax(Dc[(size_t)c * 6 + lane], cn);
        Dc[(size_t)c * 6 + lane] = d;
        if (cn > 0.0) gm = fabs(gj) / cn;
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) gm = fmax(gm, __shfl_xor_sync(0xffffffffu, gm, o));
    if (lane == 0) atomic_max_pos(gmax, gm);
}

// From the (all-rank) U, g_c: the MINPACK scaling Dc (first linearisation: column norm, 1 if zero;
// later: running max) and the gtol numerator max_j |g_j| / |J_j|.
__global__ void ba_cam_scale_kernel(const double *__restrict__ U, const double *__restrict__ gc, int nc, int first,
                                    double *__restrict__ Dc, unsigned long long *__restrict__ gmax) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nc) return;
    double gm = 0.0;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        double cn = sqrt(U[(size_t)c * 21 + sym_idx(j, j)]);
        double d = first ? (cn == 0.0 ? 1.0 : cn) : fmax(Dc[(size_t)c * 6 + j], cn);
        Dc[(size_t)c * 6 + j] = d;
        if (cn > 0.0) gm = fmax(gm, fabs(gc[(size_t)c * 6 + j]) / cn);
    }
    atomic_max_pos(gmax, gm);
}

// ------------------------------------------------------------------------------ Schur complement
// After the memset of S: diagonal blocks U_c + lambda Dc^2, right-hand-side row g_c, identity on
// the padding diagonal.
__global__ void ba_schur_init_kernel(const double *__restrict__ U, const double *__restrict__ gc,
                                     const double *__restrict__ Dc, int nc, int npad, size_t ld,
                                     double lambda, double *__restrict__ S, int padding_only = 0) {
    // (multi-GPU: the camera blocks enter the all-reduced S once, from rank 0; under the sparse all-reduce
    //  the identity on the padding diagonal does not travel and every rank writes its own)
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = 6 * nc;
    if (padding_only) {
        if (t >= n && t < npad) S[(size_t)t * ld + t] = 1.0;
        return;
    }
    if (t < nc * 36) {
        int c = t / 36, e = t - 36 * c, r = e / 6, q = e - 6 * r;
        double v = U[(size_t)c * 21 + sym_idx(r, q)];
        if (r == q) {
            double d = Dc[(size_t)c * 6 + r];
            v += lambda * d * d;
        }
        S[(size_t)(6 * c + r) * ld + 6 * c + q] = v;
    }
    if (t < n) S[(size_t)npad * ld + t] = gc[t];
    if (t >= n && t < npad) S[(size_t)t * ld + t] = 1.0;
}

constexpr int SCH_KT = 32;      // observations per point on the shared-memory path
constexpr int SCH_WARPS = 4;

__device__ __forceinline__ double sym3_at(const double *s, int r, int c) { return s[sym_idx(r, c)]; }

// K10: warp per point.
__global__ void __launch_bounds__(SCH_WARPS * 32)
ba_schur_kernel(const double *__restrict__ W, int nobs, const int32_t *__restrict__ cam_idx,
                const int32_t *__restrict__ pt_ptr, int npts, const double *__restrict__ V,
                const double *__restrict__ gp, double *__restrict__ Dp, double lambda, int update_D,
                int first, double *__restrict__ Vinv, double *__restrict__ bp, double *__restrict__ S,
                size_t ld, double *__restrict__ gred, unsigned long long *__restrict__ gmax,
                int *__restrict__ fail) {
    __shared__ double sW[SCH_WARPS][SCH_KT * 18];
    __shared__ double sY[SCH_WARPS][SCH_KT * 18];
    __shared__ int sC[SCH_WARPS][SCH_KT];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *w_ = sW[warp], *y_ = sY[warp];
    int *c_ = sC[warp];
    for (int p = blockIdx.x * SCH_WARPS + warp; p < npts; p += gridDim.x * SCH_WARPS) {
        double v[6], g[3], d[3];
#pragma unroll
        for (int i = 0; i < 6; ++i) v[i] = V[6 * (size_t)p + i];
#pragma unroll
        for (int i = 0; i < 3; ++i) g[i] = gp[3 * (size_t)p + i];
        if (update_D) {
            double gm = 0.0;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                double cn = sqrt(v[sym_idx(j, j)]);
                d[j] = first ? (cn == 0.0 ? 1.0 : cn) : fmax(Dp[3 * (size_t)p + j], cn);
                if (cn > 0.0) gm = fmax(gm, fabs(g[j]) / cn);
            }
            if (lane == 0) {
#pragma unroll
                for (int j = 0; j < 3; ++j) Dp[3 * (size_t)p + j] = d[j];
                atomic_max_pos(gmax, gm);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 3; ++j) d[j] = Dp[3 * (size_t)p + j];
        }
        __syncwarp();
        v[0] += lambda * d[0] * d[0];
        v[2] += lambda * d[1] * d[1];
        v[5] += lambda * d[2] * d[2];
        double vi[6];
        if (!spd3_inverse(v, vi)) {
            if (lane == 0) atomicOr(fail, 2);
#pragma unroll
            for (int i = 0; i < 6; ++i) vi[i] = 0.0;
        }
        double b[3];
        sym3_mul(vi, g, b);
        if (lane == 0) {
#pragma unroll
            for (int i = 0; i < 6; ++i) Vinv[6 * (size_t)p + i] = vi[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) bp[3 * (size_t)p + i] = b[i];
        }
        const int o0 = pt_ptr[p], k = pt_ptr[p + 1] - o0;
        if (k <= 0) continue;
        if (k <= SCH_KT) {
            for (int e = lane; e < 18 * k; e += 32) {
                int rc = e / k, i = e - rc * k;
                w_[i * 18 + rc] = W[(size_t)rc * nobs + o0 + i];
            }
            for (int i = lane; i < k; i += 32) c_[i] = cam_idx[o0 + i];
            __syncwarp();
            for (int e = lane; e < 6 * k; e += 32) {
                int i = e / 6, r = e - 6 * i;
                double w0 = w_[i * 18 + 3 * r], w1 = w_[i * 18 + 3 * r + 1], w2 = w_[i * 18 + 3 * r + 2];
                y_[i * 18 + 3 * r + 0] = w0 * vi[0] + w1 * vi[1] + w2 * vi[3];
                y_[i * 18 + 3 * r + 1] = w0 * vi[1] + w1 * vi[2] + w2 * vi[4];
                y_[i * 18 + 3 * r + 2] = w0 * vi[3] + w1 * vi[4] + w2 * vi[5];
                atomicAdd(&gred[6 * c_[i] + r], -(w0 * b[0] + w1 * b[1] + w2 * b[2]));
            }
            __syncwarp();
            for (int i = 0; i < k; ++i) {
                const int ci = c_[i];
                for (int idx = lane; idx < (i + 1) * 36; idx += 32) {
                    int j = idx / 36, el = idx - 36 * j, r = el / 6, q = el - 6 * r;
                    const int cj = c_[j];
                    const double *yr = y_ + i * 18 + 3 * r, *wq = w_ + j * 18 + 3 * q;
                    double val = -(yr[0] * wq[0] + yr[1] * wq[1] + yr[2] * wq[2]);
                    if (j == i) {                  // own diagonal block: K11 reads its lower triangle only
                        if (q <= r) atomicAdd(&S[(size_t)(6 * ci + r) * ld + 6 * ci + q], val);
                    } else if (ci > cj) {
                        atomicAdd(&S[(size_t)(6 * ci + r) * ld + 6 * cj + q], val);
                    } else if (ci < cj) {
                        atomicAdd(&S[(size_t)(6 * cj + q) * ld + 6 * ci + r], val);
                    } else {   // the same camera twice on one point: both orders land in the block
                        atomicAdd(&S[(size_t)(6 * ci + r) * ld + 6 * ci + q], val);
                        atomicAdd(&S[(size_t)(6 * ci + q) * ld + 6 * ci + r], val);
                    }
                }
            }
            __syncwarp();
        } else {
            // long tracks: same arithmetic straight from global memory (L1/L2 resident)
            for (int e = lane; e < 6 * k; e += 32) {
                int i = e / 6, r = e - 6 * i;
                size_t o = (size_t)o0 + i;
                double w0 = W[(size_t)(3 * r) * nobs + o], w1 = W[(size_t)(3 * r + 1) * nobs + o],
                       w2 = W[(size_t)(3 * r + 2) * nobs + o];
                atomicAdd(&gred[6 * cam_idx[o] + r], -(w0 * b[0] + w1 * b[1] + w2 * b[2]));
            }
            for (int i = 0; i < k; ++i) {
                const int ci = cam_idx[o0 + i];
                for (long long idx = lane; idx < (long long)(i + 1) * 36; idx += 32) {
                    int j = (int)(idx / 36), el = (int)(idx - 36LL * j), r = el / 6, q = el - 6 * r;
                    const int cj = cam_idx[o0 + j];
                    size_t oi = (size_t)o0 + i, oj = (size_t)o0 + j;
                    double w0 = W[(size_t)(3 * r) * nobs + oi], w1 = W[(size_t)(3 * r + 1) * nobs + oi],
                           w2 = W[(size_t)(3 * r + 2) * nobs + oi];
                    double y0 = w0 * vi[0] + w1 * vi[1] + w2 * vi[3];
                    double y1 = w0 * vi[1] + w1 * vi[2] + w2 * vi[4];
                    double y2 = w0 * vi[3] + w1 * vi[4] + w2 * vi[5];
                    double val = -(y0 * W[(size_t)(3 * q) * nobs + oj] + y1 * W[(size_t)(3 * q + 1) * nobs + oj] +
                                   y2 * W[(size_t)(3 * q + 2) * nobs + oj]);
                    if (j == i) {
                        if (q <= r) atomicAdd(&S[(size_t)(6 * ci + r) * ld + 6 * ci + q], val);
                    } else if (ci > cj) {
                        atomicAdd(&S[(size_t)(6 * ci + r) * ld + 6 * cj + q], val);
                    } else if (ci < cj) {
                        atomicAdd(&S[(size_t)(6 * cj + q) * ld + 6 * ci + r], val);
                    } else {
                        atomicAdd(&S[(size_t)(6 * ci + r) * ld + 6 * ci + q], val);
                        atomicAdd(&S[(size_t)(6 * ci + q) * ld + 6 * ci + r], val);
                    }
                }
            }
        }
    }
}

// K10a: thread per point -- MINPACK scale D_p (running maximum of the column norms), damped
// V + lambda D_p^2, its inverse and b_p = V^-1 g_p; the gtol numerator max |g_j| / |J_j|.
__global__ void __launch_bounds__(256)
ba_point_prep_kernel(const double *__restrict__ V, const double *__restrict__ gp, double *__restrict__ Dp, int npts,
                     double lambda, int update_D, int first, double *__restrict__ Vinv, double *__restrict__ bp,
                     unsigned long long *__restrict__ gmax, int *__restrict__ fail, double *__restrict__ Ginv = nullptr) {
    const int p = blockIdx.x * 256 + threadIdx.x;
    double gm = 0.0;
    if (p < npts) {
        double v[6], g[3], d[3];
#pragma unroll
        for (int i = 0; i < 6; ++i) v[i] = V[6 * (size_t)p + i];
#pragma unroll
        for (int i = 0; i < 3; ++i) g[i] = gp[3 * (size_t)p + i];
        if (update_D) {
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                double cn = sqrt(v[sym_idx(j, j)]);
                d[j] = first ? (cn == 0.0 ? 1.0 : cn) : fmax(Dp[3 * (size_t)p + j], cn);
                if (cn > 0.0) gm = fmax(gm, fabs(g[j]) / cn);
                Dp[3 * (size_t)p + j] = d[j];
            }
        } else {
#pragma unroll
            for (int j = 0; j < 3; ++j) d[j] = Dp[3 * (size_t)p + j];
        }
        v[0] += lambda * d[0] * d[0];
        v[2] += lambda * d[1] * d[1];
        v[5] += lambda * d[2] * d[2];
        double vi[6], b[3];
        if (!spd3_inverse(v, vi)) {
            atomicOr(fail, 2);
#pragma unroll
            for (int i = 0; i < 6; ++i) vi[i] = 0.0;
        }
        sym3_mul(vi, g, b);
#pragma unroll
        for (int i = 0; i < 6; ++i) Vinv[6 * (size_t)p + i] = vi[i];
#pragma unroll
        for (int i = 0; i < 3; ++i) bp[3 * (size_t)p + i] = b[i];
        if (Ginv) {
            // G = C^-1 with (V + lambda Dp^2) = C C^T, so V^-1 = G^T G (K10g works with Z = W G^T).
            // vi == 0 marks a failed factorisation: G = 0 gives the same zero contribution.
            double gi[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
            if (vi[0] != 0.0 || vi[2] != 0.0 || vi[5] != 0.0) {
                const double l00 = sqrt(v[0]), g00 = 1.0 / l00;
                const double l10 = v[1] * g00, l20 = v[3] * g00;
                const double l11 = sqrt(v[2] - l10 * l10), g11 = 1.0 / l11;
                const double l21 = (v[4] - l20 * l10) * g11;
                const double l22 = sqrt(v[5] - l20 * l20 - l21 * l21), g22 = 1.0 / l22;
                gi[0] = g00;
                gi[1] = -l10 * g00 * g11;
                gi[2] = g11;
                gi[4] = -l21 * g11 * g22;
                gi[3] = -(l20 * g00 + l21 * gi[1]) * g22;
                gi[5] = g22;
            }
#pragma unroll
            for (int i = 0; i < 6; ++i) Ginv[6 * (size_t)p + i] = gi[i];
            // h_p = G g_p: W_o b_p = W_o G^T G g_p = Z_o h_p, so the right-hand side comes out of the Z records too
            double *H = Ginv + 6 * (size_t)npts;
            H[3 * (size_t)p + 0] = gi[0] * g[0];
            H[3 * (size_t)p + 1] = gi[1] * g[0] + gi[2] * g[1];
            H[3 * (size_t)p + 2] = gi[3] * g[0] + gi[4] * g[1] + gi[5] * g[2];
        }
    }
    if (update_D) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) gm = fmax(gm, __shfl_xor_sync(0xffffffffu, gm, o));
        if ((threadIdx.x & 31) == 0) atomic_max_pos(gmax, gm);
    }
}

// ---- K10g: the Schur complement as a GATHER over a plan-time pair list --------------------------
// The point-major K10 sends 36 FP64 atomics to L2 for every camera pair of every point and runs
// at the L2 atomic ceiling (~280 G adds/s on B200). Which observation pairs meet in which block
// of S depends only on the index structure, so it is worked out once per problem:
//   plan time : every pair (o_i, o_j) of observations of one point, keyed by its block
//               (hi camera, lo camera), sorted by key with one CUB radix sort
//   per trial : Z_o = W_o G_p^T per observation (G_p = inverse Cholesky factor of the damped V_p,
//               so W_i V^-1 W_j^T = Z_i Z_j^T and ONE record type serves both sides), then a thread
//               per slice of the sorted list accumulates Z_i Z_j^T in registers and sends 36 (21
//               on the diagonal) atomics per (slice, block) instead of per (point, pair).
constexpr int SG_THREADS = 128;

constexpr uint32_t SG_SELF = 0x80000000u;     // second word of a pair entry: bit 31 set = (o, o), bits 0..30 = the point

__global__ void ba_pair_count_kernel(const int32_t *__restrict__ pt_ptr, const int32_t *__restrict__ cam_idx,
                                     int npts, unsigned long long *__restrict__ count) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npts) return;
    const int o0 = pt_ptr[p], k = pt_ptr[p + 1] - o0;
    unsigned long long n = (unsigned long long)k * (k + 1) / 2;
    for (int a = 1; a < k; ++a)                 // a camera named twice by one track: both orders are listed
        for (int b = 0; b < a; ++b) n += (cam_idx[o0 + a] == cam_idx[o0 + b]) ? 1 : 0;
    count[p] = n;
}

__global__ void ba_pair_emit_kernel(const int32_t *__restrict__ pt_ptr, const int32_t *__restrict__ cam_idx,
                                    int npts, int nc, const unsigned long long *__restrict__ offset,
                                    uint32_t *__restrict__ keys, uint2 *__restrict__ pairs) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npts) return;
    const int o0 = pt_ptr[p], k = pt_ptr[p + 1] - o0;
    size_t w = (size_t)offset[p];
    for (int a = 0; a < k; ++a) {
        const int ca = cam_idx[o0 + a];
        for (int b = 0; b <= a; ++b) {
            const int cb = cam_idx[o0 + b];
            const int hi = ca >= cb ? ca : cb, lo = ca >= cb ? cb : ca;
            // row observation = the one of the hi camera
            const int oi = ca >= cb ? o0 + a : o0 + b, oj = ca >= cb ? o0 + b : o0 + a;
            keys[w] = (uint32_t)hi * (uint32_t)nc + (uint32_t)lo;
            // an observation with itself (one per observation): the second word names the POINT instead (SG_SELF set),
            // so that the gather can fetch h_p for the right-hand side without a dependent lookup
            pairs[w] = make_uint2((unsigned)oi, a == b ? (SG_SELF | (unsigned)p) : (unsigned)oj);
            ++w;
            if (a != b && ca == cb) {
                keys[w] = (uint32_t)hi * (uint32_t)nc + (uint32_t)lo;
                pairs[w] = make_uint2((unsigned)oj, (unsigned)oi);
                ++w;
            }
        }
    }
}

// Z as an array of records the gather reads whole, at random. A record is Z_o (6 x 3 = 18 doubles) padded to 20
// doubles = 160 bytes = five whole 32-byte sectors, and it is stored at the CAMERA-MAJOR position zpos[o] of its
// observation: the pair list is sorted by (hi camera, lo camera), so consecutive list entries walk through the
// segments of two cameras -- a working set of a few hundred KB that stays in L2, where the point-major order this
// replaced scattered every block's records over the whole 0.6 GB array (2.7 GB of DRAM traffic per gather at C5,
// L2 hit rate 61 %). Thread per observation, W planes read coalesced; the 32 records of a warp go through shared
// memory and leave three at a time, ten lanes per record, so every store instruction writes whole sectors
// (scattered part-sector stores make L2 fetch the line from DRAM before merging).
constexpr int ZREC = 10;                   // 16-byte words per record
__global__ void __launch_bounds__(256)
ba_schur_z_kernel(const double *__restrict__ W, int nobs, const int32_t *__restrict__ pt_idx,
                  const int32_t *__restrict__ zpos, const double *__restrict__ Ginv, double *__restrict__ Z) {
    __shared__ __align__(16) double2 sz[8][288];
    const int o = blockIdx.x * 256 + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int zp = -1;
    if (o < nobs) {
        const int p = pt_idx[o];
        zp = zpos[o];
        double g[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) g[i] = Ginv[6 * (size_t)p + i];
        double z[18];
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            const double w0 = W[(size_t)(3 * r) * nobs + o], w1 = W[(size_t)(3 * r + 1) * nobs + o],
                         w2 = W[(size_t)(3 * r + 2) * nobs + o];
            z[3 * r + 0] = w0 * g[0];
            z[3 * r + 1] = w0 * g[1] + w1 * g[2];
            z[3 * r + 2] = w0 * g[3] + w1 * g[4] + w2 * g[5];
        }
#pragma unroll
        for (int i = 0; i < 9; ++i) sz[warp][9 * lane + i] = make_double2(z[2 * i], z[2 * i + 1]);
    }
    __syncwarp();
    double2 *Z2 = reinterpret_cast<double2 *>(Z);
    const int sub = lane / ZREC, part = lane - ZREC * sub;          // lanes 0..29: record sub of this round, word part
#pragma unroll
    for (int rnd = 0; rnd < 11; ++rnd) {
        const int rec = 3 * rnd + sub;
        const int dst = __shfl_sync(0xffffffffu, zp, rec < 32 ? rec : 0);
        if (lane < 3 * ZREC && rec < 32 && dst >= 0)
            Z2[(size_t)ZREC * dst + part] = part < 9 ? sz[warp][9 * rec + part] : make_double2(0.0, 0.0);
    }
}

__global__ void ba_invert_perm_kernel(const int32_t *__restrict__ perm, int n, int32_t *__restrict__ inv) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n) inv[perm[c]] = c;
}

__device__ __forceinline__ void sg_flush(double *__restrict__ S, size_t ld, int nc, uint32_t key, double (&acc)[36],
                                         double (&accg)[6], double *__restrict__ gred) {
    const int hi = (int)(key / (uint32_t)nc), lo = (int)(key - (uint32_t)hi * (uint32_t)nc);
    double *blk = S + (size_t)(6 * hi) * ld + 6 * lo;
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int q = 0; q < 6; ++q) {
            if (hi != lo || q <= r) atomicAdd(blk + (size_t)r * ld + q, -acc[6 * r + q]);   // K11 reads the lower triangle only
            acc[6 * r + q] = 0.0;
        }
    if (hi == lo) {
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            atomicAdd(gred + 6 * hi + r, -accg[r]);
            accg[r] = 0.0;
        }
    }
}

// Plan time: the sorted list re-laid for the gather. Slice t = entries [t seg, (t+1) seg) of the
// sorted list; entry `it` of the 32 slices of warp w sits at ((w seg + it) 32 + lane), so a warp
// reads its indices as one 128-byte and one 256-byte line per iteration. Padding entries carry
// SG_DEAD as key and observation 0.
constexpr uint32_t SG_DEAD = 0xFFFFFFFFu;
__global__ void ba_pair_interleave_kernel(const uint32_t *__restrict__ keys, const uint2 *__restrict__ pairs,
                                          const int32_t *__restrict__ zpos, size_t npairs, int seg, size_t npadded,
                                          uint32_t *__restrict__ keys_i, uint2 *__restrict__ pairs_i) {
    const size_t d = (size_t)blockIdx.x * blockDim.x + threadIdx.x;      // destination index
    if (d >= npadded) return;
    const size_t lane = d & 31, wi = d >> 5, w = wi / seg, it = wi - w * seg;
    const size_t src = (w * 32 + lane) * (size_t)seg + it;
    const bool ok = src < npairs;
    keys_i[d] = ok ? keys[src] : SG_DEAD;
    // observation indices -> positions of their records in the camera-major Z array
    const uint2 pr = ok ? pairs[src] : make_uint2(0u, 0u);
    pairs_i[d] = ok ? make_uint2((unsigned)zpos[pr.x], (pr.y & SG_SELF) ? pr.y : (unsigned)zpos[pr.y]) : make_uint2(0u, 0u);
}

// Lane = one slice of `seg` consecutive entries of the sorted list. The two 144-byte records of
// every lane's current entry are fetched by the WARP with cp.async (nine 16-byte words per record,
// consecutive lanes on consecutive words, so a request covers whole records instead of 32
// scattered sectors) straight into shared memory, one entry ahead of the arithmetic (two stages;
// the indices run two entries ahead), and read back by their owner without bank conflicts
// (record stride 9 x 16 B). The diagonal entries (o, o) -- one per observation, sorted by camera
// -- also carry the reduced right-hand side g_red[c] -= Z_o h_p, which therefore needs no pass of
// its own; their index word names the point (SG_SELF), and h_p is loaded one entry ahead like the
// records (as a dependent lookup inside the iteration it put two memory latencies on every warp's
// critical path: at least one lane of a warp holds such an entry in almost every iteration).
constexpr size_t SG_SMEM = (size_t)(SG_THREADS / 32) * 2 * 2 * 32 * 9 * sizeof(double2);

// the two record positions of a pair entry / h_p of a self entry (zero otherwise)
__device__ __forceinline__ uint2 sg_records(uint2 e) { return make_uint2(e.x, (e.y & SG_SELF) ? e.x : e.y); }
__device__ __forceinline__ void sg_load_h(const double *__restrict__ H, uint2 e, double (&h)[3]) {
    h[0] = h[1] = h[2] = 0.0;
    if (e.y & SG_SELF) {
        const double *hp = H + 3 * (size_t)(e.y & ~SG_SELF);
        h[0] = __ldg(hp); h[1] = __ldg(hp + 1); h[2] = __ldg(hp + 2);
    }
}

__device__ __forceinline__ void sg_fetch(double2 *__restrict__ dst, const double2 *__restrict__ Z2, uint2 me, int lane) {
    // dst: [2][288] words of this warp and stage
#pragma unroll
    for (int rnd = 0; rnd < 9; ++rnd) {
        const int c = lane + 32 * rnd, src = c / 9, part = c - 9 * src;
        const unsigned oi = __shfl_sync(0xffffffffu, me.x, src), oj = __shfl_sync(0xffffffffu, me.y, src);
        const unsigned s0 = (unsigned)__cvta_generic_to_shared(dst + c), s1 = (unsigned)__cvta_generic_to_shared(dst + 288 + c);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s0), "l"(Z2 + ZREC * (size_t)oi + part));
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s1), "l"(Z2 + ZREC * (size_t)oj + part));
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

__global__ void __launch_bounds__(SG_THREADS, 3)
ba_schur_gather_kernel(const uint32_t *__restrict__ keys, const uint2 *__restrict__ pairs, int seg,
                       const double *__restrict__ Z, const int32_t *__restrict__ pt_c, const double *__restrict__ H,
                       int nc, double *__restrict__ S, size_t ld, double *__restrict__ gred) {
    extern __shared__ __align__(16) double2 srec[];      // [warp][stage][2][288]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t gw = (size_t)blockIdx.x * (SG_THREADS / 32) + warp;     // the grid covers the padded list exactly
    const uint32_t *kp = keys + gw * (size_t)seg * 32 + lane;
    const uint2 *pp = pairs + gw * (size_t)seg * 32 + lane;
    double2 *wbuf = srec + (size_t)warp * (2 * 2 * 288);
    double acc[36], accg[6];
#pragma unroll
    for (int i = 0; i < 36; ++i) acc[i] = 0.0;
#pragma unroll
    for (int i = 0; i < 6; ++i) accg[i] = 0.0;
    const double2 *Z2 = reinterpret_cast<const double2 *>(Z);
    uint32_t k_cur = kp[0], k_nx = seg > 1 ? kp[32] : SG_DEAD;
    uint2 me_cur = pp[0], me_nx = seg > 1 ? pp[32] : make_uint2(0u, 0u);
    uint32_t cur = k_cur;                                 // block being accumulated (SG_DEAD: nothing yet)
    sg_fetch(wbuf, Z2, sg_records(me_cur), lane);
    double hn[3];
    sg_load_h(H, me_cur, hn);
    for (int it = 0; it < seg; ++it) {
        const uint32_t k = k_cur;
        k_cur = k_nx;
        me_cur = me_nx;
        const double h[3] = {hn[0], hn[1], hn[2]};        // of this entry; the next entry's loads go out now
        if (it + 2 < seg) { k_nx = kp[(size_t)(it + 2) * 32]; me_nx = pp[(size_t)(it + 2) * 32]; }
        if (it + 1 < seg) {
            sg_load_h(H, me_cur, hn);
            sg_fetch(wbuf + ((it + 1) & 1) * (2 * 288), Z2, sg_records(me_cur), lane);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncwarp();
        const double2 *ri = wbuf + (it & 1) * (2 * 288) + 9 * lane, *rj = ri + 288;
        if (k != SG_DEAD) {
            if (k != cur) {
                sg_flush(S, ld, nc, cur, acc, accg, gred);
                cur = k;
            }
            double b[18];
#pragma unroll
            for (int i = 0; i < 9; ++i) {
                const double2 v = rj[i];
                b[2 * i] = v.x;
                b[2 * i + 1] = v.y;
            }
#pragma unroll
            for (int hh = 0; hh < 3; ++hh) {          // two rows of Z_i (three 16-byte words) at a time
                const double2 a0 = ri[3 * hh], a1 = ri[3 * hh + 1], a2 = ri[3 * hh + 2];
                const double ar[2][3] = {{a0.x, a0.y, a1.x}, {a1.y, a2.x, a2.y}};
#pragma unroll
                for (int rr = 0; rr < 2; ++rr) {
#pragma unroll
                    for (int q = 0; q < 6; ++q)
                        acc[6 * (2 * hh + rr) + q] += ar[rr][0] * b[3 * q] + ar[rr][1] * b[3 * q + 1] + ar[rr][2] * b[3 * q + 2];
                    accg[2 * hh + rr] += ar[rr][0] * h[0] + ar[rr][1] * h[1] + ar[rr][2] * h[2];   // h = 0 off the diagonal entries
                }
            }
        }
        __syncwarp();          // the stage is refilled two iterations from now: everyone is done reading it
    }
    if (cur != SG_DEAD) sg_flush(S, ld, nc, cur, acc, accg, gred);
}

// K10b: the Schur complement by block ROW. The L2's FP64 atomic adders top out near 280 G adds/s
// on B200 (tools/ubench/bulkred.cu: the same for scalar RED and for TMA bulk reductions), and a
// point-major kernel needs 36 of them per camera pair of every point. Here a warp owns a slice
// (<= 256 observations) of ONE camera's observation list (the camera-major index K9c already
// uses) and accumulates that camera's block row S[ci, cj <= ci] in its own shared-memory window
// of WB blocks, without atomics: for observation o of point p it forms Y_o = W_o V_p^-1 once and
// adds -Y_o W_o'^T for the other observations o' of p, five at a time (lane = (o', column q)).
// The window ends at the diagonal (cj in (ci - WB, ci]); cameras further down are collected in a
// bit mask on the first pass and handled by further passes over the slice, so banded systems (large
// scenes) take one pass and small dense ones at most ceil(nc / WB). A window is flushed with one
// atomic per entry of each touched block: 36 per (slice, block) instead of 36 per (point, pair).
// Also accumulates the right-hand side g_red[ci] -= W_o b_p.
constexpr int SR_WARPS = 8;
constexpr int SR_WB_MAX = 64;

// everything one observation needs from global memory, fetched one observation ahead
struct SrLoad {
    int o0, L;
    double a0, a1, a2, b0, b1, b2;     // lanes < 18: W_o[r][:], Vinv[:][k]; lanes 18..23: W_o[r][:], b_p
    int cj[2];                         // the first two batches of the track (positions g and 5 + g)
    double w[2][3];
};

__global__ void __launch_bounds__(SR_WARPS * 32, 2)
ba_schur_rows_kernel(const double *__restrict__ W, int nobs, const int32_t *__restrict__ cam_idx,
                     const int32_t *__restrict__ pt_ptr, const int32_t *__restrict__ perm,
                     const int32_t *__restrict__ pt_c, const int32_t *__restrict__ chunk_cam,
                     const int32_t *__restrict__ chunk_beg, const int32_t *__restrict__ chunk_end,
                     const double *__restrict__ Vinv, const double *__restrict__ bp, int WB, int slice,
                     double *__restrict__ S, size_t ld, double *__restrict__ gred) {
    extern __shared__ __align__(16) double sr_smem[];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int stride = WB * 36 + 24;
    double *acc = sr_smem + (size_t)warp * stride, *Ys = acc + WB * 36;
    const int ch = blockIdx.x, ci = chunk_cam[ch];
    const int beg = chunk_beg[ch] + (blockIdx.y * SR_WARPS + warp) * slice, end = min(chunk_end[ch], beg + slice);
    if (beg >= end) return;                        // (no block-wide barrier below)
    const int g = lane / 6, q = lane - 6 * g;      // batch slot (g == 5: lanes 30, 31 idle) and block column
    unsigned todo = 0;                             // windows below the first one that hold blocks of this row
    double gacc = 0.0;                             // lanes 18..23: row ci of g_red
    const int nwin = ci / WB + 1;
    for (int d = 0; d < nwin; ++d) {
        if (d > 0 && !((todo >> (d & 31)) & 1u)) continue;
        const int w_hi = ci - d * WB, w0 = w_hi - WB + 1;     // window: cameras w0 .. w_hi (w0 may be negative)
        for (int e = lane; e < WB * 36; e += 32) acc[e] = 0.0;
        unsigned long long touched = 0ull;
        __syncwarp();

        // -Y W'^T of one batch into the window (cj < 0: no observation in this lane's slot)
        auto batch = [&](const double (&Y)[18], int cj, double wq0, double wq1, double wq2) {
            const bool lower = cj >= 0 && cj <= ci;
            const bool inwin = lower && cj <= w_hi && cj >= w0;
            if (d == 0 && lower && !inwin) todo |= 1u << (((ci - cj) / WB) & 31);
            if (!__any_sync(full, inwin)) return;
            // the same camera twice in one batch (it observed the point twice) would race
            const unsigned same = __match_any_sync(full, inwin ? cj : -2 - lane);
            const bool dup = __any_sync(full, inwin && __popc(same) > 6);
            const int slot = cj - w0;
            if (!dup) {
                if (inwin) {
                    double *a = acc + slot * 36 + q;
#pragma unroll
                    for (int r = 0; r < 6; ++r)
                        a[6 * r] -= Y[3 * r] * wq0 + Y[3 * r + 1] * wq1 + Y[3 * r + 2] * wq2;
                    touched |= 1ull << slot;
                }
            } else {
                for (int gg = 0; gg < 5; ++gg) {
                    if (inwin && g == gg) {
                        double *a = acc + slot * 36 + q;
#pragma unroll
                        for (int r = 0; r < 6; ++r)
                            a[6 * r] -= Y[3 * r] * wq0 + Y[3 * r + 1] * wq1 + Y[3 * r + 2] * wq2;
                        touched |= 1ull << slot;
                    }
                    __syncwarp();
                }
            }
        };

        for (int kb = beg; kb < end; kb += 32) {
            // lane-parallel index fetch for up to 32 observations of this camera
            int o_l = 0, p_l = 0, o0_l = 0, L_l = 0;
            if (kb + lane < end) {
                o_l = perm[kb + lane];
                p_l = pt_c[kb + lane];
                o0_l = pt_ptr[p_l];
                L_l = pt_ptr[p_l + 1] - o0_l;
            }
            const int cnt = min(32, end - kb);
            auto fetch = [&](int i, SrLoad &f) {
                const int o = __shfl_sync(full, o_l, i), p = __shfl_sync(full, p_l, i);
                f.o0 = __shfl_sync(full, o0_l, i);
                f.L = __shfl_sync(full, L_l, i);
                f.a0 = f.a1 = f.a2 = f.b0 = f.b1 = f.b2 = 0.0;
                if (lane < 24) {
                    const int r = lane < 18 ? lane / 3 : lane - 18;
                    f.a0 = W[(size_t)(3 * r) * nobs + o];
                    f.a1 = W[(size_t)(3 * r + 1) * nobs + o];
                    f.a2 = W[(size_t)(3 * r + 2) * nobs + o];
                    if (lane < 18) {
                        const int k = lane - 3 * r;
                        const double *vi = Vinv + 6 * (size_t)p;
                        f.b0 = vi[sym_idx(0, k)]; f.b1 = vi[sym_idx(1, k)]; f.b2 = vi[sym_idx(2, k)];
                    } else {
                        const double *b = bp + 3 * (size_t)p;
                        f.b0 = b[0]; f.b1 = b[1]; f.b2 = b[2];
                    }
                }
#pragma unroll
                for (int bb = 0; bb < 2; ++bb) {
                    const int t = 5 * bb + g;
                    f.cj[bb] = -1;
                    f.w[bb][0] = f.w[bb][1] = f.w[bb][2] = 0.0;
                    if (g < 5 && t < f.L) {
                        const size_t oo = (size_t)f.o0 + t;
                        f.cj[bb] = cam_idx[oo];
                        f.w[bb][0] = W[(size_t)(3 * q) * nobs + oo];
                        f.w[bb][1] = W[(size_t)(3 * q + 1) * nobs + oo];
                        f.w[bb][2] = W[(size_t)(3 * q + 2) * nobs + oo];
                    }
                }
            };
            SrLoad cur, nxt;
            fetch(0, cur);
            for (int i = 0; i < cnt; ++i) {
                if (i + 1 < cnt) fetch(i + 1, nxt);
                // lanes < 18: Y_o[r][k] = sum_m W_o[r][m] Vinv[m][k]; lanes 18..23: W_o[r][:] b_p
                const double y = cur.a0 * cur.b0 + cur.a1 * cur.b1 + cur.a2 * cur.b2;
                if (lane >= 18 && d == 0) gacc -= y;
                __syncwarp();                      // the previous observation's Y has been read by every lane
                if (lane < 18) Ys[lane] = y;
                __syncwarp();
                double Y[18];
#pragma unroll
                for (int e = 0; e < 18; ++e) Y[e] = Ys[e];
                batch(Y, cur.cj[0], cur.w[0][0], cur.w[0][1], cur.w[0][2]);
                if (cur.L > 5) batch(Y, cur.cj[1], cur.w[1][0], cur.w[1][1], cur.w[1][2]);
                for (int t0 = 10; t0 < cur.L; t0 += 5) {          // long tracks: the rest in place
                    const int t = t0 + g;
                    int cj = -1;
                    double wq0 = 0.0, wq1 = 0.0, wq2 = 0.0;
                    if (g < 5 && t < cur.L) {
                        const size_t oo = (size_t)cur.o0 + t;
                        cj = cam_idx[oo];
                        wq0 = W[(size_t)(3 * q) * nobs + oo];
                        wq1 = W[(size_t)(3 * q + 1) * nobs + oo];
                        wq2 = W[(size_t)(3 * q + 2) * nobs + oo];
                    }
                    batch(Y, cj, wq0, wq1, wq2);
                }
                cur = nxt;
            }
        }
        __syncwarp();
        if (d == 0) todo = __reduce_or_sync(full, todo);
        unsigned t_lo = __reduce_or_sync(full, (unsigned)(touched & 0xffffffffull));
        unsigned t_hi = __reduce_or_sync(full, (unsigned)(touched >> 32));
        unsigned long long tm = ((unsigned long long)t_hi << 32) | t_lo;
        while (tm) {
            const int slot = __ffsll((long long)tm) - 1;
            tm &= tm - 1;
            const int cj = w0 + slot;
            for (int e = lane; e < 36; e += 32) {
                const int r = e / 6, qq = e - 6 * r;
                atomicAdd(&S[(size_t)(6 * ci + r) * ld + 6 * cj + qq], acc[slot * 36 + e]);
            }
        }
        __syncwarp();
    }
    if (lane >= 18 && lane < 24) atomicAdd(&gred[6 * ci + lane - 18], gacc);
}

// K10s: the Schur complement accumulated in SHARED memory, for problems whose camera count lets the
// lower block triangle of S be cut into a few row ranges of <= ~190 KB (up to ~100 cameras: the
// reference's regime, 6 cameras on its own data, 50 at BASELINE config 3). CTA (range t, slice g)
// streams slice g of the point-sorted observation list through a small staging buffer -- W planes
// with coalesced loads, Y = W V^-1 computed once per observation -- and adds -Y_i W_j^T for the
// block rows of its range into its shared-memory copy; warp w of 16 owns the rows with
// (ci - first row) % 16 == w, so no two warps ever touch the same block and nothing is atomic until
// the final flush (36 atomics per touched block and CTA instead of 36 per point and camera pair:
// 23x fewer at config 3). Lane = (observation j of the track, block column q), five j per pass.
constexpr int SS_THREADS = 512;
constexpr int SS_WARPS = SS_THREADS / 32;
constexpr int SS_LDW = 97;                      // row stride of the staged W planes (odd: spreads the banks)
constexpr int SS_OB = 96;                       // observations per staged batch (whole points only)
constexpr int SS_MAX_RANGES = 8;
constexpr size_t SS_STAGE_BYTES = sizeof(double) * (18 * SS_LDW + SS_OB * 18) + sizeof(int) * (4 * SS_OB + 8);

struct SchurRanges {
    int n;                                      // number of row ranges
    int row[SS_MAX_RANGES + 1];                 // range t = block rows row[t] .. row[t+1]-1
};

// plan-time statistics: longest track and whether any track names one camera twice
__global__ void ba_track_stats_kernel(const int32_t *__restrict__ pt_ptr, const int32_t *__restrict__ cam_idx,
                                      int npts, int *__restrict__ stats) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npts) return;
    const int o0 = pt_ptr[p], k = pt_ptr[p + 1] - o0;
    atomicMax(&stats[0], k);
    if (k > 64) { atomicOr(&stats[1], 1); return; }          // not examined: assume duplicates
    for (int a = 1; a < k; ++a)
        for (int b = 0; b < a; ++b)
            if (cam_idx[o0 + a] == cam_idx[o0 + b]) { atomicOr(&stats[1], 1); return; }
}

template <bool DUPS>
__global__ void __launch_bounds__(SS_THREADS, 1)
ba_schur_smem_kernel(const double *__restrict__ W, int nobs, const int32_t *__restrict__ cam_idx,
                     const int32_t *__restrict__ pt_idx, const int32_t *__restrict__ pt_ptr, int npts,
                     const double *__restrict__ Vinv, const double *__restrict__ bp, SchurRanges rg, int nslices,
                     double *__restrict__ S, size_t ld, double *__restrict__ gred, int skip) {
    extern __shared__ __align__(16) double ss_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int t = blockIdx.x % rg.n, g = blockIdx.x / rg.n;
    const int r0 = rg.row[t], r1 = rg.row[t + 1];
    const int tri0 = r0 * (r0 + 1) / 2, nblk = r1 * (r1 + 1) / 2 - tri0;
    double *acc = ss_smem, *stW = acc + (size_t)nblk * 36, *stY = stW + 18 * SS_LDW;   // 18 * 97 doubles: stY stays 16-byte aligned
    int *stC = reinterpret_cast<int *>(stY + SS_OB * 18), *stS = stC + SS_OB, *stE = stS + SS_OB, *stPt = stE + SS_OB;
    int *ctl = stPt + SS_OB;
    for (int e = tid; e < nblk * 36; e += SS_THREADS) acc[e] = 0.0;
    const int per = (npts + nslices - 1) / nslices;
    const int p0 = g * per, p1 = min(npts, p0 + per);
    const int grp = lane / 6, q = lane - 6 * grp;            // grp == 5: lanes 30, 31 idle
    for (int pb = p0; pb < p1;) {
        __syncthreads();                                      // the previous batch is consumed (and acc is zero)
        if (warp == 0) {
            const int o0 = pt_ptr[pb];
            const int pe = min(pb + lane + 1, p1);
            const bool ok = (pb + lane < p1) && (pt_ptr[pe] - o0 <= SS_OB);
            const int nb = __popc(__ballot_sync(0xffffffffu, ok));   // whole points that fit (>= 1: the
            if (lane == 0) {                                          // host checked the longest track)
                ctl[0] = max(nb, 1);
                ctl[1] = pt_ptr[pb + max(nb, 1)] - o0;
                ctl[2] = o0;
            }
        }
        __syncthreads();
        const int np_b = ctl[0], nob = min(ctl[1], SS_OB), o0 = ctl[2];
        for (int i = tid; i < nob; i += SS_THREADS) {
            const int p = pt_idx[o0 + i];
            stC[i] = cam_idx[o0 + i];
            stPt[i] = p;
            stS[i] = pt_ptr[p] - o0;
            stE[i] = pt_ptr[p + 1] - o0;
        }
        for (int e = tid; e < nob * 18; e += SS_THREADS) {
            const int rc = e / nob, i = e - rc * nob;
            stW[rc * SS_LDW + i] = W[(size_t)rc * nobs + o0 + i];
        }
        __syncthreads();
        if (!(skip & 2))
        for (int e = tid; e < nob * 6; e += SS_THREADS) {     // Y_i = W_i V^-1 (row r), g_red -= W_i b_p
            const int i = e / 6, r = e - 6 * i, p = stPt[i];
            const double w0 = stW[(3 * r) * SS_LDW + i], w1 = stW[(3 * r + 1) * SS_LDW + i], w2 = stW[(3 * r + 2) * SS_LDW + i];
            const double *vi = Vinv + 6 * (size_t)p;
            stY[i * 18 + 3 * r + 0] = w0 * vi[0] + w1 * vi[1] + w2 * vi[3];
            stY[i * 18 + 3 * r + 1] = w0 * vi[1] + w1 * vi[2] + w2 * vi[4];
            stY[i * 18 + 3 * r + 2] = w0 * vi[3] + w1 * vi[4] + w2 * vi[5];
            if (t == 0) {
                const double *b = bp + 3 * (size_t)p;
                atomicAdd(&gred[6 * stC[i] + r], -(w0 * b[0] + w1 * b[1] + w2 * b[2]));
            }
        }
        __syncthreads();
        if (!(skip & 1))
        for (int ib = 0; ib < nob; ib += 32) {
            const int i = ib + lane;
            const int c = i < nob ? stC[i] : -1;
            const bool owned = c >= r0 && c < r1 && ((c - r0) & (SS_WARPS - 1)) == warp;
            unsigned om = __ballot_sync(0xffffffffu, owned);
            while (om) {
                const int src = __ffs(om) - 1;
                om &= om - 1;
                const int ii = ib + src;
                const int ci = __shfl_sync(0xffffffffu, c, src);
                const int s_ = stS[ii], e_ = stE[ii];
                double Y[18];
                {
                    const double2 *y2 = reinterpret_cast<const double2 *>(stY + ii * 18);
#pragma unroll
                    for (int k = 0; k < 9; ++k) { const double2 v = y2[k]; Y[2 * k] = v.x; Y[2 * k + 1] = v.y; }
                }
                double *rowbase = acc + (size_t)(ci * (ci + 1) / 2 - tri0) * 36 + q;
                for (int t0 = s_; t0 < e_; t0 += 5) {
                    const int j = t0 + grp;
                    const bool have = grp < 5 && j < e_;
                    const int cj = have ? stC[j] : 0x7fffffff;
                    const bool ok = have && cj <= ci;
                    double w0 = 0.0, w1 = 0.0, w2 = 0.0;
                    if (!__any_sync(0xffffffffu, ok)) continue;
                    if (ok) { w0 = stW[(3 * q) * SS_LDW + j]; w1 = stW[(3 * q + 1) * SS_LDW + j]; w2 = stW[(3 * q + 2) * SS_LDW + j]; }
                    if (!DUPS) {
                        if (ok) {
                            double *a = rowbase + (size_t)cj * 36;
#pragma unroll
                            for (int r = 0; r < 6; ++r) a[6 * r] -= Y[3 * r] * w0 + Y[3 * r + 1] * w1 + Y[3 * r + 2] * w2;
                        }
                    } else {                                  // a camera twice in one track: one group at a time
                        for (int gg = 0; gg < 5; ++gg) {
                            if (ok && grp == gg) {
                                double *a = rowbase + (size_t)cj * 36;
#pragma unroll
                                for (int r = 0; r < 6; ++r) a[6 * r] -= Y[3 * r] * w0 + Y[3 * r + 1] * w1 + Y[3 * r + 2] * w2;
                            }
                            __syncwarp();
                        }
                    }
                }
            }
        }
        pb += np_b;
    }
    __syncthreads();
    for (int ci = r0; ci < r1; ++ci) {
        const double *base = acc + (size_t)(ci * (ci + 1) / 2 - tri0) * 36;
        for (int e = tid; e < (ci + 1) * 36; e += SS_THREADS) {
            const double v = base[e];
            if (v != 0.0) {
                const int cj = e / 36, el = e - 36 * cj, r = el / 6, qq = el - 6 * r;
                atomicAdd(&S[(size_t)(6 * ci + r) * ld + 6 * cj + qq], v);
            }
        }
    }
}

// Multi-GPU: only the lower block triangle of S (what K10 writes and K11 reads) and the right-hand-side
// row travel through the all-reduce. Row r < npad carries its first 48 (r / 48 + 1) entries, row npad
// (g_red) all npad: half the bytes of the padded square.
__host__ __device__ inline size_t s_pack_offset(int r, int npad) {
    if (r >= npad) return (size_t)BA_NB * BA_NB * (size_t)(npad / BA_NB) * (npad / BA_NB + 1) / 2;
    const size_t b = (size_t)r / BA_NB;
    return (size_t)BA_NB * ((size_t)BA_NB * b * (b + 1) / 2 + ((size_t)r - BA_NB * b) * (b + 1));
}

template <bool PACK>
__global__ void __launch_bounds__(256)
ba_s_pack_kernel(double *__restrict__ S, size_t ld, int npad, double *__restrict__ packed) {
    const int r = blockIdx.x;                       // 0 .. npad
    const int len = r < npad ? BA_NB * (r / BA_NB + 1) : npad;
    double *row = S + (size_t)r * ld, *dst = packed + s_pack_offset(r, npad);
    for (int c = threadIdx.x; c < len; c += 256) {
        if (PACK) dst[c] = row[c];
        else row[c] = dst[c];
    }
}

// Multi-GPU with K10g: the reduced camera system of a large scene is block sparse, and the pair lists say
// exactly which 6 x 6 blocks any rank touches. The union over ranks (one all-reduce of a flag per block of
// the lower block triangle, once per solve) gives every rank the same ascending block list; per trial only
// those blocks and the right-hand-side row travel (C5: 11.5 MB instead of 145 MB).
constexpr int BA_SPARSE_AR_MAX_CAMS = 4096;        // flag array nc (nc + 1) / 2 doubles

__global__ void ba_blk_mark_kernel(const uint32_t *__restrict__ keys, size_t n_entries, int nc, double *__restrict__ flag) {
    const size_t d = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (d < (size_t)nc) flag[d * (d + 1) / 2 + d] = 1.0;           // the diagonal blocks carry U on rank 0
    if (d >= n_entries) return;
    const uint32_t k = keys[d];
    if (k == SG_DEAD) return;
    const size_t hi = k / (uint32_t)nc, lo = k - (uint32_t)hi * (uint32_t)nc;
    flag[hi * (hi + 1) / 2 + lo] = 1.0;                              // same value from every writer
}

// 48 x 48 block pattern of S (lower triangle, nb x nb doubles so that it can be sum-all-reduced) from the pair list
__global__ void ba_blk48_mark_kernel(const uint32_t *__restrict__ keys, size_t n_entries, int nc, int nb,
                                     double *__restrict__ pat) {
    const size_t d = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (d < (size_t)nb) pat[d * nb + d] = 1.0;
    if (d >= n_entries) return;
    const uint32_t k = keys[d];
    if (k == SG_DEAD) return;
    const uint32_t hi = k / (uint32_t)nc, lo = k - hi * (uint32_t)nc;
    pat[(size_t)(6 * hi / BA_NB) * nb + 6 * lo / BA_NB] = 1.0;
}

__global__ void ba_blk_flag_kernel(const double *__restrict__ flag, size_t T, int *__restrict__ f) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t <= T) f[t] = (t < T && flag[t] != 0.0) ? 1 : 0;           // f[T] = 0: the scan leaves the total there
}

__global__ void ba_blk_scatter_kernel(const int *__restrict__ f, const int *__restrict__ pos, int nc,
                                      uint32_t *__restrict__ list) {
    const int hi = blockIdx.x;                                        // one CTA per block row
    for (int lo = threadIdx.x; lo <= hi; lo += blockDim.x) {
        const size_t t = (size_t)hi * (hi + 1) / 2 + lo;
        if (f[t]) list[pos[t]] = ((uint32_t)hi << 16) | (uint32_t)lo;
    }
}

template <bool PACK>
__global__ void __launch_bounds__(256)
ba_s_pack_sparse_kernel(double *__restrict__ S, size_t ld, int npad, int n, const uint32_t *__restrict__ list,
                        int nlist, double *__restrict__ buf) {
    const size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x;
    const size_t nb_el = (size_t)nlist * 36;
    double *ptr;
    if (idx < nb_el) {
        const uint32_t b = list[idx / 36];
        const int e = (int)(idx % 36), r = e / 6, q = e - 6 * r, hi = (int)(b >> 16), lo = (int)(b & 0xffffu);
        ptr = S + (size_t)(6 * hi + r) * ld + 6 * lo + q;
    } else if (idx < nb_el + (size_t)n) {
        ptr = S + (size_t)npad * ld + (idx - nb_el);                 // right-hand-side row
    } else {
        return;
    }
    if (PACK) buf[idx] = *ptr;
    else *ptr = buf[idx];
}

// ------------------------------------------------------------------------------ update (K12)
// partial sums layout (4 doubles): [s.g, |D s|^2, |D x_new|^2, |D x_cur|^2]
struct UpdateCamsArgs {
    const double *cam, *dc, *gc, *Dc;
    int n, contribute;
    double *cam_new, *part;
};

__device__ __forceinline__ void update_cams_body(const double *__restrict__ cam, const double *__restrict__ dc,
                                                 const double *__restrict__ gc, const double *__restrict__ Dc, int n,
                                                 int contribute, double *__restrict__ cam_new, double *__restrict__ part) {
    __shared__ double sred[32];
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    for (int j = threadIdx.x; j < n; j += 256) {
        double s = dc[j], x = cam[j], xn = x - s, d = Dc[j];
        cam_new[j] = xn;
        a0 += s * gc[j];
        a1 += (d * s) * (d * s);
        a2 += (d * xn) * (d * xn);
        a3 += (d * x) * (d * x);
    }
    a0 = block_sum(a0, sred); a1 = block_sum(a1, sred); a2 = block_sum(a2, sred); a3 = block_sum(a3, sred);
    if (threadIdx.x == 0) {
        // multi-GPU: every rank updates the replicated cameras, one rank reports their sums
        part[0] = contribute ? a0 : 0.0; part[1] = contribute ? a1 : 0.0;
        part[2] = contribute ? a2 : 0.0; part[3] = contribute ? a3 : 0.0;
    }
}

// tp[p] = sum over the observations of p of W_o^T dc[cam(o)]   (thread per observation). The camera update
// (x_new = x - dc and its four partial sums; one block of work, independent of this kernel's) rides along as
// the LAST block of the grid instead of a launch of its own.
__global__ void __launch_bounds__(256)
ba_point_rhs_kernel(const double *__restrict__ W, int nobs, const int32_t *__restrict__ cam_idx,
                    const int32_t *__restrict__ pt_idx, const int32_t *__restrict__ pt_ptr,
                    const double *__restrict__ dc, double *__restrict__ tp, UpdateCamsArgs uc) {
    if (blockIdx.x == gridDim.x - 1) {
        update_cams_body(uc.cam, uc.dc, uc.gc, uc.Dc, uc.n, uc.contribute, uc.cam_new, uc.part);
        return;
    }
    const int lane = threadIdx.x & 31;
    const int o = blockIdx.x * 256 + threadIdx.x;
    int p = -1;
    double t[3] = {0.0, 0.0, 0.0};
    if (o < nobs) {
        p = pt_idx[o];
        const double *s = dc + 6 * (size_t)cam_idx[o];
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            double sr = s[r];
#pragma unroll
            for (int a = 0; a < 3; ++a) t[a] = fma(W[(size_t)(r * 3 + a) * nobs + o], sr, t[a]);
        }
    }
    seg_reduce<3>(t, p, lane);
    const int prev = __shfl_up_sync(0xffffffffu, p, 1);
    const bool head = (lane == 0) || (prev != p);
    if (head && p >= 0) {
        const bool complete = (o == pt_ptr[p]) && (pt_ptr[p + 1] <= o - lane + 32);
        double *dst = tp + 3 * (size_t)p;
        if (complete) { dst[0] = t[0]; dst[1] = t[1]; dst[2] = t[2]; }
        else { atomicAdd(dst, t[0]); atomicAdd(dst + 1, t[1]); atomicAdd(dst + 2, t[2]); }
    }
}

__global__ void __launch_bounds__(256)
ba_update_points_kernel(const double *__restrict__ pts, const double *__restrict__ Vinv,
                        const double *__restrict__ bp, const double *__restrict__ tp,
                        const double *__restrict__ gp, const double *__restrict__ Dp, int npts,
                        double *__restrict__ pts_new, double *__restrict__ part) {
    __shared__ double sred[32];
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    for (int p = blockIdx.x * 256 + threadIdx.x; p < npts; p += gridDim.x * 256) {
        double vi[6], t[3], vt[3];
#pragma unroll
        for (int i = 0; i < 6; ++i) vi[i] = Vinv[6 * (size_t)p + i];
#pragma unroll
        for (int i = 0; i < 3; ++i) t[i] = tp[3 * (size_t)p + i];
        sym3_mul(vi, t, vt);
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            double s = bp[3 * (size_t)p + i] - vt[i];
            double x = pts[3 * (size_t)p + i], xn = x - s, d = Dp[3 * (size_t)p + i];
            pts_new[3 * (size_t)p + i] = xn;
            a0 += s * gp[3 * (size_t)p + i];
            a1 += (d * s) * (d * s);
            a2 += (d * xn) * (d * xn);
            a3 += (d * x) * (d * x);
        }
    }
    a0 = block_sum(a0, sred); a1 = block_sum(a1, sred); a2 = block_sum(a2, sred); a3 = block_sum(a3, sred);
    if (threadIdx.x == 0) {
        double *dst = part + 4 * (size_t)(blockIdx.x + 1);     // slot 0 belongs to the cameras
        dst[0] = a0; dst[1] = a1; dst[2] = a2; dst[3] = a3;
    }
}

// out[0] = sum costpart; out[1..4] = sum of the update partials; out[5] = point-block failure (0/1);
// out[6] = gmax; out[7] = Cholesky failure (0/1). out[0..5] are summed across ranks, out[6..7] are not.
__global__ void __launch_bounds__(256)
ba_reduce_scalars_kernel(const double *__restrict__ costpart, int ncost, const double *__restrict__ updpart,
                         int nupd, const unsigned long long *__restrict__ gmax, const int *__restrict__ fail_s,
                         const int *__restrict__ fail_c, double *__restrict__ out) {
    __shared__ double sred[32];
    double c = 0.0, a[4] = {0, 0, 0, 0};
    for (int i = threadIdx.x; i < ncost; i += 256) c += costpart[i];
    for (int i = threadIdx.x; i < nupd; i += 256)
#pragma unroll
        for (int q = 0; q < 4; ++q) a[q] += updpart[4 * (size_t)i + q];
    c = block_sum(c, sred);
#pragma unroll
    for (int q = 0; q < 4; ++q) a[q] = block_sum(a[q], sred);
    if (threadIdx.x == 0) {
        out[0] = c;
#pragma unroll
        for (int q = 0; q < 4; ++q) out[1 + q] = a[q];
        out[5] = (fail_s && *fail_s) ? 1.0 : 0.0;
        out[6] = gmax ? __longlong_as_double((long long)*gmax) : 0.0;
        out[7] = (fail_c && *fail_c) ? 1.0 : 0.0;
    }
}

// ------------------------------------------------------------------------------ host driver
// development switches: SFM_SCHUR_ROWS=1 selects the block-row K10a/K10b pair instead of the
// point-major K10 (measured on B200: rows 0.27 ms vs points 0.17 ms at C3, 3.35 vs 2.9 ms at C5 --
// the row kernel issues ~280 instructions per observation and is latency bound at 16 warps per SM);
// SFM_SCHUR_WB / SFM_SCHUR_SLICE override its window and slice.
static bool schur_by_rows() {
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("SFM_SCHUR_ROWS");
        v = (e && e[0] == '1') ? 1 : 0;
    }
    return v == 1;
}

static int schur_window(int nc) {
    static int forced = -1, attr_done = 0;
    if (forced < 0) {
        const char *e = getenv("SFM_SCHUR_WB");
        forced = e ? std::max(1, std::min(SR_WB_MAX, atoi(e))) : 0;
    }
    if (!attr_done) {
        cudaFuncSetAttribute(ba_schur_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)(sizeof(double) * SR_WARPS * (SR_WB_MAX * 36 + 24)));
        attr_done = 1;
    }
    if (forced) return forced;
    return nc <= 48 ? nc : 40;       // 40 blocks per warp: two 8-warp CTAs per SM
}

// K10s applies when the lower block triangle of S splits into at most SS_MAX_RANGES row ranges that
// fit shared memory next to the staging buffer, and no track is longer than a staged batch.
static bool schur_smem_ranges(int nc, int max_track, SchurRanges &rg) {
    // Measured on B200 at config 3 (tools/ba_probe.py): K10a + K10s 0.22 ms per trial against 0.17 ms for
    // the point-major K10 -- the pair phase is latency bound (each warp walks its block rows serially,
    // ~3000 cycles per owned observation), so the L2 atomic path stays the default and this kernel is
    // opt-in (SFM_SCHUR_SMEM=1) until its pair phase is pipelined.
    static int off = -1;
    if (off < 0) {
        const char *e = getenv("SFM_SCHUR_SMEM");
        off = (e && e[0] == '1') ? 0 : 1;
    }
    if (off || max_track > SS_OB || max_track < 1) return false;
    const long long cap_blocks = (long long)((BA_SMEM_MAX - SS_STAGE_BYTES) / (36 * sizeof(double)));
    rg.n = 0;
    rg.row[0] = 0;
    int c = 0;
    while (c < nc) {
        if (rg.n == SS_MAX_RANGES) return false;
        long long blocks = 0;
        int e = c;
        while (e < nc && blocks + (e + 1) <= cap_blocks) { blocks += e + 1; ++e; }
        if (e == c) return false;                               // one block row alone does not fit
        rg.row[++rg.n] = e;
        c = e;
    }
    // balance: equal block counts per range instead of greedy-full ranges
    if (rg.n > 1) {
        const long long total = (long long)nc * (nc + 1) / 2;
        int cc = 0;
        for (int t = 1; t < rg.n; ++t) {
            const long long target = total * t / rg.n;
            while ((long long)cc * (cc + 1) / 2 < target) ++cc;
            rg.row[t] = cc;
        }
        for (int t = 0; t < rg.n; ++t) {
            const long long b = (long long)rg.row[t + 1] * (rg.row[t + 1] + 1) / 2 - (long long)rg.row[t] * (rg.row[t] + 1) / 2;
            if (b > cap_blocks || rg.row[t + 1] <= rg.row[t]) return false;
        }
    }
    return true;
}

// observations per warp of K10b: enough warps to fill the GPU several times over, few enough that the
// window flushes (36 atomics per touched block and warp) stay small next to the accumulation
static int schur_slice(int nobs) {
    static int forced = -1;
    if (forced < 0) {
        const char *e = getenv("SFM_SCHUR_SLICE");
        forced = e ? std::max(1, std::min(BA_CHUNK / SR_WARPS, atoi(e))) : 0;
    }
    if (forced) return forced;
    const int target_warps = sm_count() * 2 * SR_WARPS * 4;
    int sl = 32;
    while (sl < BA_CHUNK / SR_WARPS && nobs / sl > target_warps) sl *= 2;
    return sl;
}

struct BAPlan {
    int nc, np, nobs, npad, nchunks, grid_obs, grid_pts, obs_threads, max_track, has_dup;
    bool smem_cams;
    size_t ld, cam_smem;
    // index structures
    int32_t *pt_ptr, *cam_ptr, *perm, *keys_sorted, *iota, *pt_c, *chunk_cam, *chunk_beg, *chunk_end, *chunk_ptr;
    double2 *obs_c;
    void *cub_tmp;
    size_t cub_bytes;
    // state
    double *cam[2], *pts[2], *camlin[2];
    double *W, *V, *gp, *Vinv, *bp, *tp, *U, *gc, *Upart, *Dc, *Dp, *S, *Spack, *Linv, *dc;
    double *Z, *Ginv;                          // K10g: Z [nobs,18] records, G [np,6]
    const uint32_t *pair_keys;                 // K10g pair list (device memory owned by the plan cache, not the workspace)
    const uint2 *pair_vals;
    size_t npairs;                             // padded length: (warps x seg x 32) entries, warp-interleaved
    int pair_seg;
    double *costpart, *updpart, *scal, *red1;
    unsigned long long *gmax;
    int *flags;       // [0] index validation, [1] schur (vinv) failure, [2] cholesky failure
};

static size_t carve(unsigned char *&p, size_t bytes) {
    size_t off = (size_t)p;
    p += align_up(bytes, 256);
    return off;
}

static size_t cub_sort_bytes(int nobs, int nc) {
    size_t bytes = 0;
    int bits = 1;
    while ((1 << bits) < nc && bits < 31) ++bits;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const int32_t *)nullptr, (int32_t *)nullptr,
                                    (const int32_t *)nullptr, (int32_t *)nullptr, nobs, 0, bits);
    return bytes;
}

// Lays the plan out over a workspace starting at `base` (base == nullptr: size query only).
static size_t ba_layout(BAPlan &pl, unsigned char *base, int nc, int np, int nobs) {
    unsigned char *p = base;
    pl.nc = nc; pl.np = np; pl.nobs = nobs;
    pl.npad = ceil_div(6 * nc, BA_NB) * BA_NB;
    pl.ld = (size_t)pl.npad;
    const int sms = sm_count();
    pl.cam_smem = (size_t)nc * CAMLIN_STRIDE * sizeof(double);
    pl.smem_cams = pl.cam_smem <= BA_SMEM_CAM_LIMIT;
    // a camera table above ~100 KB leaves room for one CTA per SM: make it a wide one (24 warps)
    // a camera table above ~100 KB leaves room for one CTA per SM: make it as wide as the shared
    // memory left for K9p's per-warp scratch allows
    pl.obs_threads = 256;
    if (pl.smem_cams && pl.cam_smem > 90 * 1024) {
        const size_t room = BA_SMEM_MAX - pl.cam_smem, per_warp = sizeof(double) * K9P_WARP_SCRATCH;
        pl.obs_threads = room >= 24 * per_warp ? 768 : room >= 20 * per_warp ? 640 : 512;
        const char *e = getenv("SFM_K9P_THREADS");            // development switch: 512 / 640 / 768
        if (e && (atoi(e) == 512 || atoi(e) == 640 || atoi(e) == 768) && atoi(e) <= pl.obs_threads) pl.obs_threads = atoi(e);
    }
    int per_sm = pl.obs_threads >= 512 ? 1 : 2;
    pl.grid_obs = std::max(1, std::min(sms * per_sm, ceil_div(nobs, pl.obs_threads)));
    pl.grid_pts = std::max(1, std::min(sms * 4, ceil_div(np, 256)));
    const int max_chunks = nobs / BA_CHUNK + nc + 1;
    const size_t D = sizeof(double), I = sizeof(int32_t);
#define CARVE(field, type, count) pl.field = reinterpret_cast<type *>(carve(p, (size_t)(count)))
    CARVE(pt_ptr, int32_t, I * (np + 1));
    CARVE(cam_ptr, int32_t, I * (nc + 1));
    CARVE(perm, int32_t, I * nobs);
    CARVE(keys_sorted, int32_t, I * nobs);
    CARVE(iota, int32_t, I * nobs);
    CARVE(pt_c, int32_t, I * nobs);
    CARVE(chunk_cam, int32_t, I * max_chunks);
    CARVE(chunk_beg, int32_t, I * max_chunks);
    CARVE(chunk_end, int32_t, I * max_chunks);
    CARVE(chunk_ptr, int32_t, I * (nc + 1));
    CARVE(obs_c, double2, sizeof(double2) * nobs);
    pl.cub_bytes = cub_sort_bytes(nobs, nc);
    CARVE(cub_tmp, unsigned char, pl.cub_bytes + 256);
    for (int i = 0; i < 2; ++i) {
        CARVE(cam[i], double, D * 6 * nc);
        CARVE(pts[i], double, D * 3 * np);
        CARVE(camlin[i], double, D * CAMLIN_STRIDE * nc);
    }
    CARVE(W, double, D * 18 * nobs);
    CARVE(V, double, D * 9 * np);        // V [np,6] immediately followed by gp [np,3]: one memset
    pl.gp = pl.V + 6 * (size_t)np;
    CARVE(Vinv, double, D * 6 * np);
    CARVE(bp, double, D * 3 * np);
    CARVE(tp, double, D * 3 * np);
    CARVE(Z, double, D * 20 * nobs);     // 160-byte records in camera-major order
    CARVE(Ginv, double, D * 9 * np);       // G [np,6] followed by h = G g_p [np,3]
    pl.pair_keys = nullptr; pl.pair_vals = nullptr; pl.npairs = 0; pl.pair_seg = 0;
    CARVE(U, double, D * (27 * nc + 8));   // [U 21nc | g_c 6nc | cost, ...]: ONE allreduce per linearisation
    pl.gc = pl.U + 21 * (size_t)nc;
    pl.red1 = pl.gc + 6 * (size_t)nc;
    CARVE(Upart, double, D * 27 * max_chunks);
    CARVE(Dc, double, D * 6 * nc);
    CARVE(Dp, double, D * 3 * np);
    CARVE(S, double, D * (pl.npad + BA_NB) * pl.ld);
    CARVE(Spack, double, D * (s_pack_offset(pl.npad, pl.npad) + pl.npad));     // multi-GPU all-reduce staging
    CARVE(Linv, double, D * (pl.npad / BA_NB) * BA_NB * BA_NB + (size_t)(pl.npad / BA_NB + 1) * (pl.npad / BA_NB) + 64);  // + K11's block flags
    CARVE(dc, double, D * pl.npad);
    CARVE(costpart, double, D * (pl.grid_obs + 8));
    CARVE(updpart, double, D * 4 * (pl.grid_pts + 2));
    CARVE(scal, double, D * 16);
    CARVE(gmax, unsigned long long, 8);
    CARVE(flags, int, 64);
#undef CARVE
    return (size_t)(p - base);
}

template <bool SM>
static void set_obs_kernel_attrs(size_t smem) {
    cudaFuncSetAttribute(ba_point_pass_kernel<SM, 256, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(ba_point_pass_kernel<SM, 768, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(ba_point_pass_kernel<SM, 640, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(ba_point_pass_kernel<SM, 512, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(ba_cost_kernel<SM, 256, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(ba_cost_kernel<SM, 768, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}

static int launch_point_pass(const BAPlan &pl, const KParam &K, const double *camlin, const double *pts,
                             const double2 *obs, const int32_t *cam_idx, const int32_t *pt_idx, cudaStream_t st) {
    SFM_CUDA_TRY(cudaMemsetAsync(pl.V, 0, sizeof(double) * 9 * (size_t)pl.np, st));
    const size_t sm = (pl.smem_cams ? pl.cam_smem : 0) + sizeof(double) * K9P_WARP_SCRATCH * (pl.obs_threads / 32);
#define PP_ARGS K, camlin, pl.nc, pts, obs, cam_idx, pt_idx, pl.pt_ptr, pl.nobs, pl.W, pl.V, pl.gp, pl.costpart
    if (pl.smem_cams) {
        if (pl.obs_threads == 768) ba_point_pass_kernel<true, 768, 1><<<pl.grid_obs, 768, sm, st>>>(PP_ARGS);
        else if (pl.obs_threads == 640) ba_point_pass_kernel<true, 640, 1><<<pl.grid_obs, 640, sm, st>>>(PP_ARGS);
        else if (pl.obs_threads == 512) ba_point_pass_kernel<true, 512, 1><<<pl.grid_obs, 512, sm, st>>>(PP_ARGS);
        else ba_point_pass_kernel<true, 256, 2><<<pl.grid_obs, 256, sm, st>>>(PP_ARGS);
    } else {
        ba_point_pass_kernel<false, 256, 2><<<pl.grid_obs, 256, sm, st>>>(PP_ARGS);
    }
#undef PP_ARGS
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

static int launch_cost(const BAPlan &pl, const double *camlin, const double *pts, const double2 *obs,
                       const int32_t *cam_idx, const int32_t *pt_idx, double2 *r_out, cudaStream_t st) {
    const size_t sm = pl.smem_cams ? (size_t)pl.nc * K8_CAM_STRIDE * sizeof(double) : 0;
#define C_ARGS camlin, pl.nc, pts, obs, cam_idx, pt_idx, pl.nobs, pl.costpart, r_out
    if (pl.smem_cams) {
        if (pl.obs_threads >= 512) ba_cost_kernel<true, 768, 1><<<pl.grid_obs, 768, sm, st>>>(C_ARGS);
        else ba_cost_kernel<true, 256, 2><<<pl.grid_obs, 256, sm, st>>>(C_ARGS);
    } else {
        ba_cost_kernel<false, 256, 2><<<pl.grid_obs, 256, 0, st>>>(C_ARGS);
    }
#undef C_ARGS
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

static int launch_cam_linearise(const BAPlan &pl, const KParam &K, const double *cam, double *camlin,
                                cudaStream_t st) {
    ba_cam_linearise_kernel<<<ceil_div(pl.nc, 128), 128, 0, st>>>(K, cam, pl.nc, camlin);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

// Builds pt_ptr, the camera-major copy of the observations and the chunk table. Synchronises
// the stream once (the chunk table is laid out on the host from the per-camera counts).
static int ba_build_index(BAPlan &pl, const double2 *obs, const int32_t *cam_idx, const int32_t *pt_idx,
                          cudaStream_t st) {
    const int nobs = pl.nobs, nc = pl.nc, np = pl.np;
    SFM_CUDA_TRY(cudaMemsetAsync(pl.flags, 0, 64 * sizeof(int), st));
    ba_csr_from_sorted_kernel<<<ceil_div(nobs + 1, 256), 256, 0, st>>>(pt_idx, nobs, np, pl.pt_ptr, pl.flags);
    SFM_LAUNCH_CHECK();
    ba_range_check_kernel<<<ceil_div(nobs, 256), 256, 0, st>>>(cam_idx, nobs, nc, pl.flags);
    SFM_LAUNCH_CHECK();
    ba_iota_kernel<<<ceil_div(nobs, 256), 256, 0, st>>>(pl.iota, nobs);
    SFM_LAUNCH_CHECK();
    int bits = 1;
    while ((1 << bits) < nc && bits < 31) ++bits;
    size_t tmp = pl.cub_bytes;
    SFM_CUDA_TRY(cub::DeviceRadixSort::SortPairs(pl.cub_tmp, tmp, cam_idx, pl.keys_sorted,
                                                 (const int32_t *)pl.iota, pl.perm, nobs, 0, bits, st));
    count_launch();
    ba_csr_from_sorted_kernel<<<ceil_div(nobs + 1, 256), 256, 0, st>>>(pl.keys_sorted, nobs, nc, pl.cam_ptr,
                                                                        pl.flags + 3);
    SFM_LAUNCH_CHECK();
    ba_gather_cam_major_kernel<<<ceil_div(nobs, 256), 256, 0, st>>>(pl.perm, obs, pt_idx, nobs, pl.obs_c, pl.pt_c);
    SFM_LAUNCH_CHECK();
    ba_track_stats_kernel<<<ceil_div(np, 256), 256, 0, st>>>(pl.pt_ptr, cam_idx, np, pl.flags + 8);
    SFM_LAUNCH_CHECK();
    // chunk table on the host
    std::vector<int32_t> cam_ptr(nc + 1);
    int flag0 = 0, stats[2] = {0, 0};
    SFM_CUDA_TRY(cudaMemcpyAsync(stats, pl.flags + 8, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    SFM_CUDA_TRY(cudaMemcpyAsync(cam_ptr.data(), pl.cam_ptr, sizeof(int32_t) * (nc + 1), cudaMemcpyDeviceToHost, st));
    SFM_CUDA_TRY(cudaMemcpyAsync(&flag0, pl.flags, sizeof(int), cudaMemcpyDeviceToHost, st));
    SFM_CUDA_TRY(cudaStreamSynchronize(st));
    if (flag0) return SFM_E_INVALID;        // indices out of range or not sorted by point
    pl.max_track = stats[0];
    pl.has_dup = stats[1];
    std::vector<int32_t> ccam, cbeg, cend, cptr(nc + 1);
    for (int c = 0; c < nc; ++c) {
        cptr[c] = (int32_t)ccam.size();
        for (int b = cam_ptr[c]; b < cam_ptr[c + 1]; b += BA_CHUNK) {
            ccam.push_back(c);
            cbeg.push_back(b);
            cend.push_back(std::min(b + BA_CHUNK, cam_ptr[c + 1]));
        }
    }
    cptr[nc] = (int32_t)ccam.size();
    pl.nchunks = (int)ccam.size();
    if (pl.nchunks) {
        SFM_CUDA_TRY(cudaMemcpyAsync(pl.chunk_cam, ccam.data(), sizeof(int32_t) * pl.nchunks, cudaMemcpyHostToDevice, st));
        SFM_CUDA_TRY(cudaMemcpyAsync(pl.chunk_beg, cbeg.data(), sizeof(int32_t) * pl.nchunks, cudaMemcpyHostToDevice, st));
        SFM_CUDA_TRY(cudaMemcpyAsync(pl.chunk_end, cend.data(), sizeof(int32_t) * pl.nchunks, cudaMemcpyHostToDevice, st));
    }
    SFM_CUDA_TRY(cudaMemcpyAsync(pl.chunk_ptr, cptr.data(), sizeof(int32_t) * (nc + 1), cudaMemcpyHostToDevice, st));
    SFM_CUDA_TRY(cudaStreamSynchronize(st));   // the host vectors go out of scope
    return SFM_OK;
}

// K10g applies to problems large enough for the L2 atomic ceiling to matter; small ones (the wrapper's
// bundle adjustments) keep the point-major K10 and skip the plan-time sort. Development switches:
// SFM_SCHUR_PAIRS=0 turns it off, SFM_SCHUR_PAIRS_MIN=<observations> moves the threshold (default 16384).
static bool schur_pairs_wanted(int nc, int nobs) {
    const char *e = getenv("SFM_SCHUR_PAIRS");
    if (e && e[0] == '0') return false;
    if (schur_by_rows() || (getenv("SFM_SCHUR_SMEM") && getenv("SFM_SCHUR_SMEM")[0] == '1')) return false;
    const char *m = getenv("SFM_SCHUR_PAIRS_MIN");
    const int min_obs = m ? atoi(m) : 16384;
    return nc <= 65535 && nobs >= min_obs;
}

// Plan time: the sorted (block key, observation pair) list of K10g. *keys == nullptr on return
// means "not applicable" (switched off, too small, or more than 2^31 pairs).
// The stream-ordered allocations below (pair list, its sort buffers, the multi-GPU block list) come from the
// device's default memory pool, whose release threshold is 0 by default: every stream synchronisation hands
// unused pool memory back to the driver and the next solve pays cudaMalloc prices again (measured: 2.3 -> 7.2 ms
// per end-to-end solve at C3). Keep it cached instead.
static void keep_async_pool_cached() {
    static bool done = false;
    if (done) return;
    done = true;
    int dev = 0;
    cudaMemPool_t pool;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        unsigned long long thr = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    (void)cudaGetLastError();
}

static int ba_build_pairs(const BAPlan &pl, const int32_t *cam_idx, cudaStream_t st, uint32_t **keys, uint2 **vals,
                          size_t *npairs, int *seg_out) {
    *keys = nullptr; *vals = nullptr; *npairs = 0; *seg_out = 0;
    if (!schur_pairs_wanted(pl.nc, pl.nobs)) return SFM_OK;
    keep_async_pool_cached();
    const int np = pl.np;
    unsigned long long *cnt = nullptr, *off = nullptr;
    SFM_CUDA_TRY(cudaMallocAsync(&cnt, sizeof(unsigned long long) * ((size_t)np + 1), st));
    SFM_CUDA_TRY(cudaMallocAsync(&off, sizeof(unsigned long long) * ((size_t)np + 1), st));
    SFM_CUDA_TRY(cudaMemsetAsync(cnt + np, 0, sizeof(unsigned long long), st));
    ba_pair_count_kernel<<<ceil_div(np, 256), 256, 0, st>>>(pl.pt_ptr, cam_idx, np, cnt);
    SFM_LAUNCH_CHECK();
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, cnt, off, np + 1, st);
    void *tmp = nullptr;
    SFM_CUDA_TRY(cudaMallocAsync(&tmp, tmp_bytes + 256, st));
    SFM_CUDA_TRY(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, cnt, off, np + 1, st));
    count_launch();
    unsigned long long total = 0;
    SFM_CUDA_TRY(cudaMemcpyAsync(&total, off + np, sizeof(total), cudaMemcpyDeviceToHost, st));
    SFM_CUDA_TRY(cudaStreamSynchronize(st));
    SFM_CUDA_TRY(cudaFreeAsync(tmp, st));
    SFM_CUDA_TRY(cudaFreeAsync(cnt, st));
    if (total == 0 || total > 0x7fffffffull) {
        SFM_CUDA_TRY(cudaFreeAsync(off, st));
        return SFM_OK;
    }
    const size_t P = (size_t)total;
    uint32_t *k_in = nullptr, *k_out = nullptr;
    unsigned long long *v_in = nullptr, *v_out = nullptr;
    SFM_CUDA_TRY(cudaMallocAsync(&k_in, sizeof(uint32_t) * P, st));
    SFM_CUDA_TRY(cudaMallocAsync(&v_in, sizeof(unsigned long long) * P, st));
    SFM_CUDA_TRY(cudaMallocAsync(&k_out, sizeof(uint32_t) * P, st));
    SFM_CUDA_TRY(cudaMallocAsync(&v_out, sizeof(unsigned long long) * P, st));
    ba_pair_emit_kernel<<<ceil_div(np, 256), 256, 0, st>>>(pl.pt_ptr, cam_idx, np, pl.nc, off, k_in,
                                                         reinterpret_cast<uint2 *>(v_in));
    SFM_LAUNCH_CHECK();
    int bits = 1;
    while (bits < 32 && (1ull << bits) < (unsigned long long)pl.nc * (unsigned long long)pl.nc) ++bits;
    size_t sort_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, k_in, k_out, v_in, v_out, (int)P, 0, bits, st);
    SFM_CUDA_TRY(cudaMallocAsync(&tmp, sort_bytes + 256, st));
    SFM_CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, sort_bytes, k_in, k_out, v_in, v_out, (int)P, 0, bits, st));
    count_launch();
    SFM_CUDA_TRY(cudaFreeAsync(tmp, st));
    SFM_CUDA_TRY(cudaFreeAsync(k_in, st));
    SFM_CUDA_TRY(cudaFreeAsync(v_in, st));
    SFM_CUDA_TRY(cudaFreeAsync(off, st));
    // slice length: long enough that the 36 atomics of a flush (which contend when few blocks take
    // many entries, as at C3) are small next to the accumulation, short enough to leave a few hundred
    // slices per SM. Measured on B200: C3 2384 / 2509 / 2588 LM iterations/s at 8 / 16 / 32, C5 flat
    // (199-200) from 24 to 48, falling beyond 64 (development switch: SFM_SCHUR_SEG)
    static int seg_forced = getenv("SFM_SCHUR_SEG") ? atoi(getenv("SFM_SCHUR_SEG")) : 0;
    const size_t target = (size_t)sm_count() * 256;
    int seg = (int)std::min<size_t>(48, std::max<size_t>(16, (P + target - 1) / target));
    if (seg_forced > 0) seg = seg_forced;
    const size_t nslices = (P + seg - 1) / seg;
    const size_t nwarps = align_up((nslices + 31) / 32, SG_THREADS / 32);     // whole CTAs
    const size_t npadded = nwarps * 32 * (size_t)seg;
    uint32_t *k_i = nullptr;
    unsigned long long *v_i = nullptr;
    SFM_CUDA_TRY(cudaMallocAsync(&k_i, sizeof(uint32_t) * npadded, st));
    SFM_CUDA_TRY(cudaMallocAsync(&v_i, sizeof(unsigned long long) * npadded, st));
    // zpos = inverse of the camera-major permutation, kept in the index's `iota` scratch (the sort is done with it)
    ba_invert_perm_kernel<<<ceil_div(pl.nobs, 256), 256, 0, st>>>(pl.perm, pl.nobs, pl.iota);
    SFM_LAUNCH_CHECK();
    ba_pair_interleave_kernel<<<(unsigned)((npadded + 255) / 256), 256, 0, st>>>(
        k_out, reinterpret_cast<const uint2 *>(v_out), pl.iota, P, seg, npadded, k_i, reinterpret_cast<uint2 *>(v_i));
    SFM_LAUNCH_CHECK();
    SFM_CUDA_TRY(cudaFreeAsync(k_out, st));
    SFM_CUDA_TRY(cudaFreeAsync(v_out, st));
    *keys = k_i;
    *vals = reinterpret_cast<uint2 *>(v_i);
    *npairs = npadded;
    *seg_out = seg;
    return SFM_OK;
}

}  // namespace sfm

using namespace sfm;

extern "C" void sfm_ba_default_options(sfm_ba_options *o) {
    if (!o) return;
    o->max_trials = 100;     // BundleAdjustment.py:211 max_nfev=100
    o->verbose = 0;
    o->ftol = o->xtol = o->gtol = 1e-8;   // scipy least_squares defaults
    o->lambda0 = 1e-4;
    o->reuse_index = 0;
}

extern "C" size_t sfm_ba_workspace_bytes(int32_t n_cams, int32_t n_pts, int32_t n_obs) {
    if (n_cams < 1 || n_pts < 1 || n_obs < 1) return 0;
    BAPlan pl;
    return ba_layout(pl, nullptr, n_cams, n_pts, n_obs) + 256;
}

static bool g_attrs_set = false;
static double g_ba_prof[8] = {0, 0, 0, 0, 0, 0, 0, 0};

struct PlanKey {
    const void *ws, *cam_idx;
    int nc, np, nobs, nchunks, max_track, has_dup;
    uint32_t *pair_keys;      // K10g pair list of this problem (cudaMalloc'ed; freed when the entry is replaced or evicted)
    uint2 *pair_vals;
    size_t npairs;
    int pair_seg;
    CholDagPlan *dag;         // K11s plan (null: not planned yet, or the dense kernel serves this problem)
    int dag_tried;
};
static std::mutex g_plan_mu;
static std::vector<PlanKey> g_plans;


// Index structures: built once per (workspace, problem shape) when the caller vouches for
// unchanged indices (reuse), else on every call.
static int ba_acquire_index(BAPlan &pl, bool reuse, const void *d_ws, const double2 *obs, const int32_t *d_cam_idx,
                            const int32_t *d_pt_idx, int n_cams, int n_pts, int n_obs, cudaStream_t st,
                            bool refresh_obs = true) {
    bool hit = false;
    if (reuse) {
        std::lock_guard<std::mutex> g(g_plan_mu);
        for (const PlanKey &pk : g_plans)
            if (pk.ws == d_ws && pk.nc == n_cams && pk.np == n_pts && pk.nobs == n_obs && pk.cam_idx == d_cam_idx) {
                pl.nchunks = pk.nchunks;
                pl.max_track = pk.max_track;
                pl.has_dup = pk.has_dup;
                pl.pair_keys = pk.pair_keys; pl.pair_vals = pk.pair_vals; pl.npairs = pk.npairs; pl.pair_seg = pk.pair_seg;
                hit = true;
                break;
            }
    }
    if (hit && !refresh_obs) return SFM_OK;
    if (hit) {
        // the INDEX structure is vouched for, the measurements are not: the camera-major copy of the observations
        // (what the camera pass K9c reads) is refreshed from the caller's array on every reuse
        ba_gather_cam_major_kernel<<<ceil_div(n_obs, 256), 256, 0, st>>>(pl.perm, obs, d_pt_idx, n_obs, pl.obs_c, pl.pt_c);
        SFM_LAUNCH_CHECK();
        return SFM_OK;
    }
    int rc = ba_build_index(pl, obs, d_cam_idx, d_pt_idx, st);
    if (rc != SFM_OK) return rc;
    uint32_t *pkeys = nullptr;
    uint2 *pvals = nullptr;
    size_t npairs = 0;
    int pseg = 0;
    if ((rc = ba_build_pairs(pl, d_cam_idx, st, &pkeys, &pvals, &npairs, &pseg)) != SFM_OK) return rc;
    pl.pair_keys = pkeys; pl.pair_vals = pvals; pl.npairs = npairs; pl.pair_seg = pseg;
    std::lock_guard<std::mutex> g(g_plan_mu);
    PlanKey pk{d_ws, d_cam_idx, n_cams, n_pts, n_obs, pl.nchunks, pl.max_track, pl.has_dup, pkeys, pvals, npairs, pseg,
               nullptr, 0};
    auto release = [st](PlanKey &q) {           // stream-ordered: kernels already queued on st still see the arrays
        if (q.pair_keys) cudaFreeAsync(q.pair_keys, st);
        if (q.pair_vals) cudaFreeAsync(q.pair_vals, st);
        if (q.dag) chol_dag_destroy(q.dag, st);
        q.pair_keys = nullptr; q.pair_vals = nullptr; q.npairs = 0; q.dag = nullptr; q.dag_tried = 0;
    };
    bool found = false;
    for (PlanKey &q : g_plans)
        if (q.ws == d_ws) { release(q); q = pk; found = true; }
    if (!found) {
        if (g_plans.size() >= 16) { release(g_plans.front()); g_plans.erase(g_plans.begin()); }
        g_plans.push_back(pk);
    }
    return SFM_OK;
}

extern "C" int sfm_ba_release_cache(uint64_t *bytes, void *stream) {
    cudaStream_t st = as_stream(stream);
    std::lock_guard<std::mutex> g(g_plan_mu);
    uint64_t total = 0;
    for (PlanKey &q : g_plans) {
        if (q.pair_keys) { SFM_CUDA_TRY(cudaFreeAsync(q.pair_keys, st)); total += sizeof(uint32_t) * (uint64_t)q.npairs; }
        if (q.pair_vals) { SFM_CUDA_TRY(cudaFreeAsync(q.pair_vals, st)); total += sizeof(uint2) * (uint64_t)q.npairs; }
        if (q.dag) chol_dag_destroy(q.dag, st);
        q.pair_keys = nullptr; q.pair_vals = nullptr; q.npairs = 0; q.dag = nullptr;
    }
    g_plans.clear();                    // index structures in the workspaces are rebuilt on the next solve
    if (bytes) *bytes = total;
    return SFM_OK;
}

namespace sfm {
int comm_allreduce_sum_f64(void *comm, double *buf, size_t count, cudaStream_t st);
int comm_allreduce_max_u64(void *comm, unsigned long long *buf, size_t count, cudaStream_t st);
}

// K11s plan of this problem (once per index structure; cached with the pair list). The 48 x 48 block pattern of
// S comes from the K10g pair list (dense when the problem is too small for one); with several GPUs every rank
// needs the pattern of the SUMMED system, so the marks are all-reduced (a collective: every rank calls this).
// *out stays null when the dense kernel is to be used (SFM_CHOL_DAG=0, or the task graph would be too large).
static int ba_acquire_dag(const BAPlan &pl, const void *d_ws, void *comm, cudaStream_t st, CholDagPlan **out) {
    *out = nullptr;
    static int mode = -1, min_nb = 0;                 // development switches
    if (mode < 0) {
        const char *e = getenv("SFM_CHOL_DAG");       // 0 = dense kernel, 1 = auto order (default), 2 = natural, 3 = nested dissection
        mode = e ? atoi(e) : 1;
        const char *m = getenv("SFM_CHOL_DAG_MIN_NB");
        min_nb = m ? atoi(m) : 0;
    }
    const int nb = pl.npad / BA_NB;
    if (mode <= 0 || nb < min_nb) return SFM_OK;
    {
        std::lock_guard<std::mutex> g(g_plan_mu);
        for (const PlanKey &pk : g_plans)
            if (pk.ws == d_ws && pk.nc == pl.nc && pk.np == pl.np && pk.nobs == pl.nobs && pk.dag_tried) {
                *out = pk.dag;
                return SFM_OK;
            }
    }
    keep_async_pool_cached();
    double *pat = nullptr;
    const size_t cells = (size_t)nb * nb;
    SFM_CUDA_TRY(cudaMallocAsync(&pat, sizeof(double) * cells, st));
    std::vector<double> hp(cells, 1.0);
    if (pl.npairs) {
        SFM_CUDA_TRY(cudaMemsetAsync(pat, 0, sizeof(double) * cells, st));
        const size_t ne = std::max(pl.npairs, (size_t)nb);
        ba_blk48_mark_kernel<<<(unsigned)((ne + 255) / 256), 256, 0, st>>>(pl.pair_keys, pl.npairs, pl.nc, nb, pat);
        SFM_LAUNCH_CHECK();
    } else {
        SFM_CUDA_TRY(cudaMemcpyAsync(pat, hp.data(), sizeof(double) * cells, cudaMemcpyHostToDevice, st));
    }
    int rc;
    if (comm && (rc = comm_allreduce_sum_f64(comm, pat, cells, st)) != SFM_OK) return rc;
    SFM_CUDA_TRY(cudaMemcpyAsync(hp.data(), pat, sizeof(double) * cells, cudaMemcpyDeviceToHost, st));
    SFM_CUDA_TRY(cudaStreamSynchronize(st));
    SFM_CUDA_TRY(cudaFreeAsync(pat, st));
    std::vector<uint8_t> hb(cells);
    for (size_t e = 0; e < cells; ++e) hb[e] = hp[e] != 0.0;
    rc = SFM_OK;
    CholDagPlan *dag = chol_dag_create(hb.data(), nb, mode == 2 ? 1 : mode == 3 ? 2 : 0, st, &rc);
    if (rc != SFM_OK) return rc;
    std::lock_guard<std::mutex> g(g_plan_mu);
    bool stored = false;
    for (PlanKey &pk : g_plans)
        if (pk.ws == d_ws && pk.nc == pl.nc && pk.np == pl.np && pk.nobs == pl.nobs) {
            if (pk.dag) chol_dag_destroy(pk.dag, st);
            pk.dag = dag;
            pk.dag_tried = 1;
            stored = true;
        }
    if (!stored && dag) {                      // no cache entry to own it (cannot happen after ba_acquire_index)
        chol_dag_destroy(dag, st);
        dag = nullptr;
    }
    *out = dag;
    return SFM_OK;
}

// Drops what the library caches for ONE workspace (its index bookkeeping, the K10g pair list, the K11s plan), e.g.
// when the caller is about to free or reuse that memory. Stream-ordered like sfm_ba_release_cache.
extern "C" int sfm_ba_release_workspace(const void *d_ws, void *stream) {
    if (!d_ws) return SFM_E_INVALID;
    cudaStream_t st = as_stream(stream);
    std::lock_guard<std::mutex> g(g_plan_mu);
    for (size_t i = 0; i < g_plans.size(); ++i)
        if (g_plans[i].ws == d_ws) {
            PlanKey &q = g_plans[i];
            if (q.pair_keys) SFM_CUDA_TRY(cudaFreeAsync(q.pair_keys, st));
            if (q.pair_vals) SFM_CUDA_TRY(cudaFreeAsync(q.pair_vals, st));
            if (q.dag) chol_dag_destroy(q.dag, st);
            g_plans.erase(g_plans.begin() + i);
            break;
        }
    return SFM_OK;
}

namespace sfm { void record_error_text(const char *msg); }

static int ba_prepare(BAPlan &pl, void *d_ws, size_t ws_bytes, int32_t n_cams, int32_t n_pts, int32_t n_obs) {
    if (n_cams < 1 || n_pts < 1 || n_obs < 1 || !d_ws) return SFM_E_INVALID;
    {
        // One process per GPU (the deployment model: torch.distributed, one rank per device). The kernel attributes,
        // cooperative grid sizes, the side stream / events and the plan cache are set up once, on the device of the
        // first call; a call on another device would launch kernels without their shared-memory opt-in on a stream
        // of the wrong device, so it is refused instead.
        static std::mutex dev_mu;
        static int bound_dev = -1;
        int dev = -1;
        SFM_CUDA_TRY(cudaGetDevice(&dev));
        std::lock_guard<std::mutex> g(dev_mu);
        if (bound_dev < 0) bound_dev = dev;
        else if (dev != bound_dev) {
            char msg[160];
            snprintf(msg, sizeof(msg), "sfm_ba_*: this process bound the BA path to device %d on its first call; "
                     "current device is %d (one process per GPU)", bound_dev, dev);
            record_error_text(msg);
            return SFM_E_INVALID;
        }
    }
    unsigned char *base = reinterpret_cast<unsigned char *>(align_up((size_t)d_ws, 256));
    size_t need = ba_layout(pl, base, n_cams, n_pts, n_obs);
    if ((size_t)(base - (unsigned char *)d_ws) + need > ws_bytes) return SFM_E_WORKSPACE;
    if (!g_attrs_set) {
        set_obs_kernel_attrs<true>(BA_SMEM_MAX);
        (void)cudaGetLastError();
        g_attrs_set = true;
    }
    return SFM_OK;
}

extern "C" int sfm_ba_residuals(const double *d_cam, const double *d_pts, const double *d_obs,
                                const int32_t *d_cam_idx, const int32_t *d_pt_idx, int32_t n_cams,
                                int32_t n_pts, int32_t n_obs, const double *h_K, double *d_r,
                                double *d_cost, void *d_ws, size_t ws_bytes, void *stream) {
    if (!d_cam || !d_pts || !d_obs || !d_cam_idx || !d_pt_idx || !h_K) return SFM_E_INVALID;
    BAPlan pl;
    int rc = ba_prepare(pl, d_ws, ws_bytes, n_cams, n_pts, n_obs);
    if (rc != SFM_OK) return rc;
    cudaStream_t st = as_stream(stream);
    KParam K;
    for (int i = 0; i < 9; ++i) K.k[i] = h_K[i];
    SFM_CUDA_TRY(cudaMemsetAsync(pl.flags, 0, 64 * sizeof(int), st));
    ba_range_check_kernel<<<ceil_div(n_obs, 256), 256, 0, st>>>(d_cam_idx, n_obs, n_cams, pl.flags);
    SFM_LAUNCH_CHECK();
    ba_range_check_kernel<<<ceil_div(n_obs, 256), 256, 0, st>>>(d_pt_idx, n_obs, n_pts, pl.flags);
    SFM_LAUNCH_CHECK();
    int flag = 0;
    SFM_CUDA_TRY(cudaMemcpyAsync(&flag, pl.flags, sizeof(int), cudaMemcpyDeviceToHost, st));
    SFM_CUDA_TRY(cudaStreamSynchronize(st));
    if (flag) return SFM_E_INVALID;
    if ((rc = launch_cam_linearise(pl, K, d_cam, pl.camlin[0], st)) != SFM_OK) return rc;
    if ((rc = launch_cost(pl, pl.camlin[0], d_pts, reinterpret_cast<const double2 *>(d_obs), d_cam_idx,
                          d_pt_idx, reinterpret_cast<double2 *>(d_r), st)) != SFM_OK) return rc;
    ba_reduce_scalars_kernel<<<1, 256, 0, st>>>(pl.costpart, pl.grid_obs, nullptr, 0, nullptr, nullptr, nullptr,
                                                pl.scal);
    SFM_LAUNCH_CHECK();
    if (d_cost) SFM_CUDA_TRY(cudaMemcpyAsync(d_cost, pl.scal, sizeof(double), cudaMemcpyDeviceToDevice, st));
    return SFM_OK;
}

// The LM driver. comm == nullptr: single GPU. Otherwise this rank holds a contiguous range of the
// points (with all of their observations, local point indices) and ALL cameras; U/g_c/cost are
// all-reduced per linearisation, S/g_red and six scalars per trial step, and every rank runs the
// camera solve redundantly on identical bits.
static int ba_solve_impl(double *d_cam, double *d_pts, const double *d_obs, const int32_t *d_cam_idx,
                         const int32_t *d_pt_idx, int32_t n_cams, int32_t n_pts, int32_t n_obs,
                         const double *h_K, const sfm_ba_options *opts, sfm_ba_result *h_result,
                         void *d_ws, size_t ws_bytes, void *stream, void *comm, int rank) {
    if (!d_cam || !d_pts || !d_obs || !d_cam_idx || !d_pt_idx || !h_K || !h_result) return SFM_E_INVALID;
    sfm_ba_options o;
    if (opts) o = *opts; else sfm_ba_default_options(&o);
    BAPlan pl;
    int rc = ba_prepare(pl, d_ws, ws_bytes, n_cams, n_pts, n_obs);
    if (rc != SFM_OK) return rc;
    cudaStream_t st = as_stream(stream);
    KParam K;
    for (int i = 0; i < 9; ++i) K.k[i] = h_K[i];
    const double2 *obs = reinterpret_cast<const double2 *>(d_obs);
    memset(h_result, 0, sizeof(*h_result));
    h_result->status = -1;

    if ((rc = ba_acquire_index(pl, o.reuse_index != 0, d_ws, obs, d_cam_idx, d_pt_idx, n_cams, n_pts, n_obs, st)) != SFM_OK) return rc;
    const int nc = pl.nc, np = pl.np, nobs = pl.nobs, n = 6 * nc;
    CholDagPlan *dag = nullptr;
    if ((rc = ba_acquire_dag(pl, d_ws, comm, st, &dag)) != SFM_OK) return rc;
    // multi-GPU + K10g: block list of the sparse all-reduce (a collective: every rank takes this branch or none;
    // ranks whose shard is too small for K10g would break that, so the decision is all-reduced first)
    uint32_t *blk_list = nullptr;
    int n_blk = 0;
    struct BlkFree {                                   // released on every return path, stream-ordered
        uint32_t *&p; cudaStream_t s;
        ~BlkFree() { if (p) cudaFreeAsync(p, s); }
    } blk_free{blk_list, st};
    if (comm && nc <= BA_SPARSE_AR_MAX_CAMS && !(getenv("SFM_SPARSE_ALLREDUCE") && getenv("SFM_SPARSE_ALLREDUCE")[0] == '0')) {
        keep_async_pool_cached();
        const size_t T = (size_t)nc * (nc + 1) / 2;
        double *flag = nullptr;
        SFM_CUDA_TRY(cudaMallocAsync(&flag, sizeof(double) * (T + 1), st));
        SFM_CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(double) * (T + 1), st));
        // flag[T] counts the ranks that do NOT run K10g: any such rank sends the dense triangle
        if (!pl.npairs) {
            const double one = 1.0;
            SFM_CUDA_TRY(cudaMemcpyAsync(flag + T, &one, sizeof(double), cudaMemcpyHostToDevice, st));
        } else {
            const size_t ne = std::max(pl.npairs, (size_t)nc);
            ba_blk_mark_kernel<<<(unsigned)((ne + 255) / 256), 256, 0, st>>>(pl.pair_keys, pl.npairs, nc, flag);
            SFM_LAUNCH_CHECK();
        }
        if ((rc = comm_allreduce_sum_f64(comm, flag, T + 1, st)) != SFM_OK) return rc;
        double dense_ranks = 0.0;
        SFM_CUDA_TRY(cudaMemcpyAsync(&dense_ranks, flag + T, sizeof(double), cudaMemcpyDeviceToHost, st));
        SFM_CUDA_TRY(cudaStreamSynchronize(st));
        if (dense_ranks == 0.0) {
            int *f = nullptr, *pos = nullptr;
            void *tmp = nullptr;
            size_t tmp_bytes = 0;
            SFM_CUDA_TRY(cudaMallocAsync(&f, sizeof(int) * (T + 1), st));
            SFM_CUDA_TRY(cudaMallocAsync(&pos, sizeof(int) * (T + 1), st));
            ba_blk_flag_kernel<<<(unsigned)((T + 1 + 255) / 256), 256, 0, st>>>(flag, T, f);
            SFM_LAUNCH_CHECK();
            cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, f, pos, (int)(T + 1), st);
            SFM_CUDA_TRY(cudaMallocAsync(&tmp, tmp_bytes + 256, st));
            SFM_CUDA_TRY(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, f, pos, (int)(T + 1), st));
            count_launch();
            SFM_CUDA_TRY(cudaMemcpyAsync(&n_blk, pos + T, sizeof(int), cudaMemcpyDeviceToHost, st));
            SFM_CUDA_TRY(cudaStreamSynchronize(st));
            SFM_CUDA_TRY(cudaMallocAsync(&blk_list, sizeof(uint32_t) * (size_t)std::max(n_blk, 1), st));
            ba_blk_scatter_kernel<<<nc, 128, 0, st>>>(f, pos, nc, blk_list);
            SFM_LAUNCH_CHECK();
            SFM_CUDA_TRY(cudaFreeAsync(tmp, st));
            SFM_CUDA_TRY(cudaFreeAsync(f, st));
            SFM_CUDA_TRY(cudaFreeAsync(pos, st));
        }
        SFM_CUDA_TRY(cudaFreeAsync(flag, st));
    }
    SFM_CUDA_TRY(cudaMemcpyAsync(pl.cam[0], d_cam, sizeof(double) * 6 * nc, cudaMemcpyDeviceToDevice, st));
    SFM_CUDA_TRY(cudaMemcpyAsync(pl.pts[0], d_pts, sizeof(double) * 3 * (size_t)np, cudaMemcpyDeviceToDevice, st));
    int cur = 0;
    // pinned landing zone for the per-trial scalars: one per host thread, allocated once
    static thread_local double *h_scal = nullptr;
    if (!h_scal) SFM_CUDA_TRY(cudaMallocHost(&h_scal, 16 * sizeof(double)));

    // Single GPU: the camera pass (K9c + finalize/scale) and the point pass (K9p) touch disjoint outputs, so the
    // camera pass runs on a side stream forked from / joined to the caller's stream with events -- at C3 both
    // are ~14 us kernels that fill a fraction of the GPU. (Multi-GPU keeps one stream: the all-reduce follows.)
    static thread_local cudaStream_t side = nullptr;
    static thread_local cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    if (!comm && !side) {
        SFM_CUDA_TRY(cudaStreamCreateWithFlags(&side, cudaStreamNonBlocking));
        SFM_CUDA_TRY(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
        SFM_CUDA_TRY(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
    }
    // the work of one linearisation; `forked` tells the caller whether the side stream is running and unjoined
    auto linearise_body = [&](int first, bool &forked) -> int {
        int r;
        SFM_CUDA_TRY(cudaMemsetAsync(pl.gmax, 0, 8, st));
        cudaStream_t cs = st;                                  // stream of the camera pass
        if (!comm) {
            SFM_CUDA_TRY(cudaEventRecord(ev_fork, st));
            SFM_CUDA_TRY(cudaStreamWaitEvent(side, ev_fork, 0));
            cs = side;
            forked = true;
        }
        if (pl.nchunks) {
            ba_camera_pass_kernel<<<pl.nchunks, K9C_THREADS, 0, cs>>>(K, pl.camlin[cur], pl.pts[cur], pl.obs_c, pl.pt_c,
                                                              pl.chunk_cam, pl.chunk_beg, pl.chunk_end, pl.Upart);
            SFM_LAUNCH_CHECK();
        }
        // single GPU: the camera scaling rides in the finalize launch; multi-GPU: it needs the all-reduced U, g_c
        ba_cam_finalize_kernel<<<ceil_div(nc, 8), 256, 0, cs>>>(pl.Upart, pl.chunk_ptr, nc, pl.U, pl.gc, first,
                                                               comm ? nullptr : pl.Dc, pl.gmax);
        SFM_LAUNCH_CHECK();
        if ((r = launch_point_pass(pl, K, pl.camlin[cur], pl.pts[cur], obs, d_cam_idx, d_pt_idx, st)) != SFM_OK) return r;
        ba_reduce_scalars_kernel<<<1, 256, 0, st>>>(pl.costpart, pl.grid_obs, nullptr, 0, nullptr, nullptr, nullptr, pl.red1);
        SFM_LAUNCH_CHECK();
        if (comm) {
            if ((r = comm_allreduce_sum_f64(comm, pl.U, 27 * (size_t)nc + 1, st)) != SFM_OK) return r;
            ba_cam_scale_kernel<<<ceil_div(nc, 128), 128, 0, st>>>(pl.U, pl.gc, nc, first, pl.Dc, pl.gmax);
            SFM_LAUNCH_CHECK();
        }
        return SFM_OK;
    };
    auto linearise = [&](int first) -> int {
        bool forked = false;
        const int r = linearise_body(first, forked);
        if (forked) {
            // join on EVERY path: an error return must not leave the caller's stream ahead of the side stream
            // (the next call would fork from a stream the side stream is not ordered against)
            cudaError_t e1 = cudaEventRecord(ev_join, side), e2 = cudaStreamWaitEvent(st, ev_join, 0);
            if (r == SFM_OK && (e1 != cudaSuccess || e2 != cudaSuccess)) {
                record_cuda_error(e1 != cudaSuccess ? e1 : e2, "join of the camera-pass stream");
                return SFM_E_CUDA;
            }
        }
        return r;
    };

    static thread_local cudaEvent_t pe_lin0[2] = {};
    const bool prof0 = (o.reserved & 1) != 0;
    if (prof0 && !pe_lin0[0]) { SFM_CUDA_TRY(cudaEventCreate(&pe_lin0[0])); SFM_CUDA_TRY(cudaEventCreate(&pe_lin0[1])); }
    if (prof0) cudaEventRecord(pe_lin0[0], st);
    if ((rc = launch_cam_linearise(pl, K, pl.cam[0], pl.camlin[0], st)) != SFM_OK) return rc;
    if ((rc = linearise(1)) != SFM_OK) return rc;
    if (prof0) cudaEventRecord(pe_lin0[1], st);
    SFM_CUDA_TRY(cudaMemcpyAsync(h_scal, pl.red1, sizeof(double), cudaMemcpyDeviceToHost, st));
    SFM_CUDA_TRY(cudaStreamSynchronize(st));
    double cost = h_scal[0];
    h_result->cost0 = h_result->cost = cost;
    h_result->n_linearisations = 1;
    if (!(cost <= 1.7e308)) {            // NaN or inf: scipy raises, the reference returns its inputs
        h_result->status = -1;
        return SFM_OK;
    }

    // opts.reserved bit 0: phase profile of this solve (CUDA events on the caller's stream, read at the host
    // synchronisation every trial step already has): sfm_ba_last_profile()
    const bool prof = (o.reserved & 1) != 0;
    static thread_local cudaEvent_t pe[7] = {};
    if (prof && !pe[0])
        for (int i = 0; i < 7; ++i) SFM_CUDA_TRY(cudaEventCreate(&pe[i]));
    double pacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    bool lin_pending = false;
    auto mark = [&](int i) { if (prof) cudaEventRecord(pe[i], st); };
    auto span = [&](int a, int b) { float ms = 0.f; cudaEventElapsedTime(&ms, pe[a], pe[b]); return (double)ms; };
    double lambda = o.lambda0 > 0 ? o.lambda0 : 1e-4, nu = 2.0, xnorm = 0.0;
    int status = -2, trials = 0, first_lin = 1, update_D = 1, chol_fail = 0;
    const size_t s_bytes = sizeof(double) * (size_t)(pl.npad + BA_NB) * pl.ld;
    while (status == -2) {
        // ---- one trial step with damping lambda ----
        mark(0);
        // with K11s only the cells of its pattern are ever read (every Schur kernel writes inside the pattern: it was
        // built from the same pair list); the dense kernel and the dense all-reduce need the whole array
        if (dag && (!comm || blk_list)) {
            if ((rc = chol_dag_zero_system(dag, pl.S, pl.ld, st)) != SFM_OK) return rc;
        } else {
            SFM_CUDA_TRY(cudaMemsetAsync(pl.S, 0, s_bytes, st));
        }
        SFM_CUDA_TRY(cudaMemsetAsync(pl.flags + 1, 0, 2 * sizeof(int), st));
        SFM_CUDA_TRY(cudaMemsetAsync(pl.tp, 0, sizeof(double) * 3 * (size_t)np, st));
        if (rank == 0) {      // the camera blocks U + lambda Dc^2, g_c and the padding enter the summed S once
            ba_schur_init_kernel<<<ceil_div(std::max(nc * 36, pl.npad), 256), 256, 0, st>>>(pl.U, pl.gc, pl.Dc, nc, pl.npad,
                                                                                           pl.ld, lambda, pl.S);
            SFM_LAUNCH_CHECK();
        } else if (blk_list) {
            ba_schur_init_kernel<<<ceil_div(pl.npad, 256), 256, 0, st>>>(pl.U, pl.gc, pl.Dc, nc, pl.npad, pl.ld, lambda, pl.S, 1);
            SFM_LAUNCH_CHECK();
        }
        SchurRanges rg;
        if (pl.npairs) {
            // K10g: prep (V^-1, b_p, G), Z records + right-hand side, gather over the sorted pair list
            ba_point_prep_kernel<<<ceil_div(np, 256), 256, 0, st>>>(pl.V, pl.gp, pl.Dp, np, lambda, update_D, first_lin,
                                                                   pl.Vinv, pl.bp, pl.gmax, pl.flags + 1, pl.Ginv);
            SFM_LAUNCH_CHECK();
            ba_schur_z_kernel<<<ceil_div(nobs, 256), 256, 0, st>>>(pl.W, nobs, d_pt_idx, pl.iota, pl.Ginv, pl.Z);
            SFM_LAUNCH_CHECK();
            static bool sg_attr = false;
            if (!sg_attr) {
                cudaFuncSetAttribute(ba_schur_gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SG_SMEM);
                sg_attr = true;
            }
            const size_t nwarps = pl.npairs / ((size_t)pl.pair_seg * 32);
            ba_schur_gather_kernel<<<(unsigned)(nwarps / (SG_THREADS / 32)), SG_THREADS, SG_SMEM, st>>>(
                pl.pair_keys, pl.pair_vals, pl.pair_seg, pl.Z, pl.pt_c, pl.Ginv + 6 * (size_t)np, nc, pl.S, pl.ld,
                pl.S + (size_t)pl.npad * pl.ld);
            SFM_LAUNCH_CHECK();
        } else if (!schur_by_rows() && schur_smem_ranges(nc, pl.max_track, rg)) {
            ba_point_prep_kernel<<<ceil_div(np, 256), 256, 0, st>>>(pl.V, pl.gp, pl.Dp, np, lambda, update_D, first_lin,
                                                                   pl.Vinv, pl.bp, pl.gmax, pl.flags + 1);
            SFM_LAUNCH_CHECK();
            const int nslices = std::max(1, sm_count() / rg.n);
            size_t smem = 0;
            for (int t = 0; t < rg.n; ++t) {
                const size_t b = (size_t)rg.row[t + 1] * (rg.row[t + 1] + 1) / 2 - (size_t)rg.row[t] * (rg.row[t] + 1) / 2;
                smem = std::max(smem, b * 36 * sizeof(double) + SS_STAGE_BYTES);
            }
            static int ss_skip = getenv("SFM_SS_SKIP") ? atoi(getenv("SFM_SS_SKIP")) : 0;   // timing experiments only
            static bool attr = false;
            if (!attr) {
                cudaFuncSetAttribute(ba_schur_smem_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BA_SMEM_MAX);
                cudaFuncSetAttribute(ba_schur_smem_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BA_SMEM_MAX);
                attr = true;
            }
            if (pl.has_dup)
                ba_schur_smem_kernel<true><<<rg.n * nslices, SS_THREADS, smem, st>>>(
                    pl.W, nobs, d_cam_idx, d_pt_idx, pl.pt_ptr, np, pl.Vinv, pl.bp, rg, nslices, pl.S, pl.ld,
                    pl.S + (size_t)pl.npad * pl.ld, ss_skip);
            else
                ba_schur_smem_kernel<false><<<rg.n * nslices, SS_THREADS, smem, st>>>(
                    pl.W, nobs, d_cam_idx, d_pt_idx, pl.pt_ptr, np, pl.Vinv, pl.bp, rg, nslices, pl.S, pl.ld,
                    pl.S + (size_t)pl.npad * pl.ld, ss_skip);
            SFM_LAUNCH_CHECK();
        } else if (schur_by_rows()) {
            ba_point_prep_kernel<<<ceil_div(np, 256), 256, 0, st>>>(pl.V, pl.gp, pl.Dp, np, lambda, update_D, first_lin,
                                                                   pl.Vinv, pl.bp, pl.gmax, pl.flags + 1);
            SFM_LAUNCH_CHECK();
            if (pl.nchunks) {
                const int WB = schur_window(nc);
                const size_t smem = sizeof(double) * SR_WARPS * ((size_t)WB * 36 + 24);
                const int slice = schur_slice(nobs);
                dim3 grid(pl.nchunks, ceil_div(BA_CHUNK, SR_WARPS * slice));
                ba_schur_rows_kernel<<<grid, SR_WARPS * 32, smem, st>>>(
                    pl.W, nobs, d_cam_idx, pl.pt_ptr, pl.perm, pl.pt_c, pl.chunk_cam, pl.chunk_beg, pl.chunk_end,
                    pl.Vinv, pl.bp, WB, slice, pl.S, pl.ld, pl.S + (size_t)pl.npad * pl.ld);
                SFM_LAUNCH_CHECK();
            }
        } else {
            int grid = std::max(1, std::min(sm_count() * 8, ceil_div(np, SCH_WARPS)));
            ba_schur_kernel<<<grid, SCH_WARPS * 32, 0, st>>>(pl.W, nobs, d_cam_idx, pl.pt_ptr, np, pl.V, pl.gp, pl.Dp,
                                                             lambda, update_D, first_lin, pl.Vinv, pl.bp, pl.S, pl.ld,
                                                             pl.S + (size_t)pl.npad * pl.ld, pl.gmax, pl.flags + 1);
            SFM_LAUNCH_CHECK();
        }
        mark(1);
        if (comm && blk_list) {
            const size_t packed = (size_t)n_blk * 36 + n;
            const unsigned grid = (unsigned)((packed + 255) / 256);
            ba_s_pack_sparse_kernel<true><<<grid, 256, 0, st>>>(pl.S, pl.ld, pl.npad, n, blk_list, n_blk, pl.Spack);
            SFM_LAUNCH_CHECK();
            if ((rc = comm_allreduce_sum_f64(comm, pl.Spack, packed, st)) != SFM_OK) return rc;
            ba_s_pack_sparse_kernel<false><<<grid, 256, 0, st>>>(pl.S, pl.ld, pl.npad, n, blk_list, n_blk, pl.Spack);
            SFM_LAUNCH_CHECK();
        } else if (comm) {
            const size_t packed = s_pack_offset(pl.npad, pl.npad) + pl.npad;
            ba_s_pack_kernel<true><<<pl.npad + 1, 256, 0, st>>>(pl.S, pl.ld, pl.npad, pl.Spack);
            SFM_LAUNCH_CHECK();
            if ((rc = comm_allreduce_sum_f64(comm, pl.Spack, packed, st)) != SFM_OK) return rc;
            ba_s_pack_kernel<false><<<pl.npad + 1, 256, 0, st>>>(pl.S, pl.ld, pl.npad, pl.Spack);
            SFM_LAUNCH_CHECK();
        }
        if (update_D && (rc = comm_allreduce_max_u64(comm, pl.gmax, 1, st)) != SFM_OK) return rc;
        mark(2);
        static unsigned long long *d_cprof = nullptr;          // SFM_BA_CHOL_PROF=1: K11 phase profile (diagnostic)
        static int cprof = -1;
        if (cprof < 0) {
            cprof = getenv("SFM_BA_CHOL_PROF") ? 1 : 0;
            if (cprof && (cudaMalloc(&d_cprof, 8 * sizeof(unsigned long long)) != cudaSuccess ||
                          cudaMemset(d_cprof, 0, 8 * sizeof(unsigned long long)) != cudaSuccess)) cprof = 0;
        }
        if (dag) rc = chol_dag_launch(dag, pl.S, pl.ld, pl.npad, pl.dc, pl.flags + 2, st);
        else rc = chol_solve_launch(pl.S, pl.ld, pl.npad, pl.Linv, pl.dc, pl.flags + 2, st, cprof ? d_cprof : nullptr);
        if (rc != SFM_OK) return rc;
        mark(3);
        if (cprof) {
            unsigned long long hp[8];
            cudaStreamSynchronize(st);
            cudaMemcpy(hp, d_cprof, sizeof(hp), cudaMemcpyDeviceToHost);
            cudaMemset(d_cprof, 0, sizeof(hp));
            fprintf(stderr, "K11 phases [us]: diag %.1f trsm %.1f sync1 %.1f update %.1f sync2 %.1f back %.1f\n", hp[0] * 1e-3,
                    hp[1] * 1e-3, hp[2] * 1e-3, hp[3] * 1e-3, hp[4] * 1e-3, hp[5] * 1e-3);
        }
        const int nxt = cur ^ 1;
        const UpdateCamsArgs uc{pl.cam[cur], pl.dc, pl.gc, pl.Dc, n, rank == 0, pl.cam[nxt], pl.updpart};
        ba_point_rhs_kernel<<<ceil_div(nobs, 256) + 1, 256, 0, st>>>(pl.W, nobs, d_cam_idx, d_pt_idx, pl.pt_ptr, pl.dc, pl.tp, uc);
        SFM_LAUNCH_CHECK();
        ba_update_points_kernel<<<pl.grid_pts, 256, 0, st>>>(pl.pts[cur], pl.Vinv, pl.bp, pl.tp, pl.gp, pl.Dp, np,
                                                             pl.pts[nxt], pl.updpart);
        SFM_LAUNCH_CHECK();
        if ((rc = launch_cam_linearise(pl, K, pl.cam[nxt], pl.camlin[nxt], st)) != SFM_OK) return rc;
        if ((rc = launch_cost(pl, pl.camlin[nxt], pl.pts[nxt], obs, d_cam_idx, d_pt_idx, nullptr, st)) != SFM_OK) return rc;
        ba_reduce_scalars_kernel<<<1, 256, 0, st>>>(pl.costpart, pl.grid_obs, pl.updpart, pl.grid_pts + 1, pl.gmax,
                                                    pl.flags + 1, pl.flags + 2, pl.scal);
        SFM_LAUNCH_CHECK();
        if ((rc = comm_allreduce_sum_f64(comm, pl.scal, 6, st)) != SFM_OK) return rc;
        mark(4);
        SFM_CUDA_TRY(cudaMemcpyAsync(h_scal, pl.scal, 8 * sizeof(double), cudaMemcpyDeviceToHost, st));
        SFM_CUDA_TRY(cudaStreamSynchronize(st));
        if (prof) {
            pacc[0] += span(0, 1); pacc[1] += span(1, 2); pacc[2] += span(2, 3); pacc[3] += span(3, 4);
            if (lin_pending) { pacc[4] += span(5, 0); lin_pending = false; }
        }
        ++trials;
        const double cost_new = h_scal[0], sg = h_scal[1], Ds2 = h_scal[2], Dx2n = h_scal[3], Dx2c = h_scal[4];
        const double gmax = h_scal[6];
        const int fail = (h_scal[5] != 0.0 ? 2 : 0) | (h_scal[7] != 0.0 ? 4 : 0);
        const double fnorm2 = 2.0 * cost;
        if (update_D) {
            // MINPACK gtol test (lmder, top of the outer loop): max_j |J_j^T f| / (|J_j| |f|)
            double gnorm = fnorm2 > 0.0 ? gmax / sqrt(fnorm2) : 0.0;
            h_result->gnorm = gnorm;
            if (gnorm <= o.gtol) { status = 1; break; }
        }
        update_D = 0;
        bool solve_ok = fail == 0 && cost_new == cost_new && Ds2 == Ds2 && sg == sg;
        if (fail & 4) ++chol_fail;
        double actred = -1.0, prered = 0.0, ratio = 0.0;
        if (solve_ok) {
            if (cost_new <= 1.7e308 && 0.01 * cost_new < cost) actred = 1.0 - cost_new / cost;
            prered = (sg + lambda * Ds2) / fnorm2;
            ratio = prered > 0.0 ? actred / prered : 0.0;
        }
        const bool accept = solve_ok && ratio >= 1e-4;
        if (o.verbose)
            fprintf(stderr, "[sfm_ba] lin %d trial %d lambda %.3e cost %.12g -> %.12g ratio %.4f |Ds| %.3e fail %d %s\n",
                    h_result->n_linearisations, trials, lambda, cost, cost_new, ratio, sqrt(Ds2 > 0 ? Ds2 : 0), fail,
                    accept ? "accept" : "reject");
        if (accept) {
            cur = nxt;
            cost = cost_new;
            xnorm = sqrt(Dx2n);
            double t = 2.0 * ratio - 1.0;
            lambda *= fmax(1.0 / 3.0, 1.0 - t * t * t);
            nu = 2.0;
        } else {
            xnorm = sqrt(Dx2c);
            lambda *= nu;
            nu *= 2.0;
        }
        if (lambda < 1e-20) lambda = 1e-20;
        if (solve_ok) {
            bool f = fabs(actred) <= o.ftol && prered <= o.ftol && 0.5 * ratio <= 1.0;
            bool x = sqrt(Ds2) <= o.xtol * xnorm;
            if (f && x) status = 4; else if (f) status = 2; else if (x) status = 3;
        }
        if (status == -2 && (trials >= o.max_trials || lambda > 1e30)) status = 0;
        if (status != -2) break;
        if (accept) {
            first_lin = 0;
            update_D = 1;
            mark(5);
            if ((rc = linearise(0)) != SFM_OK) return rc;
            lin_pending = prof;
            h_result->n_linearisations += 1;
        }
    }
    SFM_CUDA_TRY(cudaMemcpyAsync(d_cam, pl.cam[cur], sizeof(double) * 6 * nc, cudaMemcpyDeviceToDevice, st));
    SFM_CUDA_TRY(cudaMemcpyAsync(d_pts, pl.pts[cur], sizeof(double) * 3 * (size_t)np, cudaMemcpyDeviceToDevice, st));
    SFM_CUDA_TRY(cudaStreamSynchronize(st));
    if (prof) {
        float ms0 = 0.f;
        cudaEventElapsedTime(&ms0, pe_lin0[0], pe_lin0[1]);
        pacc[4] += ms0;
        pacc[5] = trials; pacc[6] = h_result->n_linearisations;
        pacc[7] = blk_list ? ((double)n_blk * 36 + n) * 8.0 : (comm ? (double)(s_pack_offset(pl.npad, pl.npad) + pl.npad) * 8.0 : 0.0);
        std::lock_guard<std::mutex> g(g_plan_mu);
        for (int i = 0; i < 8; ++i) g_ba_prof[i] = pacc[i];
    }
    h_result->cost = cost;
    h_result->status = status;
    h_result->n_trials = trials;
    h_result->n_chol_fail = chol_fail;
    h_result->lambda = lambda;
    return SFM_OK;
}

extern "C" int sfm_ba_last_profile(double *h_out8) {
    if (!h_out8) return SFM_E_INVALID;
    std::lock_guard<std::mutex> g(g_plan_mu);
    for (int i = 0; i < 8; ++i) h_out8[i] = g_ba_prof[i];
    return SFM_OK;
}

extern "C" int sfm_ba_solve(double *d_cam, double *d_pts, const double *d_obs, const int32_t *d_cam_idx,
                            const int32_t *d_pt_idx, int32_t n_cams, int32_t n_pts, int32_t n_obs,
                            const double *h_K, const sfm_ba_options *opts, sfm_ba_result *h_result,
                            void *d_ws, size_t ws_bytes, void *stream) {
    return ba_solve_impl(d_cam, d_pts, d_obs, d_cam_idx, d_pt_idx, n_cams, n_pts, n_obs, h_K, opts, h_result, d_ws,
                         ws_bytes, stream, nullptr, 0);
}

extern "C" int sfm_ba_solve_mgpu(double *d_cam, double *d_pts_local, const double *d_obs_local,
                                 const int32_t *d_cam_idx_local, const int32_t *d_pt_idx_local, int32_t n_cams,
                                 int32_t n_pts_local, int32_t n_obs_local, const double *h_K,
                                 const sfm_ba_options *opts, sfm_ba_result *h_result, void *d_ws, size_t ws_bytes,
                                 void *comm, int32_t rank, int32_t world, void *stream) {
    if (world < 1 || rank < 0 || rank >= world || (world > 1 && !comm)) return SFM_E_INVALID;
    return ba_solve_impl(d_cam, d_pts_local, d_obs_local, d_cam_idx_local, d_pt_idx_local, n_cams, n_pts_local,
                         n_obs_local, h_K, opts, h_result, d_ws, ws_bytes, stream, world > 1 ? comm : nullptr, rank);
}

// One linearisation pass in isolation (bench / profiling): which = bit 0 K9p point pass, bit 1 K9c
// camera pass + finalize, bit 2 K8 cost pass, bit 3 skip the per-camera linearisation kernel (reuse
// the table of the previous call). Reads d_cam / d_pts in place; outputs stay in the workspace.
extern "C" int sfm_ba_linearise(const double *d_cam, const double *d_pts, const double *d_obs,
                                const int32_t *d_cam_idx, const int32_t *d_pt_idx, int32_t n_cams,
                                int32_t n_pts, int32_t n_obs, const double *h_K, int32_t which,
                                int32_t reuse_index, void *d_ws, size_t ws_bytes, void *stream) {
    if (!d_cam || !d_pts || !d_obs || !d_cam_idx || !d_pt_idx || !h_K) return SFM_E_INVALID;
    BAPlan pl;
    int rc = ba_prepare(pl, d_ws, ws_bytes, n_cams, n_pts, n_obs);
    if (rc != SFM_OK) return rc;
    cudaStream_t st = as_stream(stream);
    KParam K;
    for (int i = 0; i < 9; ++i) K.k[i] = h_K[i];
    const double2 *obs = reinterpret_cast<const double2 *>(d_obs);
    // bit 3 (same state as the previous call, for timing one pass alone) also skips the refresh of the camera-major
    // observation copy
    if ((rc = ba_acquire_index(pl, reuse_index != 0, d_ws, obs, d_cam_idx, d_pt_idx, n_cams, n_pts, n_obs, st,
                               !(which & 8))) != SFM_OK)
        return rc;
    if (!(which & 8) && (rc = launch_cam_linearise(pl, K, d_cam, pl.camlin[0], st)) != SFM_OK) return rc;
    if (which & 1) {
        if ((rc = launch_point_pass(pl, K, pl.camlin[0], d_pts, obs, d_cam_idx, d_pt_idx, st)) != SFM_OK) return rc;
    }
    if (which & 2) {
        SFM_CUDA_TRY(cudaMemsetAsync(pl.gmax, 0, 8, st));
        if (pl.nchunks) {
            ba_camera_pass_kernel<<<pl.nchunks, K9C_THREADS, 0, st>>>(K, pl.camlin[0], d_pts, pl.obs_c, pl.pt_c, pl.chunk_cam,
                                                              pl.chunk_beg, pl.chunk_end, pl.Upart);
            SFM_LAUNCH_CHECK();
        }
        ba_cam_finalize_kernel<<<ceil_div(pl.nc, 8), 256, 0, st>>>(pl.Upart, pl.chunk_ptr, pl.nc, pl.U, pl.gc, 1, pl.Dc,
                                                                  pl.gmax);
        SFM_LAUNCH_CHECK();
    }
    if (which & 4) {
        if ((rc = launch_cost(pl, pl.camlin[0], d_pts, obs, d_cam_idx, d_pt_idx, nullptr, st)) != SFM_OK) return rc;
    }
    return SFM_OK;
}
