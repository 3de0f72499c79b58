// h_ransac.cu -- homography RANSAC on the K1..K4 skeleton of f_ransac.cu.
//
//   h4pt_kernel     find_homography (GetHomographyInliers.py:4-82) for every sampled 4-tuple
//   h_score_kernel  :139-150: inlier count of every hypothesis over all correspondences
//   argmax          :153 strict '>' = first maximum (sfm_argmax_first)
//   h_mask_kernel   :149,155 for the winner
//
// It precedes every F-RANSAC call of the wrapper (Wrapper_dev.py:87,105) and draws from the same
// random.sample stream (4 indices per iteration), which the host replays (sampler.py).
#include "h_math.cuh"
#include "../../include/sfm_b200.h"

namespace sfm {

__global__ void __launch_bounds__(128)
h4pt_kernel(const double *__restrict__ x1x2, const int32_t *__restrict__ sample, int h, double *__restrict__ Hout) {
    const int hyp = blockIdx.x * blockDim.x + threadIdx.x;
    if (hyp >= h) return;
    const int4 s = __ldg(reinterpret_cast<const int4 *>(sample + (size_t)hyp * 4));
    const int idx[4] = {s.x, s.y, s.z, s.w};
    double px1[4], py1[4], px2[4], py2[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const double2 *row = reinterpret_cast<const double2 *>(x1x2 + (size_t)idx[i] * 4);
        const double2 a = __ldg(row), b = __ldg(row + 1);
        px1[i] = a.x; py1[i] = a.y; px2[i] = b.x; py2[i] = b.y;
    }
    double H[9];
    h4pt_from_sample(px1, py1, px2, py2, H);
#pragma unroll
    for (int i = 0; i < 9; ++i) Hout[(size_t)hyp * 9 + i] = H[i];
}

constexpr int HK2_THREADS = 256;

// thread per hypothesis (H in 9 registers) x block-staged chunk of correspondences, as K2
__global__ void __launch_bounds__(HK2_THREADS)
h_score_kernel(const double *__restrict__ x1x2, int n, const double *__restrict__ Hall, int h, double thr2,
               int chunk, int32_t *__restrict__ counts) {
    extern __shared__ double2 spts[];
    const int p0 = blockIdx.y * chunk;
    const int np = min(chunk, n - p0);
    {
        const double2 *src = reinterpret_cast<const double2 *>(x1x2) + (size_t)p0 * 2;
        for (int i = threadIdx.x; i < np * 2; i += HK2_THREADS) spts[i] = __ldg(src + i);
    }
    const int hyp = blockIdx.x * HK2_THREADS + threadIdx.x;
    HMat H;
    {
        const double *f = Hall + (size_t)min(hyp, h - 1) * 9;
        H.h0 = __ldg(f + 0); H.h1 = __ldg(f + 1); H.h2 = __ldg(f + 2);
        H.h3 = __ldg(f + 3); H.h4 = __ldg(f + 4); H.h5 = __ldg(f + 5);
        H.h6 = __ldg(f + 6); H.h7 = __ldg(f + 7); H.h8 = __ldg(f + 8);
    }
    __syncthreads();
    int cnt = 0;
#pragma unroll 4
    for (int i = 0; i < np; ++i) {
        const double2 a = spts[2 * i], b = spts[2 * i + 1];
        cnt += h_is_inlier(H, a.x, a.y, b.x, b.y, thr2) ? 1 : 0;
    }
    if (hyp < h && cnt) atomicAdd(&counts[hyp], cnt);
}

// winner through a device index: Hone = Hbase + 9 * best[0] when best is given
__global__ void __launch_bounds__(256)
h_mask_kernel(const double *__restrict__ x1x2, int n, const double *__restrict__ Hbase,
              const int32_t *__restrict__ best, int h, double thr2, uint8_t *__restrict__ mask,
              double *__restrict__ err, double *__restrict__ Hcopy) {
    int sel = 0;
    bool valid = true;
    if (best) {
        sel = best[0];
        valid = best[1] > 0 && sel >= 0 && sel < h;
        if (!valid) sel = 0;
    }
    const double *f = Hbase + (size_t)sel * 9;
    const HMat H = {f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7], f[8]};
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (Hcopy && i < 9) Hcopy[i] = f[i];
    if (i >= n) return;
    const double2 *row = reinterpret_cast<const double2 *>(x1x2) + (size_t)i * 2;
    const double2 a = __ldg(row), b = __ldg(row + 1);
    mask[i] = (valid && h_is_inlier(H, a.x, a.y, b.x, b.y, thr2)) ? 1 : 0;
    if (err) err[i] = h_error_exact(H, a.x, a.y, b.x, b.y);
}

static int h_pick_chunk(int n, int h) {
    int tiles = ceil_div(h, HK2_THREADS);
    int chunk = 1024;
    while (chunk > 128 && (long long)tiles * ceil_div(n, chunk) < 4LL * sm_count()) chunk >>= 1;
    return chunk;
}

static double thr_squared(double thr) { return thr > 0.0 ? thr * thr : -1.0; }

}  // namespace sfm

using namespace sfm;

extern "C" int sfm_h4pt_batch(const double *d_x1x2, int32_t n, const int32_t *d_sample, int32_t h, double *d_H,
                              void *stream) {
    if (!d_x1x2 || !d_sample || !d_H || n < 4 || h < 0) return SFM_E_INVALID;
    if (h == 0) return SFM_OK;
    h4pt_kernel<<<ceil_div(h, 128), 128, 0, as_stream(stream)>>>(d_x1x2, d_sample, h, d_H);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_h_score_count(const double *d_x1x2, int32_t n, const double *d_H, int32_t h, double threshold,
                                 int32_t *d_counts, void *stream) {
    if (!d_x1x2 || !d_H || !d_counts || n < 0 || h < 0) return SFM_E_INVALID;
    cudaStream_t st = as_stream(stream);
    if (h == 0) return SFM_OK;
    SFM_CUDA_TRY(cudaMemsetAsync(d_counts, 0, sizeof(int32_t) * (size_t)h, st));
    if (n == 0) return SFM_OK;
    const int chunk = h_pick_chunk(n, h);
    dim3 grid(ceil_div(h, HK2_THREADS), ceil_div(n, chunk));
    h_score_kernel<<<grid, HK2_THREADS, sizeof(double2) * 2 * (size_t)chunk, st>>>(d_x1x2, n, d_H, h,
                                                                                    thr_squared(threshold), chunk, d_counts);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_h_inlier_mask(const double *d_x1x2, int32_t n, const double *d_Hone, double threshold,
                                 uint8_t *d_mask, double *d_err, void *stream) {
    if (!d_x1x2 || !d_Hone || !d_mask || n < 0) return SFM_E_INVALID;
    if (n == 0) return SFM_OK;
    h_mask_kernel<<<ceil_div(n, 256), 256, 0, as_stream(stream)>>>(d_x1x2, n, d_Hone, nullptr, 1,
                                                                  thr_squared(threshold), d_mask, d_err, nullptr);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_h_ransac(const double *d_x1x2, int32_t n, const int32_t *d_sample, int32_t h, double threshold,
                            double *d_H, int32_t *d_counts, int32_t *d_best, double *d_Hbest, uint8_t *d_mask,
                            double *d_err, void *stream) {
    if (!d_x1x2 || !d_sample || !d_H || !d_counts || !d_best || !d_Hbest || !d_mask || n < 4 || h < 1)
        return SFM_E_INVALID;
    int rc;
    if ((rc = sfm_h4pt_batch(d_x1x2, n, d_sample, h, d_H, stream)) != SFM_OK) return rc;
    if ((rc = sfm_h_score_count(d_x1x2, n, d_H, h, threshold, d_counts, stream)) != SFM_OK) return rc;
    if ((rc = sfm_argmax_first(d_counts, h, 0, d_best, nullptr, stream)) != SFM_OK) return rc;
    h_mask_kernel<<<ceil_div(std::max(n, 9), 256), 256, 0, as_stream(stream)>>>(
        d_x1x2, n, d_H, d_best, h, thr_squared(threshold), d_mask, d_err, d_Hbest);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}
