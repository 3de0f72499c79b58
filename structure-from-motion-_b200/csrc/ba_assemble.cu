// ba_assemble.cu -- SURVEY.md 8(f) N3: the observation list of a bundle-adjustment problem built on the
// device from the wrapper's dense (n_features x n_images) arrays.
//
// Replaces BundleAdjustment.py:148-176: the reference walks every valid feature row and every camera in
// Python and locates the point's compact index with np.where inside the loop (O(Nobs * Np)). Here:
//   1. thread per feature row: valid = (filtered_world_coords == 1); count of cameras j < n_cams with
//      filtered_feature_flags[r, j] == 1
//   2. two exclusive scans (CUB): compact point index of every valid row, first observation of every row
//   3. thread per row: scatter (camera, point, x, y) in row order, cameras ascending -- the reference's
//      order (point-major), so the result feeds sfm_ba_solve directly
// Flags are uint8 on the device (the host wrapper converts whatever dtype numpy holds with `== 1`).
#include "sfm_common.cuh"
#include <cub/device/device_scan.cuh>

namespace sfm {

__global__ void __launch_bounds__(256)
asm_count_kernel(const uint8_t *__restrict__ valid, const uint8_t *__restrict__ flags, int F, int stride, int n_cams,
                 int32_t *__restrict__ rowcnt, int32_t *__restrict__ rowvalid) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= F) return;
    const int v = valid[r] == 1;
    int c = 0;
    if (v)
        for (int j = 0; j < n_cams; ++j) c += flags[(size_t)r * stride + j] == 1;
    rowcnt[r] = c;
    rowvalid[r] = v;
}

__global__ void __launch_bounds__(256)
asm_scatter_kernel(const uint8_t *__restrict__ flags, const double *__restrict__ fx, const double *__restrict__ fy,
                   const double *__restrict__ world, int F, int stride, int n_cams,
                   const int32_t *__restrict__ rowcnt, const int32_t *__restrict__ rowvalid,
                   const int32_t *__restrict__ rowoff, const int32_t *__restrict__ ptlocal,
                   int32_t *__restrict__ cam_idx, int32_t *__restrict__ pt_idx, double *__restrict__ obs,
                   double *__restrict__ pts, int32_t *__restrict__ valid_rows, int32_t *__restrict__ totals) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= F) return;
    if (r == F - 1) {
        totals[0] = ptlocal[r] + rowvalid[r];       // number of points
        totals[1] = rowoff[r] + rowcnt[r];          // number of observations
    }
    if (!rowvalid[r]) return;
    const int p = ptlocal[r];
    valid_rows[p] = r;
    pts[3 * (size_t)p] = world[3 * (size_t)r];
    pts[3 * (size_t)p + 1] = world[3 * (size_t)r + 1];
    pts[3 * (size_t)p + 2] = world[3 * (size_t)r + 2];
    int o = rowoff[r];
    for (int j = 0; j < n_cams; ++j) {
        if (flags[(size_t)r * stride + j] == 1) {
            cam_idx[o] = j;
            pt_idx[o] = p;
            obs[2 * (size_t)o] = fx[(size_t)r * stride + j];
            obs[2 * (size_t)o + 1] = fy[(size_t)r * stride + j];
            ++o;
        }
    }
}

static size_t asm_scan_bytes(int F) {
    size_t b = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, b, (const int32_t *)nullptr, (int32_t *)nullptr, F);
    return b;
}

}  // namespace sfm

using namespace sfm;

extern "C" size_t sfm_ba_assemble_workspace_bytes(int32_t n_features) {
    if (n_features <= 0) return 0;
    return 4 * align_up(sizeof(int32_t) * (size_t)n_features, 256) + align_up(asm_scan_bytes(n_features), 256) + 256;
}

extern "C" int sfm_ba_assemble(const uint8_t *d_valid, const uint8_t *d_flags, const double *d_fx, const double *d_fy,
                               const double *d_world, int32_t n_features, int32_t row_stride, int32_t n_cams,
                               int32_t *d_cam_idx, int32_t *d_pt_idx, double *d_obs, double *d_pts,
                               int32_t *d_valid_rows, int32_t *d_totals2, void *d_ws, size_t ws_bytes, void *stream) {
    if (!d_valid || !d_flags || !d_fx || !d_fy || !d_world || !d_cam_idx || !d_pt_idx || !d_obs || !d_pts ||
        !d_valid_rows || !d_totals2 || !d_ws || n_features < 1 || n_cams < 0 || row_stride < n_cams)
        return SFM_E_INVALID;
    if (ws_bytes < sfm_ba_assemble_workspace_bytes(n_features)) return SFM_E_WORKSPACE;
    cudaStream_t st = as_stream(stream);
    const size_t seg = align_up(sizeof(int32_t) * (size_t)n_features, 256);
    unsigned char *p = reinterpret_cast<unsigned char *>(d_ws);
    int32_t *rowcnt = reinterpret_cast<int32_t *>(p), *rowvalid = reinterpret_cast<int32_t *>(p + seg);
    int32_t *rowoff = reinterpret_cast<int32_t *>(p + 2 * seg), *ptlocal = reinterpret_cast<int32_t *>(p + 3 * seg);
    void *tmp = p + 4 * seg;
    size_t tmp_bytes = asm_scan_bytes(n_features);
    const int grid = ceil_div(n_features, 256);
    asm_count_kernel<<<grid, 256, 0, st>>>(d_valid, d_flags, n_features, row_stride, n_cams, rowcnt, rowvalid);
    SFM_LAUNCH_CHECK();
    SFM_CUDA_TRY(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, rowcnt, rowoff, n_features, st));
    count_launch();
    SFM_CUDA_TRY(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, rowvalid, ptlocal, n_features, st));
    count_launch();
    asm_scatter_kernel<<<grid, 256, 0, st>>>(d_flags, d_fx, d_fy, d_world, n_features, row_stride, n_cams, rowcnt,
                                            rowvalid, rowoff, ptlocal, d_cam_idx, d_pt_idx, d_obs, d_pts,
                                            d_valid_rows, d_totals2);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}
