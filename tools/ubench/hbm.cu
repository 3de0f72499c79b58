// HBM stream ceilings on B200: read-only, write-only, copy, and the K9p write shape (18 SoA planes).
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_read(const double2 *a, size_t n, double *out) {
    double s = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) { double2 v = a[i]; s += v.x + v.y; }
    if (s == 1.2345) out[0] = s;
}
__global__ void k_write(double2 *a, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) a[i] = make_double2(1.0, 2.0);
}
__global__ void k_copy(const double2 *a, double2 *b, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) b[i] = a[i];
}
template <bool CS>
__global__ void k_planes(double *w, size_t nobs) {
    for (size_t o = blockIdx.x * (size_t)blockDim.x + threadIdx.x; o < nobs; o += (size_t)gridDim.x * blockDim.x)
#pragma unroll
        for (int r = 0; r < 18; ++r) { if (CS) __stcs(&w[r * nobs + o], 1.0 + r); else w[r * nobs + o] = 1.0 + r; }
}
int main() {
    const size_t nobs = 4000000, bytes = nobs * 144;
    double *a, *b, *out;
    cudaMalloc(&a, bytes); cudaMalloc(&b, bytes); cudaMalloc(&out, 8);
    cudaMemset(a, 0, bytes); cudaMemset(b, 0, bytes);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const size_t n2 = bytes / 16;
    for (int grid : {148 * 2, 148 * 4, 148 * 8, 148 * 16}) for (int threads : {256, 768}) {
        float ms[5];
        for (int which = 0; which < 5; ++which) {
            for (int rep = 0; rep < 3; ++rep) {
                cudaEventRecord(e0);
                if (which == 0) k_read<<<grid, threads>>>((double2 *)a, n2, out);
                if (which == 1) k_write<<<grid, threads>>>((double2 *)a, n2);
                if (which == 2) k_copy<<<grid, threads>>>((double2 *)a, (double2 *)b, n2);
                if (which == 3) k_planes<false><<<grid, threads>>>(a, nobs);
                if (which == 4) k_planes<true><<<grid, threads>>>(a, nobs);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                cudaEventElapsedTime(&ms[which], e0, e1);
            }
        }
        printf("grid %5d x %3d: read %6.0f GB/s  write %6.0f GB/s  copy %6.0f GB/s (r+w)  18-plane write %6.0f GB/s  (.cs) %6.0f GB/s\n", grid, threads,
               bytes / ms[0] / 1e6, bytes / ms[1] / 1e6, 2.0 * bytes / ms[2] / 1e6, bytes / ms[3] / 1e6, bytes / ms[4] / 1e6);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
