// Throughput of block-scatter accumulation into a large FP64 matrix on sm_100a (B200):
//   mode 0: 36 scalar RED.ADD.F64 per 6x6 block (row-major S, 6 rows of 48 B)      [K10 today]
//   mode 1: 6 x cp.reduce.async.bulk add.f64 of 48 B per block (row-major S)
//   mode 2: 1 x cp.reduce.async.bulk add.f64 of 288 B per block (block-contiguous S)
// Blocks are drawn pseudo-randomly from an nb x nb block grid (lower triangle), like the Schur
// complement's camera pairs. Diagnostic only.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_red_f64(double *g, const double *s, uint32_t bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;"
                 :: "l"(g), "r"(smem_u32(s)), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t lcg(uint32_t &s) { s = s * 1664525u + 1013904223u; return s >> 8; }

template <int MODE>
__global__ void __launch_bounds__(256) k(double *S, size_t ld, int nb, int iters, int band) {
    __shared__ __align__(16) double stage[8][2][36];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t seed = (blockIdx.x * 8 + warp) * 2654435761u + 12345u;
    for (int it = 0; it < iters; ++it) {
        int ci = lcg(seed) % nb, cj = band ? max(0, ci - (int)(lcg(seed) % band)) : (int)(lcg(seed) % (ci + 1));
        double val = 1e-9 * (lane + it);
        if (MODE == 0) {
            for (int e = lane; e < 36; e += 32) {
                int r = e / 6, q = e - 6 * r;
                atomicAdd(&S[(size_t)(6 * ci + r) * ld + 6 * cj + q], val);
            }
        } else {
            double *st = stage[warp][it & 1];
            if ((it & 1) == 0 && it >= 2) {       // both stages were handed to the bulk engine: wait until read
                if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                __syncwarp();
            }
            for (int e = lane; e < 36; e += 32) st[e] = val;
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                if (MODE == 1) {
                    for (int r = 0; r < 6; ++r) bulk_red_f64(&S[(size_t)(6 * ci + r) * ld + 6 * cj], st + 6 * r, 48);
                } else {
                    bulk_red_f64(&S[((size_t)ci * nb + cj) * 36], st, 288);
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            __syncwarp();
        }
    }
    if (MODE != 0 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

int main() {
    const int nb = 750, iters = 2000;
    const size_t ld = 6 * nb;
    double *S;
    cudaMalloc(&S, ld * ld * 8);
    cudaMemset(S, 0, ld * ld * 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int band : {0, 40}) {
        for (int mode = 0; mode < 3; ++mode) {
            for (int grid : {148, 296, 592}) {
                float ms = 0;
                for (int rep = 0; rep < 2; ++rep) {
                    cudaEventRecord(e0);
                    if (mode == 0) k<0><<<grid, 256>>>(S, ld, nb, iters, band);
                    if (mode == 1) k<1><<<grid, 256>>>(S, ld, nb, iters, band);
                    if (mode == 2) k<2><<<grid, 256>>>(S, ld, nb, iters, band);
                    cudaEventRecord(e1);
                    cudaEventSynchronize(e1);
                    cudaEventElapsedTime(&ms, e0, e1);
                }
                double blocks = (double)grid * 8 * iters;
                printf("band %3d mode %d grid %4d: %8.3f ms  %7.2f M blocks/s  (%6.1f G f64 adds/s)  %s\n", band, mode, grid, ms,
                       blocks / ms / 1e3, blocks * 36 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
            }
        }
    }
    // correctness spot check of mode 2 vs expectation: sum of S must equal sum of all values added in the last run
    return 0;
}
