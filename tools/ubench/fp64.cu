// FP64 peak micro-benchmarks on sm_100a (B200): DFMA from the CUDA cores and DMMA (mma.sync.m8n8k4.f64, the FP64
// tensor path K11 / K11s use), both from registers only, at 8 / 16 / 32 warps per SM. The denominators for K11's
// "TFLOP/s" ("FP64 peak is not in MEASURED_PEAKS -- measure a DFMA micro-benchmark on the box").
// Diagnostic only; results recorded under profiles/.   nvcc -arch=sm_100a -O3 -o fp64 fp64.cu && ./fp64
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITER = 2048;

// CHAINS independent DFMA chains per thread
template <int CHAINS>
__global__ void dfma_kernel(double *out, double seed) {
    double x[CHAINS];
    const double b = seed * 1.0000001, c = seed * 0.5;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) x[i] = seed + threadIdx.x * 1e-3 + i;
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) x[i] = fma(x[i], b, c);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += x[i];
    if (s == 12345.678) out[0] = s;
}

__device__ __forceinline__ void dmma_884(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// TILES independent 8x8 accumulator tiles per warp
template <int TILES>
__global__ void dmma_kernel(double *out, double seed) {
    double acc[TILES][2];
    const double a = seed * 1.0000001 + (threadIdx.x & 31) * 1e-6, b = seed * 0.5;
#pragma unroll
    for (int i = 0; i < TILES; ++i) { acc[i][0] = seed + i; acc[i][1] = seed - i; }
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < TILES; ++i) dmma_884(acc[i][0], acc[i][1], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < TILES; ++i) s += acc[i][0] + acc[i][1];
    if (s == 12345.678) out[0] = s;
}

template <class F>
static float time_ms(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        best = ms < best ? ms : best;
    }
    return best;
}

int main() {
    int dev = 0, sms = 0, khz = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    double *out; cudaMalloc(&out, 64);
    printf("device: %d SMs, max SM clock %.0f MHz\n", sms, khz / 1e3);
    const int warps_per_sm[3] = {8, 16, 32};
    for (int w = 0; w < 3; ++w) {
        const int threads = 256, blocks = sms * warps_per_sm[w] * 32 / threads;
        {
            float ms = time_ms([&] { dfma_kernel<8><<<blocks, threads>>>(out, 1.0); });
            double fma = (double)blocks * threads * ITER * 8;
            printf("DFMA x8 chains, %2d warps/SM   %8.3f ms  %7.2f TFLOP/s  (%5.1f DFMA/clk/SM at max clock)\n", warps_per_sm[w], ms,
                   2 * fma / ms / 1e9, fma / (ms * 1e-3) / sms / (khz * 1e3));
        }
        {
            float ms = time_ms([&] { dmma_kernel<8><<<blocks, threads>>>(out, 1.0); });
            double mma = (double)blocks * (threads / 32) * ITER * 8;           // warp-level m8n8k4: 256 FMA each
            printf("DMMA x8 tiles,  %2d warps/SM   %8.3f ms  %7.2f TFLOP/s  (%5.1f FMA/clk/SM at max clock)\n", warps_per_sm[w], ms,
                   2 * mma * 256 / ms / 1e9, mma * 256 / (ms * 1e-3) / sms / (khz * 1e3));
        }
    }
    {
        const int threads = 256, blocks = sms * 4;
        float ms = time_ms([&] { dfma_kernel<1><<<blocks, threads>>>(out, 1.0); });
        printf("DFMA dependent chain latency: %.1f cycles at max clock\n", ms * 1e-3 * khz * 1e3 / ITER);
        ms = time_ms([&] { dmma_kernel<1><<<blocks, threads>>>(out, 1.0); });
        printf("DMMA dependent chain latency: %.1f cycles at max clock\n", ms * 1e-3 * khz * 1e3 / ITER);
    }
    cudaFree(out);
    return 0;
}
