// MUFU issue-rate micro-benchmarks on sm_100a (B200): which special-function ops cost FMA-pipe
// issue cycles when mixed with packed FP32 work (the shape of K2's inner loop)?
// Diagnostic only; results recorded under profiles/.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a);
    unsigned long long rb = *reinterpret_cast<unsigned long long *>(&b);
    unsigned long long rc = *reinterpret_cast<unsigned long long *>(&c);
    unsigned long long rd;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2 *>(&rd);
}
template <int OP>
__device__ __forceinline__ float mufu(float s) {
    float r;
    if (OP == 0) asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));
    if (OP == 1) asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));
    if (OP == 2) asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));
    if (OP == 3) asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));
    if (OP == 4) asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));
    return r;
}

constexpr int ITER = 4096;

// NF packed FMAs (broadcast-operand form, full-rate) + NM independent MUFU ops per iteration
template <int OP, int NF, int NM>
__global__ void __launch_bounds__(256) k(float *out, float seed) {
    float b = seed * 1.0000001f, c = seed * 0.5f;
    float2 y[4], bb = make_float2(b, b), cc = make_float2(c, c);
    float m[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        y[i] = make_float2(seed + threadIdx.x * 1e-3f + i, seed + i);
        m[i] = seed + 2.f + i;
    }
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < NF; ++i) y[i & 3] = ffma2(y[i & 3], bb, cc);
#pragma unroll
        for (int i = 0; i < NM; ++i) m[i & 3] = mufu<OP>(m[i & 3]);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) s += y[i].x + y[i].y + m[i];
    if (s == 123.456f) out[0] = s;
}

template <int OP, int NF, int NM>
void run(const char *name, float *d) {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    dim3 grid(sms * 4), block(256);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    k<OP, NF, NM><<<grid, block>>>(d, 1.0f);
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    k<OP, NF, NM><<<grid, block>>>(d, 1.0f);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    // cycles per iteration per SMSP-resident warp slot: 8 warps per SMSP share the issue port
    double cyc = ms * 1e-3 * 1.965e9 / ITER / 8.0;
    printf("%-40s %8.3f ms  %6.2f issue cycles per iteration per warp\n", name, ms, cyc);
}

int main() {
    float *d;
    cudaMalloc(&d, 1024);
    run<0, 12, 0>("12 FFMA2", d);
    run<0, 0, 4>("4 MUFU.SQRT", d);
    run<1, 0, 4>("4 MUFU.RSQ", d);
    run<2, 0, 4>("4 MUFU.RCP", d);
    run<3, 0, 4>("4 MUFU.EX2", d);
    run<4, 0, 4>("4 MUFU.LG2", d);
    run<0, 12, 2>("12 FFMA2 + 2 MUFU.SQRT", d);
    run<1, 12, 2>("12 FFMA2 + 2 MUFU.RSQ", d);
    run<2, 12, 2>("12 FFMA2 + 2 MUFU.RCP", d);
    run<3, 12, 2>("12 FFMA2 + 2 MUFU.EX2", d);
    run<0, 12, 1>("12 FFMA2 + 1 MUFU.SQRT", d);
    run<1, 12, 1>("12 FFMA2 + 1 MUFU.RSQ", d);
    run<0, 12, 4>("12 FFMA2 + 4 MUFU.SQRT", d);
    run<1, 12, 4>("12 FFMA2 + 4 MUFU.RSQ", d);
    return 0;
}
