// FP32 issue-rate micro-benchmarks on sm_100a (B200): scalar FFMA against the packed FFMA2
// (fma.rn.f32x2), alone and mixed with MUFU.SQRT / FSETP, at full occupancy (8 warps per SMSP).
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
__device__ __forceinline__ float sqrt_approx(float s) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));
    return r;
}

constexpr int ITER = 4096;

// mode 0: 8 independent scalar FFMA chains, all-register operands
// mode 1: 4 independent FFMA2 chains (same flops as mode 0)
// mode 2: mode 0 + one MUFU.SQRT per 12 FFMA
// mode 3: mode 1 + one MUFU.SQRT per 6 FFMA2
// mode 4: mode 0 + one FSETP/IADD pair per 8 FFMA
// mode 5: mode 1 + one FSETP/IADD pair per 4 FFMA2
// mode 6: 2 FFMA2 chains + 4 scalar FFMA chains interleaved (same flops as mode 0): do the packed
//         and the scalar instructions share one pipe or run side by side?
// mode 7: 4 FFMA2 chains whose multiplier/addend have equal halves (ptxas emits the .F32
//         broadcast operand form, as in f_score_packed_kernel)
// mode 8: 3 FFMA2 chains + 2 scalar FFMA chains (same flops as mode 0)
template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, float seed) {
    float x[8];
    float b = seed * 1.0000001f, c = seed * 0.5f, d = seed * 0.25f;
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = seed + threadIdx.x * 1e-3f + i;
    float2 y[4], b2 = make_float2(b, d), c2 = make_float2(c, b);
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] = make_float2(x[2 * i], x[2 * i + 1]);
    float m = seed + 2.f;
    int cnt = 0;
    for (int it = 0; it < ITER; ++it) {
        if (MODE == 6) {
#pragma unroll
            for (int r = 0; r < 3; ++r) {
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    y[i] = ffma2(y[i], (i & 1) ? b2 : c2, (i & 1) ? c2 : b2);
                    x[4 + 2 * i] = fmaf(x[4 + 2 * i], b, c);
                    x[5 + 2 * i] = fmaf(x[5 + 2 * i], d, b);
                }
            }
        } else if (MODE == 8) {
#pragma unroll
            for (int r = 0; r < 3; ++r) {
#pragma unroll
                for (int i = 0; i < 3; ++i) y[i] = ffma2(y[i], (i & 1) ? b2 : c2, (i & 1) ? c2 : b2);
                x[6] = fmaf(x[6], b, c);
                x[7] = fmaf(x[7], d, b);
            }
        } else if (MODE == 7) {
            float2 bb = make_float2(b, b), cc2 = make_float2(c, c);
#pragma unroll
            for (int r = 0; r < 3; ++r) {
#pragma unroll
                for (int i = 0; i < 4; ++i) y[i] = ffma2(y[i], bb, cc2);
            }
        } else if (MODE == 0 || MODE == 2 || MODE == 4) {
#pragma unroll
            for (int r = 0; r < 3; ++r) {
#pragma unroll
                for (int i = 0; i < 8; ++i) x[i] = fmaf(x[i], (i & 1) ? b : d, (i & 2) ? c : b);
                if (MODE == 4) cnt += (x[r] < c) ? 1 : 0;
            }
            if (MODE == 2) { m = sqrt_approx(m + x[0]); m = sqrt_approx(m + x[4]); }
        } else {
#pragma unroll
            for (int r = 0; r < 3; ++r) {
#pragma unroll
                for (int i = 0; i < 4; ++i) y[i] = ffma2(y[i], (i & 1) ? b2 : c2, (i & 2) ? c2 : b2);
                if (MODE == 5) cnt += (y[r].x < c) ? 1 : 0;
            }
            if (MODE == 3) { m = sqrt_approx(m + y[0].x); m = sqrt_approx(m + y[2].y); }
        }
    }
    float s = m + cnt;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += x[i];
#pragma unroll
    for (int i = 0; i < 4; ++i) s += y[i].x + y[i].y;
    if (s == 123.456f) out[0] = s;
}

template <int MODE>
void run(const char *name, float *d) {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    dim3 grid(sms * 4), block(256);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    k<MODE><<<grid, block>>>(d, 1.0f);
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    k<MODE><<<grid, block>>>(d, 1.0f);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    double fma_lane = (double)grid.x * 256 * ITER * 24;  // scalar-FMA equivalents
    printf("%-44s %8.3f ms  %7.2f T scalar-FMA/s  (%5.1f FMA/clk/SM at 1.965 GHz)\n", name, ms,
           fma_lane / (ms * 1e-3) / 1e12, fma_lane / (ms * 1e-3) / sms / 1.965e9);
}

int main() {
    float *d;
    cudaMalloc(&d, 1024);
    run<0>("FFMA x24", d);
    run<1>("FFMA2 x12", d);
    run<2>("FFMA x24 + 2 MUFU.SQRT", d);
    run<3>("FFMA2 x12 + 2 MUFU.SQRT", d);
    run<4>("FFMA x24 + 3 (FSETP+IADD)", d);
    run<5>("FFMA2 x12 + 3 (FSETP+IADD)", d);
    run<6>("FFMA2 x6 + FFMA x12 interleaved", d);
    run<8>("FFMA2 x9 + FFMA x6 interleaved", d);
    run<7>("FFMA2 x12, broadcast .F32 operands", d);
    return 0;
}
