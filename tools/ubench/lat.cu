// Latency / throughput micro-benchmarks for the FP64 pipe, shuffles, shared memory and barriers on
// sm_100a (B200). Diagnostic only; results recorded in profiles/r01_ubench_b200.txt.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_lat(double *out, long long *cyc, int nwarps_active) {
    __shared__ double sm[1024];
    int tid = threadIdx.x;
    sm[tid & 1023] = tid * 1e-3;
    __syncthreads();
    double a = out[0] + tid, b = 1.0000001, c = 0.5;
    long long t0, t1;
    // 1) dependent DFMA chain
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 256; ++i) a = fma(a, b, c);
    t1 = clock64();
    if (tid == 0) cyc[0] = (t1 - t0);
    // 2) dependent DADD chain
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 256; ++i) a = a + c;
    t1 = clock64();
    if (tid == 0) cyc[1] = (t1 - t0);
    // 3) dependent DMUL chain
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 256; ++i) a = a * b;
    t1 = clock64();
    if (tid == 0) cyc[2] = (t1 - t0);
    // 4) 8 independent DFMA chains (throughput per warp)
    double x[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = a + k;
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 64; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k) x[k] = fma(x[k], b, c);
    t1 = clock64();
    if (tid == 0) cyc[3] = (t1 - t0);
#pragma unroll
    for (int k = 0; k < 8; ++k) a += x[k];
    // 5) dependent f64 shuffle + add chain
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 64; ++i) a += __shfl_xor_sync(0xffffffffu, a, 1);
    t1 = clock64();
    if (tid == 0) cyc[4] = (t1 - t0);
    // 6) dependent LDS.64 chain (pointer chase through values)
    int idx = tid & 1023;
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 64; ++i) { double v = sm[idx]; idx = ((int)(v * 0.0) + idx + 33) & 1023; a += v; }
    t1 = clock64();
    if (tid == 0) cyc[5] = (t1 - t0);
    // 7) __syncthreads chain
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 64; ++i) __syncthreads();
    t1 = clock64();
    if (tid == 0) cyc[6] = (t1 - t0);
    // 8) STS -> barrier -> LDS (another thread's value) -> DADD chain
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 64; ++i) { sm[tid] = a; __syncthreads(); a += sm[(tid + 32) % blockDim.x]; __syncthreads(); }
    t1 = clock64();
    if (tid == 0) cyc[7] = (t1 - t0);
    // 9) rsqrt seed + 2 newton
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 64; ++i) {
        double y; asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
        double h = 0.5 * a; y = y * fma(-h * y, y, 1.5); y = y * fma(-h * y, y, 1.5); a = a + y;
    }
    t1 = clock64();
    if (tid == 0) cyc[8] = (t1 - t0);
    // 10) sqrt + div
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 64; ++i) { a = sqrt(a) ; a = 1.0 / a + 2.0; }
    t1 = clock64();
    if (tid == 0) cyc[9] = (t1 - t0);
    // 11) MUFU.RCP64H + DADD
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 64; ++i) { double y; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a)); a = a + y; }
    t1 = clock64();
    if (tid == 0) cyc[10] = (t1 - t0);
    // 12) rcp seed + cubic correction (3 dependent DFMA) + DADD
    t0 = clock64();
#pragma unroll
    for (int i = 0; i < 64; ++i) {
        double y0; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(a));
        double e = fma(-a, y0, 1.0), q = fma(e, e, e); a = a + fma(-y0, q, -y0);
    }
    t1 = clock64();
    if (tid == 0) cyc[11] = (t1 - t0);
    // 13) dependent LDS.128 chain
    {
        const double2 *s2 = reinterpret_cast<const double2 *>(sm);
        int j = tid & 511;
        t0 = clock64();
#pragma unroll
        for (int i = 0; i < 64; ++i) { double2 v = s2[j]; j = ((int)(v.x * 0.0) + j + 17) & 511; a += v.y; }
        t1 = clock64();
        if (tid == 0) cyc[12] = (t1 - t0);
    }
    // 14) one warp: STS.128 x2 -> barrier -> everyone LDS.128 -> DADD (no second barrier: two buffers)
    {
        double2 *s2 = reinterpret_cast<double2 *>(sm);
        t0 = clock64();
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            if (tid < 32) { s2[(i & 1) * 256 + 2 * tid] = make_double2(a, a); s2[(i & 1) * 256 + 2 * tid + 1] = make_double2(a, a); }
            __syncthreads();
            double2 v = s2[(i & 1) * 256 + (tid & 31)];
            a += v.x;
        }
        t1 = clock64();
        if (tid == 0) cyc[13] = (t1 - t0);
    }
    // 15) same with a single STS.64 by one warp
    {
        t0 = clock64();
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            if (tid < 32) sm[(i & 1) * 512 + tid] = a;
            __syncthreads();
            a += sm[(i & 1) * 512 + (tid & 31)];
        }
        t1 = clock64();
        if (tid == 0) cyc[14] = (t1 - t0);
    }
    // 16) barrier -> LDS -> DADD only (no store)
    {
        t0 = clock64();
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            __syncthreads();
            a += sm[(i & 1) * 512 + (tid & 31) + (a > 1e300 ? 1 : 0)];
        }
        t1 = clock64();
        if (tid == 0) cyc[15] = (t1 - t0);
    }
    // 17..20) store-shape variants of 14: one warp stores, barrier, everyone loads
    {
        double2 *s2 = reinterpret_cast<double2 *>(sm);
        t0 = clock64();
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            if (tid < 32) s2[(i & 1) * 256 + tid] = make_double2(a, a);
            __syncthreads();
            a += s2[(i & 1) * 256 + (tid & 31)].x;
        }
        t1 = clock64();
        if (tid == 0) cyc[16] = (t1 - t0);
        t0 = clock64();
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            if (tid < 32) { sm[(i & 1) * 512 + tid] = a; sm[(i & 1) * 512 + 32 + tid] = a; }
            __syncthreads();
            a += sm[(i & 1) * 512 + (tid & 31)];
        }
        t1 = clock64();
        if (tid == 0) cyc[17] = (t1 - t0);
        t0 = clock64();
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            if (tid < 32) { sm[(i & 1) * 512 + tid] = a; sm[(i & 1) * 512 + 32 + tid] = a; sm[(i & 1) * 512 + 64 + tid] = a; sm[(i & 1) * 512 + 96 + tid] = a; }
            __syncthreads();
            a += sm[(i & 1) * 512 + (tid & 31)];
        }
        t1 = clock64();
        if (tid == 0) cyc[18] = (t1 - t0);
        // 4 lanes of each of the first 6 warps store one STS.128 pair (the factor's publish shape)
        t0 = clock64();
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            if ((tid & 7) == 0 && tid < 192) { s2[(i & 1) * 256 + (tid >> 3) * 2] = make_double2(a, a); s2[(i & 1) * 256 + (tid >> 3) * 2 + 1] = make_double2(a, a); }
            __syncthreads();
            a += s2[(i & 1) * 256 + (tid & 31)].x;
        }
        t1 = clock64();
        if (tid == 0) cyc[19] = (t1 - t0);
    }
    out[tid] = a;
}
int main() {
    double *out; long long *cyc, h[32];
    cudaMalloc(&out, 8 * 2048); cudaMemset(out, 0, 8 * 2048); cudaMalloc(&cyc, 8 * 32);
    const char *names[] = {"dep DFMA", "dep DADD", "dep DMUL", "8x indep DFMA (per instr)", "dep SHFL64+DADD", "dep LDS64+DADD", "__syncthreads", "STS/bar/LDS/DADD/bar", "rsqrt+2 newton+DADD", "sqrt+div+DADD", "MUFU.RCP64H+DADD", "rcp+cubic+DADD", "dep LDS128+DADD", "STS128x2(1 warp)/bar/LDS128/DADD", "STS64(1 warp)/bar/LDS64/DADD", "bar/LDS64/DADD", "STS128x1/bar/LDS128/DADD", "STS64x2/bar/LDS64/DADD", "STS64x4/bar/LDS64/DADD", "STS128x2 (4 lanes x 6 warps)/bar/LDS128"};
    const int cnt[] = {256, 256, 256, 512, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64};
    for (int threads : {32, 192, 256}) {
        k_lat<<<1, threads>>>(out, cyc, 0); cudaDeviceSynchronize();
        k_lat<<<1, threads>>>(out, cyc, 0); cudaDeviceSynchronize();
        cudaMemcpy(h, cyc, 8 * 32, cudaMemcpyDeviceToHost);
        printf("threads per CTA = %d (1 CTA)\n", threads);
        for (int i = 0; i < 20; ++i) printf("  %-28s %8.1f cycles each\n", names[i], (double)h[i] / cnt[i]);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
