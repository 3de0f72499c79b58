// f_ransac.cu -- K1..K4 of the F-matrix RANSAC path for sm_100a.
//
//   K1 f8pt_kernel        EstimateFundamentalMatrix.py:30-78 for every sampled 8-tuple
//   K2 f_score_packed_kernel GetInliersRANSAC.py:67-82: inlier count of every hypothesis
//      (f_score_kernel = the all-FP64 form it must reproduce; f_score_mixed_kernel = the scalar
//       FP32 filter it replaced, kept behind SFM_F_SCORE_SCALAR32 for comparison)
//   K3 argmax_first_kernel GetInliersRANSAC.py:85: first hypothesis with the maximum count
//   K4 f_mask_kernel      GetInliersRANSAC.py:81,87 for the winner
//
// Layout: correspondences [n,4] f64 rows (x1,y1,x2,y2) = two 16-byte vector loads per row;
// samples [h,8] i32; hypotheses [h,9] f64; counts [h] i32.
//
// K2 is the hot kernel: h x n point-hypothesis tests. One thread owns one hypothesis (its F
// stays in registers), a block stages a chunk of correspondences in shared memory and every
// warp reads them back as broadcasts, so the inner loop is pure arithmetic-pipe work: 31 FP64
// instructions per test in f_score_kernel; in the shipped kernel a certified FP32 filter on the
// packed FP32 pipe (11.5 FFMA2-class + 2 MUFU per test) with an FP64 redo of the ~2 tests in
// 10^4 it cannot decide, which gives the FP64 kernel's counts bit for bit. The grid is
// (hypothesis tiles) x (point chunks); partial counts are merged with integer atomics, so the
// result is order-independent and bit-reproducible.
#include "f_math.cuh"

namespace sfm {

// ------------------------------------------------------------------------------------ K1
__global__ void __launch_bounds__(128)
f8pt_kernel(const double *__restrict__ x1x2, const int32_t *__restrict__ sample, int h,
            double *__restrict__ Fout, const unsigned long long *__restrict__ key = nullptr,
            int32_t *__restrict__ best = nullptr) {
    int hyp = blockIdx.x * blockDim.x + threadIdx.x;
    double *out = Fout + (size_t)hyp * 9;
    if (key) {
        // multi-GPU tail (sfm_f_winner_from_key): the ONE hypothesis named by the all-reduced (count, ~index) key, F to
        // Fout[0..9). The same kernel as the batch on purpose: the winner's F then has the bits a single GPU scoring
        // the whole stream stores for it, whatever ptxas contracts.
        if (hyp != 0) return;
        const unsigned long long k = *key;
        const int idx = (int)(0xFFFFFFFFu - (unsigned int)(k & 0xFFFFFFFFull)), cnt = (int)(k >> 32);
        best[0] = idx;
        best[1] = cnt;
        if (cnt <= 0 || idx < 0 || idx >= h) {
            best[1] = 0;
            for (int i = 0; i < 9; ++i) Fout[i] = 0.0;
            return;
        }
        hyp = idx;
        out = Fout;
    } else if (hyp >= h) {
        return;
    }
    double px1[8], py1[8], px2[8], py2[8];
    const int4 *sp = reinterpret_cast<const int4 *>(sample + (size_t)hyp * 8);
    int4 s0 = __ldg(sp), s1 = __ldg(sp + 1);
    int idx[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const double2 *row = reinterpret_cast<const double2 *>(x1x2 + (size_t)idx[i] * 4);
        double2 a = __ldg(row), b = __ldg(row + 1);
        px1[i] = a.x; py1[i] = a.y; px2[i] = b.x; py2[i] = b.y;
    }
    double F[9];
    f8pt_from_sample(px1, py1, px2, py2, F);
#pragma unroll
    for (int i = 0; i < 9; ++i) out[i] = F[i];
}

// ------------------------------------------------------------------------------------ K2
constexpr int K2_THREADS = 256;

__global__ void __launch_bounds__(K2_THREADS)
f_score_kernel(const double *__restrict__ x1x2, int n, const double *__restrict__ Fall, int h,
               double two_thr, int chunk, int32_t *__restrict__ counts) {
    extern __shared__ double2 spts[];  // chunk rows, two double2 per row
    const int p0 = blockIdx.y * chunk;
    const int np = min(chunk, n - p0);
    {
        const double2 *src = reinterpret_cast<const double2 *>(x1x2) + (size_t)p0 * 2;
        for (int i = threadIdx.x; i < np * 2; i += K2_THREADS) spts[i] = __ldg(src + i);
    }
    const int hyp = blockIdx.x * K2_THREADS + threadIdx.x;
    FMat F;
    {
        // out-of-range threads score hypothesis h-1 again and drop the result: keeps the
        // warp convergent around __syncthreads and the loop branch-free.
        const double *f = Fall + (size_t)min(hyp, h - 1) * 9;
        F.f0 = __ldg(f + 0); F.f1 = __ldg(f + 1); F.f2 = __ldg(f + 2);
        F.f3 = __ldg(f + 3); F.f4 = __ldg(f + 4); F.f5 = __ldg(f + 5);
        F.f6 = __ldg(f + 6); F.f7 = __ldg(f + 7); F.f8 = __ldg(f + 8);
    }
    __syncthreads();
    int cnt = 0;
#pragma unroll 4
    for (int i = 0; i < np; ++i) {
        double2 a = spts[2 * i], b = spts[2 * i + 1];
        cnt += f_is_inlier(F, a.x, a.y, b.x, b.y, two_thr) ? 1 : 0;
    }
    if (hyp < h && cnt) atomicAdd(&counts[hyp], cnt);
}

// K2 with the certified FP32 filter (f_math.cuh: f_cert_margin). Same thread/hypothesis x
// staged-chunk decomposition; shared memory holds the chunk rounded to FP32 (16 B per row). A
// test whose FP32 margin does not clear its error bound is redone in FP64 from global memory
// (cold path), so the counts are those of f_score_kernel bit for bit. Per test: 1 LDS.128,
// 27 FP32-pipe + 2 MUFU instructions and 4 predicate/integer ones instead of 31 FP64.
constexpr int K2M_GROUP = 4;
__device__ unsigned long long g_k2_fp64_tests = 0;  // diagnostic: tests that took the cold path

template <int G>
__device__ __forceinline__ void score_group(const FCert &c, const float4 *__restrict__ rows,
                                            const double *__restrict__ x1x2_rows,
                                            const double *__restrict__ Frow, double two_thr,
                                            int &cnt) {
    unsigned unc = 0;
#pragma unroll
    for (int k = 0; k < G; ++k) {
        float4 r = rows[k];
        float tol;
        float d = f_cert_margin(c, r.x, r.y, r.z, r.w, tol);
        cnt += (d < -tol) ? 1 : 0;
        unc |= (fabsf(d) > tol) ? 0u : (1u << k);
    }
    if (__any_sync(0xffffffffu, unc != 0)) {
        if (unc) {
            // opaque pointer: keeps ptxas from merging these loads with the prologue's and then
            // holding F in FP64 registers across the hot loop
            const double *Fr = Frow;
            asm volatile("" : "+l"(Fr));
            FMat F;
            F.f0 = __ldcg(Fr + 0); F.f1 = __ldcg(Fr + 1); F.f2 = __ldcg(Fr + 2);
            F.f3 = __ldcg(Fr + 3); F.f4 = __ldcg(Fr + 4); F.f5 = __ldcg(Fr + 5);
            F.f6 = __ldcg(Fr + 6); F.f7 = __ldcg(Fr + 7); F.f8 = __ldcg(Fr + 8);
            int redo = 0;
#pragma unroll
            for (int k = 0; k < G; ++k) {
                if (unc & (1u << k)) {
                    const double2 *row = reinterpret_cast<const double2 *>(x1x2_rows) + 2 * k;
                    double2 a = __ldg(row), b = __ldg(row + 1);
                    cnt += f_is_inlier(F, a.x, a.y, b.x, b.y, two_thr) ? 1 : 0;
                    ++redo;
                }
            }
            atomicAdd(&g_k2_fp64_tests, (unsigned long long)redo);
        }
    }
}

__global__ void __launch_bounds__(K2_THREADS)
f_score_mixed_kernel(const double *__restrict__ x1x2, int n, const double *__restrict__ Fall, int h,
                     double two_thr, int chunk, int32_t *__restrict__ counts) {
    extern __shared__ float4 spts32[];  // chunk rows (x1,y1,x2,y2) rounded to FP32
    __shared__ float s_m1[K2_THREADS / 32], s_m2[K2_THREADS / 32];
    const int p0 = blockIdx.y * chunk;
    const int np = min(chunk, n - p0);
    {
        const double2 *src = reinterpret_cast<const double2 *>(x1x2) + (size_t)p0 * 2;
        float m1 = 0.f, m2 = 0.f;
        for (int i = threadIdx.x; i < np; i += K2_THREADS) {
            double2 a = __ldg(src + 2 * i), b = __ldg(src + 2 * i + 1);
            float4 r = make_float4((float)a.x, (float)a.y, (float)b.x, (float)b.y);
            spts32[i] = r;
            // NaN coordinates: fmaxf drops them here, and they make d NaN -> FP64 path
            m1 = fmaxf(m1, fmaxf(fabsf(r.x), fabsf(r.y)));
            m2 = fmaxf(m2, fmaxf(fabsf(r.z), fabsf(r.w)));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, o));
            m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
        }
        if ((threadIdx.x & 31) == 0) {
            s_m1[threadIdx.x >> 5] = m1;
            s_m2[threadIdx.x >> 5] = m2;
        }
    }
    const int hyp = blockIdx.x * K2_THREADS + threadIdx.x;
    const double *Frow = Fall + (size_t)min(hyp, h - 1) * 9;
    __syncthreads();
    FCert c;
    {
        float m1 = 0.f, m2 = 0.f;
#pragma unroll
        for (int w = 0; w < K2_THREADS / 32; ++w) {
            m1 = fmaxf(m1, s_m1[w]);
            m2 = fmaxf(m2, s_m2[w]);
        }
        FMat F;
        F.f0 = __ldg(Frow + 0); F.f1 = __ldg(Frow + 1); F.f2 = __ldg(Frow + 2);
        F.f3 = __ldg(Frow + 3); F.f4 = __ldg(Frow + 4); F.f5 = __ldg(Frow + 5);
        F.f6 = __ldg(Frow + 6); F.f7 = __ldg(Frow + 7); F.f8 = __ldg(Frow + 8);
        // the FP32 rounding of a coordinate can exceed the FP64 value by one part in 2^24
        c = f_cert_setup(F, (double)m1 * 1.0000002, (double)m2 * 1.0000002, two_thr);
        // pin the FP32 constants in registers: ptxas otherwise keeps F in FP64 registers and
        // re-converts inside the loop (F2F shares the quarter-rate pipe with the two MUFU.SQRT).
        // z is 0 at run time (chunk > 0) but not at compile time.
        const int z = chunk >> 31;
#define SFM_PIN(x) x = __int_as_float(__float_as_int(x) | z)
        SFM_PIN(c.f0); SFM_PIN(c.f1); SFM_PIN(c.f2); SFM_PIN(c.f3); SFM_PIN(c.f4);
        SFM_PIN(c.f5); SFM_PIN(c.f6); SFM_PIN(c.f7); SFM_PIN(c.f8); SFM_PIN(c.T);
        SFM_PIN(c.DeT); SFM_PIN(c.Dn12); SFM_PIN(c.C0);
#undef SFM_PIN
    }
    int cnt = 0;
    const double *grow = x1x2 + (size_t)p0 * 4;
    int i = 0;
    for (; i + K2M_GROUP <= np; i += K2M_GROUP)
        score_group<K2M_GROUP>(c, spts32 + i, grow + (size_t)i * 4, Frow, two_thr, cnt);
    for (; i < np; ++i) score_group<1>(c, spts32 + i, grow + (size_t)i * 4, Frow, two_thr, cnt);
    if (hyp < h && cnt) atomicAdd(&counts[hyp], cnt);
}

// K2 on Blackwell's packed FP32 pipe (FFMA2 / FMUL2 / FADD2 = fma|mul|add.rn.f32x2, sm_100+): the
// certified filter above for TWO correspondences per instruction. One thread still owns one
// hypothesis; shared memory holds the chunk pair-interleaved, (x1a,x1b,y1a,y1b),(x2a,x2b,y2a,y2b),
// so one LDS.128 lands each operand pair in an aligned 64-bit register pair, and the nine F
// entries and the bound's constants sit in registers duplicated in both halves. Every packed
// operation rounds each half exactly like the scalar fmaf/mul/add of f_cert_margin, so d and tol
// are bit-identical to the scalar filter and the same certification argument holds; only the
// square roots, |e| and the comparisons stay scalar. Per pair of tests: 2 LDS.128, 23 packed
// FP32, 4 MUFU, 2 LOP and ~8 predicate/integer instructions (20 issue slots per test instead
// of 34).
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk2(float lo, float hi) {
    u64 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void upk2(u64 v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
    u64 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b) {
    u64 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ u64 add2(u64 a, u64 b) {
    u64 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

struct FCert2 {
    u64 f0, f1, f2, f3, f4, f5, f6, f7, f8;
    u64 nT;                     // -2 thr
    u64 DeT, Dn12, C0, U65;     // f_cert_margin's bound constants, 65u
    u64 nDe, Ea1, Ea0;          // f_cert_reject's constants
};

template <int PG, bool EARLY = true>  // pairs per uncertainty check
__device__ __forceinline__ void score_pairs(const FCert2 &c, const float4 *__restrict__ rows,
                                            const double *__restrict__ x1x2_rows,
                                            const double *__restrict__ Frow, double two_thr,
                                            int &cnt, int &vote_fail) {
    unsigned unc = 0;
    if (EARLY && PG > 1) {
        // the first-stage certificate (f_cert_reject_stage1_from) for ALL pairs of the group with one vote: the
        // pairs' chains interleave, and the common case (every test of the group a gross outlier: 86 % of the
        // groups at C2) costs 13 packed operations per pair, 4 compares and one vote -- no F^T x2, no second norm,
        // no square root. A failed vote falls through to the per-pair path below, which repeats the stage for
        // its pair (cheaper than keeping the intermediates alive: 64 registers at 4 CTAs per SM is the measured
        // optimum; 3 CTAs 2.96 ms, 5 CTAs 3.26 ms, 4 pairs per vote 3.08 ms, 2 pairs 2.82 ms)
        bool ok = true;
#pragma unroll
        for (int k = 0; k < PG; ++k) {
            float4 A = rows[2 * k], B = rows[2 * k + 1];
            u64 x1 = pk2(A.x, A.y), y1 = pk2(A.z, A.w), x2 = pk2(B.x, B.y), y2 = pk2(B.z, B.w);
            u64 a = fma2(c.f0, x1, fma2(c.f1, y1, c.f2));
            u64 b = fma2(c.f3, x1, fma2(c.f4, y1, c.f5));
            u64 cc = fma2(c.f6, x1, fma2(c.f7, y1, c.f8));
            u64 e = fma2(x2, a, fma2(y2, b, cc));
            u64 q1 = fma2(a, a, mul2(b, b));
            float el, eh, tl, th, t2l, t2h, rl, rh;
            upk2(e, el, eh);
            const u64 t = add2(pk2(fabsf(el), fabsf(eh)), c.nDe), rhs1 = fma2(c.Ea1, q1, c.Ea0);
            upk2(t, tl, th);
            // t |t| instead of (t > 0 && t t > rhs): the right-hand side is positive (Ea0 >= 1e-36) or +inf, so a
            // non-positive t fails the comparison by itself, and for t > 0 the product is the same t t
            const u64 t2 = mul2(t, pk2(fabsf(tl), fabsf(th)));
            upk2(t2, t2l, t2h);
            upk2(rhs1, rl, rh);
            ok = ok && t2l > rl && t2h > rh;
        }
        if (__all_sync(0xffffffffu, ok)) return;
    }
#pragma unroll
    for (int k = 0; k < PG; ++k) {
        float4 A = rows[2 * k], B = rows[2 * k + 1];
        u64 x1 = pk2(A.x, A.y), y1 = pk2(A.z, A.w), x2 = pk2(B.x, B.y), y2 = pk2(B.z, B.w);
        u64 a = fma2(c.f0, x1, fma2(c.f1, y1, c.f2));
        u64 b = fma2(c.f3, x1, fma2(c.f4, y1, c.f5));
        u64 cc = fma2(c.f6, x1, fma2(c.f7, y1, c.f8));
        u64 e = fma2(x2, a, fma2(y2, b, cc));
        u64 q1 = fma2(a, a, mul2(b, b));
        float q1l, q1h, q2l, q2h, el, eh;
        upk2(q1, q1l, q1h);
        upk2(e, el, eh);
        u64 ea = pk2(fabsf(el), fabsf(eh));
        u64 p, q, q2;
        if (EARLY) {
            // certified early reject (f_math.cuh f_cert_reject_from, the same operations on both halves): more than
            // 95 % of the tests of a RANSAC run are gross outliers, |e| >= T min(N1, N2) by a wide margin; when
            // every lane of the warp can certify that for both of its tests, the two square roots, the margin and
            // its error bound are skipped for the pair. Two stages: first against the norm already at hand
            // (f_cert_reject_stage1_from: implies the full certificate, passes ~93 % of the votes at C2, and spares
            // F^T x2 and the second norm), then against min(q1, q2).
            const u64 t = add2(ea, c.nDe), t2 = mul2(t, t);
            float tl, th, t2l, t2h, rl, rh;
            upk2(t, tl, th);
            upk2(t2, t2l, t2h);
            const bool pos = tl > 0.0f && th > 0.0f;
            {
                const u64 rhs1 = fma2(c.Ea1, q1, c.Ea0);
                upk2(rhs1, rl, rh);
                if (__all_sync(0xffffffffu, pos && t2l > rl && t2h > rh)) continue;     // strict: see f_cert_reject_from
            }
            p = fma2(c.f0, x2, fma2(c.f3, y2, c.f6));
            q = fma2(c.f1, x2, fma2(c.f4, y2, c.f7));
            q2 = fma2(p, p, mul2(q, q));
            upk2(q2, q2l, q2h);
            const u64 rhs = fma2(c.Ea1, pk2(fminf(q1l, q2l), fminf(q1h, q2h)), c.Ea0);
            upk2(rhs, rl, rh);
            if (__all_sync(0xffffffffu, pos && t2l > rl && t2h > rh)) continue;
            ++vote_fail;                       // warp-uniform
        } else {
            p = fma2(c.f0, x2, fma2(c.f3, y2, c.f6));
            q = fma2(c.f1, x2, fma2(c.f4, y2, c.f7));
            q2 = fma2(p, p, mul2(q, q));
            upk2(q2, q2l, q2h);
        }
        u64 n1 = pk2(sqrt_approx_f32(q1l), sqrt_approx_f32(q1h));
        u64 n2 = pk2(sqrt_approx_f32(q2l), sqrt_approx_f32(q2h));
        u64 s = add2(n1, n2);
        u64 lhs = mul2(ea, s);
        u64 d = fma2(mul2(c.nT, n1), n2, lhs);
        u64 tol = fma2(c.U65, lhs, c.C0);
        tol = fma2(c.DeT, s, tol);
        tol = fma2(ea, c.Dn12, tol);
        float d0, d1, t0, t1;
        upk2(d, d0, d1);
        upk2(tol, t0, t1);
        cnt += (d0 < -t0) ? 1 : 0;
        cnt += (d1 < -t1) ? 1 : 0;
        unc |= (fabsf(d0) > t0) ? 0u : (1u << (2 * k));
        unc |= (fabsf(d1) > t1) ? 0u : (2u << (2 * k));
    }
    if (__any_sync(0xffffffffu, unc != 0)) {
        if (unc) {
            const double *Fr = Frow;  // opaque, see score_group
            asm volatile("" : "+l"(Fr));
            FMat F;
            F.f0 = __ldcg(Fr + 0); F.f1 = __ldcg(Fr + 1); F.f2 = __ldcg(Fr + 2);
            F.f3 = __ldcg(Fr + 3); F.f4 = __ldcg(Fr + 4); F.f5 = __ldcg(Fr + 5);
            F.f6 = __ldcg(Fr + 6); F.f7 = __ldcg(Fr + 7); F.f8 = __ldcg(Fr + 8);
            int redo = 0;
#pragma unroll
            for (int t = 0; t < 2 * PG; ++t) {
                if (unc & (1u << t)) {
                    const double2 *row = reinterpret_cast<const double2 *>(x1x2_rows) + 2 * t;
                    double2 a = __ldg(row), b = __ldg(row + 1);
                    cnt += f_is_inlier(F, a.x, a.y, b.x, b.y, two_thr) ? 1 : 0;
                    ++redo;
                }
            }
            atomicAdd(&g_k2_fp64_tests, (unsigned long long)redo);
        }
    }
}

// Two pairs per vote and 4 CTAs per SM: measured best of {1,2,4} pairs x {3,4,5,6} CTAs (round 1: 4.76-5.45 ms at
// C2; with the two-stage early reject 2.82 ms against 2.96-3.26 ms for the neighbours; more CTAs per SM spill).
constexpr int K2P_PAIRS = 2;
constexpr int K2P_CTAS = 4;

template <bool EARLY>
__global__ void __launch_bounds__(K2_THREADS, K2P_CTAS)
f_score_packed_kernel(const double *__restrict__ x1x2, int n, const double *__restrict__ Fall, int h,
                      double two_thr, int chunk, int32_t *__restrict__ counts) {
    extern __shared__ float4 spair[];  // chunk/2 pairs x 2 float4, pair-interleaved FP32 rows
    __shared__ float s_m1[K2_THREADS / 32], s_m2[K2_THREADS / 32];
    const int p0 = blockIdx.y * chunk;  // even: chunk is
    const int np = min(chunk, n - p0);
    {
        const double2 *src = reinterpret_cast<const double2 *>(x1x2) + (size_t)p0 * 2;
        float *sf = reinterpret_cast<float *>(spair);
        float m1 = 0.f, m2 = 0.f;
        for (int i = threadIdx.x; i < np; i += K2_THREADS) {
            double2 a = __ldg(src + 2 * i), b = __ldg(src + 2 * i + 1);
            float4 r = make_float4((float)a.x, (float)a.y, (float)b.x, (float)b.y);
            float *dst = sf + (i >> 1) * 8 + (i & 1);
            dst[0] = r.x; dst[2] = r.y; dst[4] = r.z; dst[6] = r.w;
            m1 = fmaxf(m1, fmaxf(fabsf(r.x), fabsf(r.y)));
            m2 = fmaxf(m2, fmaxf(fabsf(r.z), fabsf(r.w)));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, o));
            m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
        }
        if ((threadIdx.x & 31) == 0) {
            s_m1[threadIdx.x >> 5] = m1;
            s_m2[threadIdx.x >> 5] = m2;
        }
    }
    const int hyp = blockIdx.x * K2_THREADS + threadIdx.x;
    const double *Frow = Fall + (size_t)min(hyp, h - 1) * 9;
    __syncthreads();
    FCert c;
    {
        float m1 = 0.f, m2 = 0.f;
#pragma unroll
        for (int w = 0; w < K2_THREADS / 32; ++w) {
            m1 = fmaxf(m1, s_m1[w]);
            m2 = fmaxf(m2, s_m2[w]);
        }
        FMat F;
        F.f0 = __ldg(Frow + 0); F.f1 = __ldg(Frow + 1); F.f2 = __ldg(Frow + 2);
        F.f3 = __ldg(Frow + 3); F.f4 = __ldg(Frow + 4); F.f5 = __ldg(Frow + 5);
        F.f6 = __ldg(Frow + 6); F.f7 = __ldg(Frow + 7); F.f8 = __ldg(Frow + 8);
        c = f_cert_setup(F, (double)m1 * 1.0000002, (double)m2 * 1.0000002, two_thr);
    }
    FCert2 c2;
    {
        // runtime-opaque zero: keeps ptxas from rematerialising the packed constants from the
        // FP64 F inside the loop (see f_score_mixed_kernel)
        const int z = chunk >> 31;
#define SFM_DUP(x) pk2(__int_as_float(__float_as_int(x) | z), __int_as_float(__float_as_int(x) | z))
        c2.f0 = SFM_DUP(c.f0); c2.f1 = SFM_DUP(c.f1); c2.f2 = SFM_DUP(c.f2);
        c2.f3 = SFM_DUP(c.f3); c2.f4 = SFM_DUP(c.f4); c2.f5 = SFM_DUP(c.f5);
        c2.f6 = SFM_DUP(c.f6); c2.f7 = SFM_DUP(c.f7); c2.f8 = SFM_DUP(c.f8);
        float nT = -c.T, u65 = 3.874301910400391e-6f;
        c2.nT = SFM_DUP(nT); c2.DeT = SFM_DUP(c.DeT); c2.Dn12 = SFM_DUP(c.Dn12);
        c2.C0 = SFM_DUP(c.C0); c2.U65 = SFM_DUP(u65);
        c2.nDe = SFM_DUP(c.nDe); c2.Ea1 = SFM_DUP(c.Ea1); c2.Ea0 = SFM_DUP(c.Ea0);
#undef SFM_DUP
    }
    int cnt = 0;
    const double *grow = x1x2 + (size_t)p0 * 4;
    const int npair = np >> 1;
    int j = 0;
    int unused = 0;
    for (; j + K2P_PAIRS <= npair; j += K2P_PAIRS)
        score_pairs<K2P_PAIRS, EARLY>(c2, spair + 2 * j, grow + (size_t)j * 8, Frow, two_thr, cnt, unused);
    for (; j < npair; ++j) score_pairs<1, EARLY>(c2, spair + 2 * j, grow + (size_t)j * 8, Frow, two_thr, cnt, unused);
    if (np & 1) {
        // odd tail of the last chunk: the scalar filter on lane 0 of the last pair
        float4 A = spair[2 * npair], B = spair[2 * npair + 1];
        float4 r = make_float4(A.x, A.z, B.x, B.z);
        score_group<1>(c, &r, grow + (size_t)(np - 1) * 4, Frow, two_thr, cnt);
    }
    if (hyp < h && cnt) atomicAdd(&counts[hyp], cnt);
}

// ------------------------------------------------------------------------------------ K3
// key = (count << 32) | (0xFFFFFFFF - global index): max(key) = largest count, and among equal
// counts the LOWEST index, which is the reference's strict '>' update rule.
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long w = __shfl_xor_sync(0xffffffffu, v, o);
        v = w > v ? w : v;
    }
    return v;
}

__global__ void __launch_bounds__(1024)
argmax_first_kernel(const int32_t *__restrict__ counts, int h, int idx_offset,
                    int32_t *__restrict__ best, unsigned long long *__restrict__ key_out) {
    __shared__ unsigned long long swarp[32];
    // one block per segment of h counts (sfm_argmax_first_batch; a single block for sfm_argmax_first)
    counts += (size_t)blockIdx.x * h;
    best += 2 * (size_t)blockIdx.x;
    if (key_out) key_out += blockIdx.x;
    unsigned long long k = 0ull;  // count 0 at "index 0xFFFFFFFF": loses to any real entry
    for (int i = threadIdx.x; i < h; i += blockDim.x) {
        unsigned long long c = (unsigned long long)(unsigned int)max(counts[i], 0);
        unsigned long long cand = (c << 32) | (unsigned long long)(0xFFFFFFFFu - (unsigned int)(i + idx_offset));
        k = cand > k ? cand : k;
    }
    k = warp_max_u64(k);
    if ((threadIdx.x & 31) == 0) swarp[threadIdx.x >> 5] = k;
    __syncthreads();
    if (threadIdx.x < 32) {
        k = threadIdx.x < (blockDim.x >> 5) ? swarp[threadIdx.x] : 0ull;
        k = warp_max_u64(k);
        if (threadIdx.x == 0) {
            best[0] = (int32_t)(0xFFFFFFFFu - (unsigned int)(k & 0xFFFFFFFFull));
            best[1] = (int32_t)(k >> 32);
            if (key_out) *key_out = k;
        }
    }
}

// ------------------------------------------------------------------------------------ K4
// F is read through a device pointer + optional device index so the winner never has to make
// a round trip to the host: Fone = Fbase + 9 * (best ? best[0] - idx_offset : 0).
__global__ void __launch_bounds__(256)
f_mask_kernel(const double *__restrict__ x1x2, int n, const double *__restrict__ Fbase,
              const int32_t *__restrict__ best, int idx_offset, int h, double two_thr,
              uint8_t *__restrict__ mask, double *__restrict__ err, double *__restrict__ Fcopy) {
    int sel = 0;
    bool valid = true;
    if (best) {
        // idx_offset < 0: Fbase already IS the winner's F (multi-GPU tail); only validity is read
        sel = idx_offset < 0 ? 0 : best[0] - idx_offset;
        valid = best[1] > 0 && sel >= 0 && sel < h;
        if (!valid) sel = 0;
    }
    const double *f = Fbase + (size_t)sel * 9;
    FMat F = {f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7], f[8]};
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (Fcopy && i < 9) Fcopy[i] = f[i];
    if (i >= n) return;
    const double2 *row = reinterpret_cast<const double2 *>(x1x2) + (size_t)i * 2;
    double2 a = __ldg(row), b = __ldg(row + 1);
    mask[i] = (valid && f_is_inlier(F, a.x, a.y, b.x, b.y, two_thr)) ? 1 : 0;
    if (err) err[i] = f_error_exact(F, a.x, a.y, b.x, b.y);
}

// chunk of correspondences per block: large enough to amortise the F load and the atomics,
// small enough that (tiles x chunks) covers the 148 SMs a few times over.
static int pick_chunk(int n, int h) {
    int tiles = ceil_div(h, K2_THREADS);
    int chunk = 1024;  // 32 KB of shared memory
    while (chunk > 128 && (long long)tiles * ceil_div(n, chunk) < 4LL * sm_count()) chunk >>= 1;
    return chunk;
}

// The same for the filter kernels, which run exactly `per_sm` CTAs per SM: all CTAs do the same
// work, so the grid should fill whole waves. Picks the smallest number of waves w for which
// floor(w * slots / tiles) point chunks of at most 1024 (even) correspondences cover n. Matters for
// the drop-in's chunked launches (2k-25k hypotheses each): a 2,048-hypothesis launch was 1.32
// waves with the power-of-two rule above.
static int pick_chunk_waves(int n, int h, int per_sm) {
    const long long tiles = ceil_div(h, K2_THREADS), slots = (long long)per_sm * sm_count();
    for (long long w = 1; w <= 4096; ++w) {
        const long long nchunks = w * slots / tiles;
        if (nchunks < 1) continue;
        long long chunk = (n + nchunks - 1) / nchunks;
        chunk += chunk & 1;                    // even: the packed kernel works on pairs
        if (chunk < 128) chunk = 128;
        if (chunk <= 1024) return (int)chunk;
    }
    return 1024;
}

}  // namespace sfm

using namespace sfm;

extern "C" int sfm_f8pt_batch(const double *d_x1x2, int32_t n, const int32_t *d_sample,
                              int32_t h, double *d_F, void *stream) {
    if (!d_x1x2 || !d_sample || !d_F || n < 8 || h < 0) return SFM_E_INVALID;
    if (h == 0) return SFM_OK;
    f8pt_kernel<<<ceil_div(h, 128), 128, 0, as_stream(stream)>>>(d_x1x2, d_sample, h, d_F);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

static int f_score_launch(const double *d_x1x2, int32_t n, const double *d_F, int32_t h,
                          double threshold, int32_t *d_counts, void *stream, bool all_fp64) {
    if (!d_x1x2 || !d_F || !d_counts || n < 0 || h < 0) return SFM_E_INVALID;
    if (h == 0) return SFM_OK;
    cudaStream_t st = as_stream(stream);
    SFM_CUDA_TRY(cudaMemsetAsync(d_counts, 0, sizeof(int32_t) * (size_t)h, st));
    if (n == 0) return SFM_OK;
    int chunk = all_fp64 ? pick_chunk(n, h) : pick_chunk_waves(n, h, K2P_CTAS);
    dim3 grid(ceil_div(h, K2_THREADS), ceil_div(n, chunk));
    if (!all_fp64) {
        // development switch: the scalar-FP32 filter the packed kernel replaced
        static const bool scalar32 = getenv("SFM_F_SCORE_SCALAR32") != nullptr;
        if (scalar32)
            f_score_mixed_kernel<<<grid, K2_THREADS, sizeof(float4) * (size_t)chunk, st>>>(
                d_x1x2, n, d_F, h, 2.0 * threshold, chunk, d_counts);
        else {
            // The early reject pays when whole warps certify their pairs as gross outliers: 97 % of the votes at
            // the reference's 0.06 px (K2 at C2: 4.82 -> 3.97 ms with one stage, 2.82 ms with the two-stage form),
            // but a third to a half of them fail at 1.0 px and then the extra test costs (4.83 -> 5.18 ms with one
            // stage; a per-warp adaptive switch lost on both: 4.20 / 5.04). Two-stage form against the plain kernel's
            // 4.82 ms (tools/k2_threshold_sweep.py, C2): 2.75 / 3.19 / 3.80 / 4.66 / 5.36 / 6.40 ms at
            // 0.06 / 0.15 / 0.3 / 0.6 / 1.0 / 3.0 px -- the crossover sits near 0.65 px.
            // Thresholds are in pixels in the reference's domain, so the launch picks by threshold. Either kernel
            // gives the FP64 kernel's counts. SFM_F_SCORE_NO_EARLY / SFM_F_SCORE_EARLY force one of them.
            static const int forced = getenv("SFM_F_SCORE_NO_EARLY") ? 0 : getenv("SFM_F_SCORE_EARLY") ? 1 : -1;
            const bool early = forced >= 0 ? forced == 1 : threshold < 0.65;
            if (!early)
                f_score_packed_kernel<false><<<grid, K2_THREADS, sizeof(float4) * (size_t)chunk, st>>>(
                    d_x1x2, n, d_F, h, 2.0 * threshold, chunk, d_counts);
            else
                f_score_packed_kernel<true><<<grid, K2_THREADS, sizeof(float4) * (size_t)chunk, st>>>(
                    d_x1x2, n, d_F, h, 2.0 * threshold, chunk, d_counts);
        }
        SFM_LAUNCH_CHECK();
        return SFM_OK;
    }
    size_t smem = (size_t)chunk * 32;
    static bool attr_set = false;
    if (!attr_set) {
        SFM_CUDA_TRY(cudaFuncSetAttribute(f_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          1024 * 32));
        attr_set = true;
    }
    f_score_kernel<<<grid, K2_THREADS, smem, st>>>(d_x1x2, n, d_F, h, 2.0 * threshold, chunk,
                                                   d_counts);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_f_score_count(const double *d_x1x2, int32_t n, const double *d_F, int32_t h,
                                 double threshold, int32_t *d_counts, void *stream) {
    static const bool all_fp64 = getenv("SFM_F_SCORE_F64") != nullptr;  // development switch
    return f_score_launch(d_x1x2, n, d_F, h, threshold, d_counts, stream, all_fp64);
}

// Diagnostics: the all-FP64 scoring kernel (what the certified filter must reproduce), and the
// number of point-hypothesis tests the filter handed to the FP64 path since the last reset.
extern "C" int sfm_f_score_count_fp64(const double *d_x1x2, int32_t n, const double *d_F, int32_t h,
                                      double threshold, int32_t *d_counts, void *stream) {
    return f_score_launch(d_x1x2, n, d_F, h, threshold, d_counts, stream, true);
}

extern "C" int sfm_f_score_fp64_tests(uint64_t *out, int32_t reset) {
    unsigned long long v = 0;
    SFM_CUDA_TRY(cudaDeviceSynchronize());
    SFM_CUDA_TRY(cudaMemcpyFromSymbol(&v, g_k2_fp64_tests, sizeof(v)));
    if (out) *out = v;
    if (reset) {
        v = 0;
        SFM_CUDA_TRY(cudaMemcpyToSymbol(g_k2_fp64_tests, &v, sizeof(v)));
    }
    return SFM_OK;
}

extern "C" int sfm_argmax_first(const int32_t *d_counts, int32_t h, int32_t idx_offset,
                                int32_t *d_best, uint64_t *d_key, void *stream) {
    if (!d_counts || !d_best || h < 0) return SFM_E_INVALID;
    argmax_first_kernel<<<1, 1024, 0, as_stream(stream)>>>(
        d_counts, h, idx_offset, d_best, reinterpret_cast<unsigned long long *>(d_key));
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_argmax_first_batch(const int32_t *d_counts, int32_t n_seg, int32_t seg_len, int32_t *d_best,
                                      void *stream) {
    if (!d_counts || !d_best || n_seg < 0 || seg_len < 0) return SFM_E_INVALID;
    if (n_seg == 0) return SFM_OK;
    argmax_first_kernel<<<n_seg, 256, 0, as_stream(stream)>>>(d_counts, seg_len, 0, d_best, nullptr);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_f_inlier_mask(const double *d_x1x2, int32_t n, const double *d_Fone,
                                 double threshold, uint8_t *d_mask, double *d_err, void *stream) {
    if (!d_x1x2 || !d_Fone || !d_mask || n < 0) return SFM_E_INVALID;
    if (n == 0) return SFM_OK;
    f_mask_kernel<<<ceil_div(n, 256), 256, 0, as_stream(stream)>>>(
        d_x1x2, n, d_Fone, nullptr, 0, 1, 2.0 * threshold, d_mask, d_err, nullptr);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

extern "C" int sfm_f_ransac(const double *d_x1x2, int32_t n, const int32_t *d_sample, int32_t h,
                            double threshold, double *d_F, int32_t *d_counts, int32_t *d_best,
                            double *d_Fbest, uint8_t *d_mask, double *d_err, void *stream) {
    if (!d_x1x2 || !d_sample || !d_F || !d_counts || !d_best || !d_Fbest || !d_mask || n < 8 ||
        h < 1)
        return SFM_E_INVALID;
    int rc;
    if ((rc = sfm_f8pt_batch(d_x1x2, n, d_sample, h, d_F, stream)) != SFM_OK) return rc;
    if ((rc = sfm_f_score_count(d_x1x2, n, d_F, h, threshold, d_counts, stream)) != SFM_OK) return rc;
    if ((rc = sfm_argmax_first(d_counts, h, 0, d_best, nullptr, stream)) != SFM_OK) return rc;
    f_mask_kernel<<<ceil_div(max(n, 9), 256), 256, 0, as_stream(stream)>>>(
        d_x1x2, n, d_F, d_best, 0, h, 2.0 * threshold, d_mask, d_err, d_Fbest);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

// ---- multi-GPU tail ---------------------------------------------------------------------------
// After the 8-byte max-allreduce of the packed key every rank decodes the global winner and
// recomputes its F from the winning sample (all ranks hold the whole sample stream), so no
// broadcast of F or of the mask is needed and all ranks end with identical bits.

extern "C" int sfm_f_winner_from_key(const double *d_x1x2, int32_t n, const int32_t *d_sample_all,
                                     int32_t h_total, const uint64_t *d_key, double threshold,
                                     int32_t *d_best, double *d_Fbest, uint8_t *d_mask,
                                     double *d_err, void *stream) {
    if (!d_x1x2 || !d_sample_all || !d_key || !d_best || !d_Fbest || !d_mask || n < 8 || h_total < 1)
        return SFM_E_INVALID;
    f8pt_kernel<<<1, 32, 0, as_stream(stream)>>>(d_x1x2, d_sample_all, h_total, d_Fbest,
                                                 reinterpret_cast<const unsigned long long *>(d_key), d_best);
    SFM_LAUNCH_CHECK();
    // Fbase = d_Fbest, selected entry 0 (idx_offset = best[0]); valid iff best[1] > 0
    f_mask_kernel<<<ceil_div(n, 256), 256, 0, as_stream(stream)>>>(
        d_x1x2, n, d_Fbest, d_best, /*idx_offset (device-side)*/ -1, 1, 2.0 * threshold, d_mask, d_err,
        nullptr);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}
