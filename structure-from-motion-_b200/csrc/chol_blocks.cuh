// chol_blocks.cuh -- 48 x 48 block primitives shared by the two K11 kernels (chol.cu: dense right-looking
// panels; chol_dag.cu: structure-aware left-looking task graph): shared-memory block views, the diagonal
// factor + inverse, and the 48 x 48 x 48 products on the DFMA and the FP64 tensor (DMMA) path.
#pragma once
#include "sfm_common.cuh"

namespace sfm {

constexpr int CH_NB = 48;          // panel width: a multiple of 6 (camera blocks) and of 8
constexpr int CH_LDS = CH_NB + 4;  // padded shared-memory row: stride 52 = 4 mod 16 keeps the 8 x 4 DMMA
                                   // fragment reads (row = lane>>2, col = lane&3) conflict-free per half-warp
constexpr int CH_THREADS = 256;
constexpr int CH_SUPER_MIN = 20;   // trailing blocks from which phase 2 uses 96 x 96 super-tiles
constexpr int CH_SINGLE_BACK = 16;  // up to this many panels the back substitution runs in one CTA

struct CholSmem {
    // phase 1 / small phase 2: three NB x NB block views L | B | C; large phase 2: panel stages
    double buf[2 * 2 * 2 * CH_NB * CH_LDS];   // two stages of two 96 x 48 panels
    double x[CH_NB];
    short act[1024];      // phase 2: super-rows of the trailing matrix whose panel blocks are not all zero
    int nact, nact_cols;
    int fail;
};
static_assert(2 * 2 * 2 * CH_NB * CH_LDS >= 3 * CH_NB * CH_LDS, "block views must fit");

// load an NB x NB block (rows r0.., cols c0..) of the row-major matrix into padded smem
__device__ __forceinline__ void load_block(double *s, const double *A, size_t ld, int r0, int c0) {
    for (int e = threadIdx.x; e < CH_NB * CH_NB; e += CH_THREADS) {
        int r = e / CH_NB, c = e - r * CH_NB;
        s[r * CH_LDS + c] = A[(size_t)(r0 + r) * ld + c0 + c];
    }
}
__device__ __forceinline__ void store_block(const double *s, double *A, size_t ld, int r0, int c0) {
    for (int e = threadIdx.x; e < CH_NB * CH_NB; e += CH_THREADS) {
        int r = e / CH_NB, c = e - r * CH_NB;
        A[(size_t)(r0 + r) * ld + c0 + c] = s[r * CH_LDS + c];
    }
}

// 1/sqrt(d) to full double precision: MUFU seed + two Newton steps (the DP sqrt and divide
// sequences are several hundred cycles of dependent latency on the critical path of every panel).
__device__ __forceinline__ double rsqrt_full(double d) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    double h = 0.5 * d;
    y = y * fma(-h * y, y, 1.5);
    y = y * fma(-h * y, y, 1.5);
    return y;
}

// -1/d to full double precision: MUFU seed (about 20 bits) and one cubic correction,
// r = -y0 (1 + e + e^2) with e = 1 - d y0: three dependent DFMAs after the MUFU.
__device__ __forceinline__ double neg_rcp_full(double d) {
    double y0;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(d));
    const double e = fma(-d, y0, 1.0);
    const double q = fma(e, e, e);
    return fma(-y0, q, -y0);
}

// Cholesky factor L (first block view of sm.buf, lower, strict upper part zero) and L^-1 (second
// view) of the NB x NB diagonal block at Ablk (row-major, leading dimension ld, lower triangle read).
//
// Sweep: right-looking, two columns per step, the block held in REGISTERS of six warps. The sweep is
// one long dependent chain; tools/diag_bench.py shows its cost is that chain plus the time the block
// barrier waits for the shared-memory stores issued before it (roughly 30 cycles per STS
// instruction, and twice that for a 16-byte one), so a step crosses ONE barrier and issues four
// 8-byte stores per publishing thread for TWO columns. Thread (ty, tx) owns rows 2ty, 2ty+1 and columns 6tx .. 6tx+5; a warp
// holds eight consecutive rows and drops out once the sweep has passed them. For the column pair
// (2p, 2p+1) the owners published, at the end of step p-1, both columns as updated through column
// 2p-1 into P2[p][col][row]; every thread then derives by itself -1/d1, the multiplier
// t = -a_{2p+1,2p}/d1, the second pivot d2 = a_{2p+1,2p+1} + t a_{2p+1,2p}, -1/d2, and the updated
// second column, and applies the rank-2 update to its tile without predicates (entries that are
// already final just collect garbage nothing reads). P2 keeps every published pair, so the scaled
// factor is written by all threads after the sweep. Dependent chain per PAIR of columns:
// barrier -> LDS -> rcp -> 2 DP ops -> rcp -> DMUL -> DFMA -> STS.
//
// Inverse: blocked, after the sweep, all threads: the three 16 x 16 diagonal blocks by column-wise
// forward substitution in registers, then X10 = -X11 (L10 X00), X21 = -X22 (L21 X11),
// X20 = -X22 (L20 X00 + L21 X10) as small dense products.
// Sets sm.fail on a pivot that is not positive and finite (NaNs then fill the block; callers test
// the flag). Uses the first four block views of sm.buf.
template <int V = 0>
__device__ void factor_diag_block(CholSmem &sm, const double *Ablk, size_t ld, unsigned long long *prof = nullptr) {
    const long long c0 = clock64();
    double *Lout = sm.buf, *Xout = sm.buf + CH_NB * CH_LDS;
    double *P2 = sm.buf + 2 * CH_NB * CH_LDS;     // [NB/2 pairs][2 cols][NB rows]
    double *sinv = sm.buf + 3 * CH_NB * CH_LDS;   // [NB] 1/sqrt(pivot)
    double *st = sinv + CH_NB;                     // [NB/2] multiplier t of each pair
    double *T = sinv + 2 * CH_NB;                  // 2 x 16 x 16 scratch for the inverse
    constexpr int HP = CH_NB / 2, PROW = HP * 4;   // 24 pairs, 96 doubles per published pair
    const bool sweeper = threadIdx.x < 192;
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;      // sweepers: ty in [0, 24)
    const int last_p = sweeper ? 4 * (int)(threadIdx.x >> 5) + 3 : -1;   // this warp's rows end at 8w+7
    double A[2][6];
    if (sweeper) {
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 6; b += 2) {
                const int r = 2 * ty + a, c = 6 * tx + b;
                const double2 v = *reinterpret_cast<const double2 *>(&Ablk[(size_t)r * ld + c]);
                A[a][b] = (c <= r) ? v.x : 0.0;
                A[a][b + 1] = (c + 1 <= r) ? v.y : 0.0;
            }
        if (tx == 0) {                             // pair 0
            P2[2 * ty] = A[0][0]; P2[2 * ty + 1] = A[1][0];
            P2[CH_NB + 2 * ty] = A[0][1]; P2[CH_NB + 2 * ty + 1] = A[1][1];
        }
    }
    for (int e = threadIdx.x; e < CH_NB * CH_NB; e += CH_THREADS) {
        int r = e / CH_NB, c = e - r * CH_NB;
        Xout[r * CH_LDS + c] = 0.0;
    }
    __syncthreads();
    const long long c1 = clock64();
    // software pipeline: a step applies its rank-2 update only to the two tile columns published
    // next; the other four columns take it at the top of the following step, where those DFMAs
    // fill the latency slots of the pivot chain instead of delaying this step's barrier
    double pxs[4] = {0.0, 0.0, 0.0, 0.0}, py1[6], py2[6];
#pragma unroll
    for (int b = 0; b < 6; ++b) py1[b] = py2[b] = 0.0;
    for (int p0 = 0; p0 < HP; p0 += 3) {
#pragma unroll
        for (int u = 0; u < 3; ++u) {
            const int p = p0 + u;
            if (p <= last_p) {                     // warp-uniform: rows above 2p are finished
                const double *col0 = P2 + p * PROW, *col1 = col0 + CH_NB;
                const double2 dA = *reinterpret_cast<const double2 *>(col0 + 2 * p);   // d1, c
                const double d2p = col1[2 * p + 1];
                const double2 x1 = *reinterpret_cast<const double2 *>(col0 + 2 * ty);
                const double2 x2p = *reinterpret_cast<const double2 *>(col1 + 2 * ty);
                double y1[6], y2[6];
#pragma unroll
                for (int b = 0; b < 6; b += 2) {
                    const double2 ya = *reinterpret_cast<const double2 *>(col0 + 6 * tx + b);
                    const double2 yb = *reinterpret_cast<const double2 *>(col1 + 6 * tx + b);
                    y1[b] = ya.x; y1[b + 1] = ya.y;
                    y2[b] = yb.x; y2[b + 1] = yb.y;
                }
                // deferred part of the previous step (its published columns were 2*(u%3), +1)
                const int bp = 2 * (u % 3);
#pragma unroll
                for (int b = 0; b < 6; ++b) {
                    if (b == bp || b == bp + 1) continue;
                    A[0][b] = fma(pxs[2], py2[b], fma(pxs[0], py1[b], A[0][b]));
                    A[1][b] = fma(pxs[3], py2[b], fma(pxs[1], py1[b], A[1][b]));
                }
                const double r1 = neg_rcp_full(dA.x);
                const double t = dA.y * r1;
                const double r2 = neg_rcp_full(fma(t, dA.y, d2p));
                pxs[0] = x1.x * r1; pxs[1] = x1.y * r1;
                pxs[2] = fma(x1.x, t, x2p.x) * r2; pxs[3] = fma(x1.y, t, x2p.y) * r2;
#pragma unroll
                for (int b = 0; b < 6; ++b) { py1[b] = y1[b]; py2[b] = fma(y1[b], t, y2[b]); }
                const int bn = 2 * ((u + 1) % 3);
#pragma unroll
                for (int b = bn; b < bn + 2; ++b) {
                    A[0][b] = fma(pxs[2], py2[b], fma(pxs[0], py1[b], A[0][b]));
                    A[1][b] = fma(pxs[3], py2[b], fma(pxs[1], py1[b], A[1][b]));
                }
                if (V != 4 && tx == p0 / 3 + (u + 1) / 3) {     // 8-byte stores: STS.128 costs ~25 cycles more each
                    volatile double *n0 = P2 + (p + 1) * PROW + 2 * ty;   // volatile: keeps ptxas from pairing them
                    n0[0] = A[0][bn]; n0[1] = A[1][bn];
                    n0[CH_NB] = A[0][bn + 1]; n0[CH_NB + 1] = A[1][bn + 1];
                }
            }
            if (V == 5) __syncwarp();
            else __syncthreads();
        }
    }
    if (V == 5) __syncthreads();
    // pivots -> 1/sqrt, multipliers; then L = published columns scaled (strict upper part zero)
    if (threadIdx.x < HP) {
        const int p = threadIdx.x;
        const double *col0 = P2 + p * PROW;
        const double d1 = col0[2 * p], c = col0[2 * p + 1];
        const double t = c * neg_rcp_full(d1);
        const double d2 = fma(t, c, col0[CH_NB + 2 * p + 1]);
        const bool bad1 = !(d1 > 1e-290) || !(d1 < 1e290), bad2 = !(d2 > 1e-290) || !(d2 < 1e290);
        sinv[2 * p] = bad1 ? 1.0 : rsqrt_full(d1);
        sinv[2 * p + 1] = bad2 ? 1.0 : rsqrt_full(d2);
        st[p] = t;
        if (bad1 || bad2) sm.fail = 1;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < CH_NB * CH_NB; e += CH_THREADS) {
        const int j = e / CH_NB, i = e - j * CH_NB, p = j >> 1;
        const double *col0 = P2 + p * PROW;
        const double v = (j & 1) ? fma(col0[i], st[p], col0[CH_NB + i]) : col0[i];
        Lout[i * CH_LDS + j] = (j <= i) ? v * sinv[j] : 0.0;
    }
    __syncthreads();
    const long long c2 = clock64();
    // ---- inverse ----
    // diagonal 16 x 16 blocks: thread (blk, c) builds column c of L_bb^-1 in registers
    if (threadIdx.x < CH_NB) {
        const int blk = threadIdx.x >> 4, c = threadIdx.x & 15, o = 16 * blk;
        double x[16], acc[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) { x[i] = 0.0; acc[i] = (i == c) ? 1.0 : 0.0; }
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            // x_i = (e_c - sum_{m<i} L[i][m] x_m)_i / L[i][i]; zero above the diagonal (i < c)
            const double xv = (i >= c) ? acc[i] * sinv[o + i] : 0.0;
            x[i] = xv;
#pragma unroll
            for (int r = i + 1; r < 16; ++r) acc[r] = fma(-Lout[(o + r) * CH_LDS + o + i], xv, acc[r]);
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) Xout[(o + i) * CH_LDS + o + c] = x[i];
    }
    __syncthreads();
    {
        const int r = threadIdx.x >> 4, c = threadIdx.x & 15;      // one output of a 16 x 16 product each
        // T0 = L10 X00, T1 = L21 X11
        double t0 = 0.0, t1 = 0.0;
#pragma unroll
        for (int m = 0; m < 16; ++m) {
            t0 = fma(Lout[(16 + r) * CH_LDS + m], Xout[m * CH_LDS + c], t0);
            t1 = fma(Lout[(32 + r) * CH_LDS + 16 + m], Xout[(16 + m) * CH_LDS + 16 + c], t1);
        }
        T[r * 16 + c] = t0;
        T[256 + r * 16 + c] = t1;
        __syncthreads();
        // X10 = -X11 T0, X21 = -X22 T1
        double u0 = 0.0, u1 = 0.0;
#pragma unroll
        for (int m = 0; m < 16; ++m) {
            u0 = fma(Xout[(16 + r) * CH_LDS + 16 + m], T[m * 16 + c], u0);
            u1 = fma(Xout[(32 + r) * CH_LDS + 32 + m], T[256 + m * 16 + c], u1);
        }
        __syncthreads();
        Xout[(16 + r) * CH_LDS + c] = -u0;
        Xout[(32 + r) * CH_LDS + 16 + c] = -u1;
        __syncthreads();
        // T0 = L20 X00 + L21 X10
        double t2 = 0.0;
#pragma unroll
        for (int m = 0; m < 16; ++m) {
            t2 = fma(Lout[(32 + r) * CH_LDS + m], Xout[m * CH_LDS + c], t2);
            t2 = fma(Lout[(32 + r) * CH_LDS + 16 + m], Xout[(16 + m) * CH_LDS + c], t2);
        }
        T[r * 16 + c] = t2;
        __syncthreads();
        double u2 = 0.0;
#pragma unroll
        for (int m = 0; m < 16; ++m) u2 = fma(Xout[(32 + r) * CH_LDS + 32 + m], T[m * 16 + c], u2);
        Xout[(32 + r) * CH_LDS + c] = -u2;
    }
    __syncthreads();
    if (prof && threadIdx.x == 0) {          // diagnostic: SM cycles of the column sweep / whole factor
        prof[6] += (unsigned long long)(c2 - c1);
        prof[7] += (unsigned long long)(clock64() - c0);
    }
}

// C(48x48) = or -= X * Y^T with X, Y, C in padded smem; thread (ty, tx) of a 16 x 16 layout owns
// the 3 x 3 outputs (ty + 16a, tx + 16b). MODE 0: C = X Y^T; MODE 1: C -= X Y^T.
template <int MODE>
__device__ __forceinline__ void block_xyT(const double *X, const double *Y, double *C) {
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    double acc[3][3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) acc[a][b] = 0.0;
    // each group of four lanes walks the reduction index rotated by one more column: with the row
    // stride of 52 (= 4 mod 16 banks) this makes the 16 row reads of a half-warp conflict-free
    const int skew = tx >> 2;
#pragma unroll 4
    for (int m0 = 0; m0 < CH_NB; ++m0) {
        int m = m0 + skew;
        m = m >= CH_NB ? m - CH_NB : m;
        double xv[3], yv[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) xv[a] = X[(ty + 16 * a) * CH_LDS + m];
#pragma unroll
        for (int b = 0; b < 3; ++b) yv[b] = Y[(tx + 16 * b) * CH_LDS + m];
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int b = 0; b < 3; ++b) acc[a][b] = fma(xv[a], yv[b], acc[a][b]);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            double *c = &C[(ty + 16 * a) * CH_LDS + tx + 16 * b];
            if (MODE == 0) *c = acc[a][b]; else *c -= acc[a][b];
        }
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ void cp_async_8(double *smem_dst, const double *gmem_src) {
    unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// D(8x8) = A(8x4) B(4x8) + C on the FP64 tensor path (DMMA). Fragments: a = A[lane>>2][lane&3],
// b = B[lane&3][lane>>2], c/d = C[lane>>2][2*(lane&3) + {0,1}].
__device__ __forceinline__ void dmma_884(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// block_xyT on the FP64 tensor path: warp w < 6 owns rows 8w .. 8w+7 of C (six 8 x 8 tiles), one A and six B
// fragment loads (conflict-free at the row stride of 52) for six DMMAs per k-step. The scalar form above issues
// 288 LDS.64 and 432 DFMAs per thread and takes ~2 us per 48 x 48 x 48 product; two of them sit on CTA 0's chain
// of every panel.
template <int MODE>
__device__ __forceinline__ void block_xyT_mma(const double *X, const double *Y, double *C) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp >= 6) return;
    const int g = lane >> 2, q = lane & 3;
    double acc[6][2];
    double *pc = C + (8 * warp + g) * CH_LDS + 2 * q;
#pragma unroll
    for (int t = 0; t < 6; ++t) {
        if (MODE == 0) {
            acc[t][0] = 0.0; acc[t][1] = 0.0;
        } else {
            const double2 v = *reinterpret_cast<const double2 *>(pc + 8 * t);
            acc[t][0] = v.x; acc[t][1] = v.y;
        }
    }
    const double *pa = X + (8 * warp + g) * CH_LDS + q;
    const double *pb = Y + g * CH_LDS + q;
#pragma unroll 4
    for (int k4 = 0; k4 < CH_NB; k4 += 4) {
        const double a = (MODE == 0) ? pa[k4] : -pa[k4];
        double bf[6];
#pragma unroll
        for (int t = 0; t < 6; ++t) bf[t] = pb[8 * t * CH_LDS + k4];
#pragma unroll
        for (int t = 0; t < 6; ++t) dmma_884(acc[t][0], acc[t][1], a, bf[t]);
    }
#pragma unroll
    for (int t = 0; t < 6; ++t) *reinterpret_cast<double2 *>(pc + 8 * t) = make_double2(acc[t][0], acc[t][1]);
}

}  // namespace sfm
