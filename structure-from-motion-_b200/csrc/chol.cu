// chol.cu -- K11: blocked right-looking Cholesky factorisation and solve of the reduced camera
// system S dc = g_red, as ONE cooperative persistent kernel (grid-wide barriers between panel
// steps instead of ~3 launches per panel).
//
// In the reference this solve is hidden inside MINPACK (dense pivoted QR of the full Jacobian,
// BundleAdjustment.py:205-212 -> scipy least_squares(method='lm') -> lmder/lmpar/qrsolv); here it
// is the dense SPD solve of the Schur complement.
//
// Layout: A is row-major with leading dimension ld; only the lower triangle of its leading
// npad x npad part is read or written (npad = multiple of NB; rows/cols >= n are identity
// padding supplied by the caller). The right-hand side rides along as block-row nb (rows
// npad .. npad+NB-1, one meaningful row): it goes through the panel triangular solves and
// the trailing updates like any other block-row, which leaves y = L^-1 g in it for free; the
// backward substitution L^T x = y then runs block by block from the bottom.
//
// Per panel step k (two grid barriers):
//   phase 1  every CTA factors the NB x NB diagonal block in shared memory (redundantly: cheaper
//            than a barrier), inverts the factor, and solves its share of the block-rows below:
//            A_ik <- A_ik L_kk^-T as a dense product with the inverse;
//   phase 2  trailing update A_ij -= L_ik L_jk^T over the lower-triangle tiles, round-robin.
#include "chol_blocks.cuh"
#include <cooperative_groups.h>
#include <stdlib.h>

namespace cg = cooperative_groups;

namespace sfm {

// ---- phase-2 super-tile: a 96 x 96 block of the trailing matrix (2 x 2 panel blocks) per CTA ----
// C -= P_I P_J^T with P_I, P_J the 96 x 48 slices of panel k. The panels of the CTA's NEXT tile
// stream into the other shared-memory stage with cp.async while the current tile is computed;
// the 6 x 6 accumulators per thread start from C (read straight from global memory before the
// wait) and go straight back. Rows >= row_lim (beyond the right-hand-side block-row), columns >=
// col_lim (beyond the matrix) and rows / columns < lo (the panel itself, when the trailing matrix
// starts on an odd block) are masked.
constexpr int CH_ST = 2 * CH_NB;               // 96
constexpr int CH_PANEL = CH_ST * CH_LDS;       // doubles per staged 96 x 48 panel


__device__ __forceinline__ void issue_panel96(double *s, const double *A, size_t ld, int r0, int c0, int row_lim) {
    for (int e = threadIdx.x; e < CH_ST * CH_NB; e += CH_THREADS) {
        int r = e / CH_NB, c = e - r * CH_NB;
        if (r0 + r < row_lim) cp_async_8(&s[r * CH_LDS + c], &A[(size_t)(r0 + r) * ld + c0 + c]);
        else s[r * CH_LDS + c] = 0.0;
    }
}

struct SuperTiles {
    int nJ, tri, ntiles;
    const short *act;
    // Builds (all threads of the CTA, identical result in every CTA) the ascending list of ACTIVE
    // super-rows of the trailing matrix at panel k: pairs of block-rows 2s, 2s+1 of which at least
    // one has a non-zero panel block L_ik. The first super-row (it holds the next diagonal block,
    // whose owner factors it) and the right-hand-side block-row are always active.
    __device__ __forceinline__ void init(CholSmem &sm, const unsigned char *nz, int k, int nb) {
        const int b0 = (k + 1) >> 1;             // first super-row/col (may start one block early)
        const int nI = ((nb + 2) >> 1) - b0;     // super-rows covering block-rows .. nb (RHS)
        __syncthreads();
        if (threadIdx.x < 32) {
            int base = 0, cols = 0;
            for (int s0 = 0; s0 < nI; s0 += 32) {
                const int sidx = b0 + s0 + (int)threadIdx.x;
                bool a = false;
                if (s0 + (int)threadIdx.x < nI) {
                    const int i0 = 2 * sidx, i1 = 2 * sidx + 1;
                    a = (sidx == b0) || i0 == nb || i1 == nb;
                    if (i0 > k && i0 < nb) a = a || nz[(size_t)i0 * nb + k];
                    if (i1 > k && i1 < nb) a = a || nz[(size_t)i1 * nb + k];
                }
                const unsigned m = __ballot_sync(0xffffffffu, a);
                if (a) sm.act[base + __popc(m & ((1u << threadIdx.x) - 1u))] = (short)sidx;
                const bool col_ok = 2 * sidx < nb;      // holds at least one real block column
                cols += __popc(__ballot_sync(0xffffffffu, a && col_ok));
                base += __popc(m);
            }
            if (threadIdx.x == 0) { sm.nact = base; sm.nact_cols = cols; }
        }
        __syncthreads();
        act = sm.act;
        const int c = sm.nact;
        nJ = sm.nact_cols;                       // the column-capable ones come first (ascending order)
        tri = nJ * (nJ + 1) / 2;
        ntiles = tri + (c - nJ) * nJ;
    }
    __device__ __forceinline__ void decode(int t, int &I, int &J) const {
        int Ip, Jp;
        if (t < tri) {
            Ip = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
            while ((Ip + 1) * (Ip + 2) / 2 <= t) ++Ip;
            while (Ip * (Ip + 1) / 2 > t) --Ip;
            Jp = t - Ip * (Ip + 1) / 2;
        } else {
            Ip = nJ + (t - tri) / nJ;
            Jp = (t - tri) - (Ip - nJ) * nJ;
        }
        I = act[Ip];
        J = act[Jp];
    }
};


// Phase-2 work split (G > 1): CTA 0 is the critical path -- it updates ONLY the next diagonal block
// (one 48 x 48 product with the block it solved in phase 1) and factors it right away; the
// super-tiles, with that block masked out of tile 0, go round-robin over CTAs 1 .. G-1. (CTA 0 used
// to take the whole 96 x 96 tile 0 before factoring: ~4 us more on the chain of every panel.)
__device__ __forceinline__ int st_first_tile(int cta, int G) { return G == 1 ? 0 : (cta == 0 ? 0x3fffffff : cta - 1); }
__device__ __forceinline__ int st_next_tile(int t, int cta, int G) {
    if (G == 1) return t + 1;
    return t + (G - 1);
}

template <bool DMMA, class FactorFn>
__device__ void supertile_phase(CholSmem &sm, double *A, size_t ld, const unsigned char *nz, int k, int nb,
                                int cta, int G, FactorFn &factor_next) {
    double *buf = sm.buf;
    const int kcol = k * CH_NB, lo = (k + 1) * CH_NB, row_lim = (nb + 1) * CH_NB, col_lim = nb * CH_NB;
    const bool split = G > 1;
    if (split && cta == 0) {
        double *const sL = buf, *const sC = buf + 2 * CH_NB * CH_LDS;
        __syncthreads();                          // the phase-1 views of buf are dead from here on
        load_block(sL, A, ld, lo, kcol);
        load_block(sC, A, ld, lo, lo);
        __syncthreads();
        block_xyT<1>(sL, sL, sC);                 // A_{k+1,k+1} -= L_{k+1,k} L_{k+1,k}^T
        __syncthreads();
        store_block(sC, A, ld, lo, lo);
        factor_next(k + 1);
        return;
    }
    const int dlim = split ? lo + CH_NB : lo;     // rows and columns both below dlim: CTA 0's block
    SuperTiles st;
    st.init(sm, nz, k, nb);
    int t = st_first_tile(cta, G), I = 0, J = 0, stage = 0;
    __syncthreads();                              // the phase-1 views of buf are dead from here on
    if (t < st.ntiles) {
        st.decode(t, I, J);
        issue_panel96(buf, A, ld, I * CH_ST, kcol, row_lim);
        if (I != J) issue_panel96(buf + CH_PANEL, A, ld, J * CH_ST, kcol, row_lim);
    }
    cp_async_commit();
    while (t < st.ntiles) {
        const int tn = st_next_tile(t, cta, G);
        int In = 0, Jn = 0;
        if (tn < st.ntiles) {
            st.decode(tn, In, Jn);
            double *nxt = buf + (stage ^ 1) * 2 * CH_PANEL;
            issue_panel96(nxt, A, ld, In * CH_ST, kcol, row_lim);
            if (In != Jn) issue_panel96(nxt + CH_PANEL, A, ld, Jn * CH_ST, kcol, row_lim);
        }
        cp_async_commit();
        const double *sI = buf + stage * 2 * CH_PANEL;
        const double *sJ = (I != J) ? sI + CH_PANEL : sI;
        if (DMMA) {
            // warp (wr, wc) of a 2 x 4 grid owns rows wr*48.., cols wc*24..: 6 x 3 DMMA tiles
            const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
            const int wr = warp >> 2, wc = warp & 3, g = lane >> 2, q = lane & 3;
            const int r0 = I * CH_ST + wr * 48 + g, c0 = J * CH_ST + wc * 24 + 2 * q;
            double acc[6][3][2];
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const int r = r0 + 8 * a;
#pragma unroll
                for (int b = 0; b < 3; ++b) {
                    const int c = c0 + 8 * b;       // c even, col_lim and lo even: the pair is in or out together
                    if (r < row_lim && r >= lo && c < col_lim && c >= lo && !(r < dlim && c < dlim)) {
                        const double2 v = *reinterpret_cast<const double2 *>(&A[(size_t)r * ld + c]);
                        acc[a][b][0] = v.x; acc[a][b][1] = v.y;
                    } else {
                        acc[a][b][0] = 0.0; acc[a][b][1] = 0.0;
                    }
                }
            }
            cp_async_wait<1>();                   // everything but the group just committed has landed
            __syncthreads();
            const double *pa = sI + (wr * 48 + g) * CH_LDS + q;
            const double *pb = sJ + (wc * 24 + g) * CH_LDS + q;
#pragma unroll 3
            for (int k4 = 0; k4 < CH_NB; k4 += 4) {
                double af[6], bf[3];
#pragma unroll
                for (int a = 0; a < 6; ++a) af[a] = -pa[a * 8 * CH_LDS + k4];
#pragma unroll
                for (int b = 0; b < 3; ++b) bf[b] = pb[b * 8 * CH_LDS + k4];
#pragma unroll
                for (int a = 0; a < 6; ++a)
#pragma unroll
                    for (int b = 0; b < 3; ++b) dmma_884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
            }
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const int r = r0 + 8 * a;
                if (r >= row_lim || r < lo) continue;
#pragma unroll
                for (int b = 0; b < 3; ++b) {
                    const int c = c0 + 8 * b;
                    if (c < col_lim && c >= lo && !(r < dlim && c < dlim))
                        *reinterpret_cast<double2 *>(&A[(size_t)r * ld + c]) = make_double2(acc[a][b][0], acc[a][b][1]);
                }
            }
        } else {
            const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
            double acc[6][6];
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const int r = I * CH_ST + ty + 16 * a;
#pragma unroll
                for (int b = 0; b < 6; ++b) {
                    const int c = J * CH_ST + tx + 16 * b;
                    acc[a][b] = (r < row_lim && r >= lo && c < col_lim && c >= lo && !(r < dlim && c < dlim)) ? A[(size_t)r * ld + c] : 0.0;
                }
            }
            cp_async_wait<1>();
            __syncthreads();
            const int skew = tx >> 2;
#pragma unroll 2
            for (int m0 = 0; m0 < CH_NB; ++m0) {
                int m = m0 + skew;
                m = m >= CH_NB ? m - CH_NB : m;
                double xv[6], yv[6];
#pragma unroll
                for (int a = 0; a < 6; ++a) xv[a] = -sI[(ty + 16 * a) * CH_LDS + m];
#pragma unroll
                for (int b = 0; b < 6; ++b) yv[b] = sJ[(tx + 16 * b) * CH_LDS + m];
#pragma unroll
                for (int a = 0; a < 6; ++a)
#pragma unroll
                    for (int b = 0; b < 6; ++b) acc[a][b] = fma(xv[a], yv[b], acc[a][b]);
            }
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const int r = I * CH_ST + ty + 16 * a;
                if (r >= row_lim || r < lo) continue;
#pragma unroll
                for (int b = 0; b < 6; ++b) {
                    const int c = J * CH_ST + tx + 16 * b;
                    if (c < col_lim && c >= lo && !(r < dlim && c < dlim)) A[(size_t)r * ld + c] = acc[a][b];
                }
            }
        }
        __syncthreads();                          // this stage may be refilled by the next iteration's loads
        if (t == 0 && !split) factor_next(k + 1); // single CTA: tile 0 holds the next diagonal block; its
                                                  // factor works in this (now free) stage's first 3 views
        stage ^= 1;
        t = tn;
        I = In;
        J = Jn;
    }
    cp_async_wait<0>();
}

// ---- phase 2 for block-sparse panels (CTAs 1 .. G-1 of a split grid) ----
// When only a few block-rows below the diagonal hold a non-zero panel block (banded / arrowhead reduced
// camera systems: ~10 of 125 at C5), 96 x 96 super-tiles mostly multiply zeros and their ~13 us per tile
// outlast CTA 0's syrk + factor chain (10 us). Here the active block-rows are listed at 48-block
// granularity and the tiles (I, J), I >= J both active, are plain 48 x 48 products, one or two per CTA.
// Returns false (nothing done) when more than CH_SPARSE_ROWS rows are active: the caller then runs the
// super-tile phase. Every CTA reaches the same verdict from the same flags. Measured at C5 (a trailing
// CTA's phase 2): 3.6 us per panel; the panel step is then bounded by CTA 0's chain alone (syrk + 48 x 48
// factor + commit, 10.5 us) plus the two grid barriers.
constexpr int CH_SPARSE_ROWS = 24;

__device__ bool sparse_tile_phase(CholSmem &sm, double *A, size_t ld, const unsigned char *nz, int k, int nb,
                                  int cta, int G) {
    double *const sL = sm.buf, *const sB = sm.buf + CH_NB * CH_LDS, *const sC = sm.buf + 2 * CH_NB * CH_LDS;
    __syncthreads();                              // the phase-1 views of buf and sm.act are free from here on
    if (threadIdx.x < 32) {
        int base = 0;
        for (int i0 = k + 1; i0 < nb && base <= CH_SPARSE_ROWS; i0 += 32) {
            const int i = i0 + (int)threadIdx.x;
            const bool a = i < nb && nz[(size_t)i * nb + k];
            const unsigned m = __ballot_sync(0xffffffffu, a);
            const int at = base + __popc(m & ((1u << threadIdx.x) - 1u));
            if (a && at <= CH_SPARSE_ROWS) sm.act[at] = (short)i;
            base += __popc(m);
        }
        if (threadIdx.x == 0) sm.nact = base;     // block columns; the right-hand-side row is implicit
    }
    __syncthreads();
    const int nJ = sm.nact;
    if (nJ > CH_SPARSE_ROWS) return false;
    const int tri = nJ * (nJ + 1) / 2, ntiles = tri + nJ, lo = k + 1;
    for (int t = cta - 1; t < ntiles; t += G - 1) {
        int Ip, Jp;
        if (t < tri) {
            Ip = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
            while ((Ip + 1) * (Ip + 2) / 2 <= t) ++Ip;
            while (Ip * (Ip + 1) / 2 > t) --Ip;
            Jp = t - Ip * (Ip + 1) / 2;
        } else {
            Ip = nJ;
            Jp = t - tri;
        }
        const int I = Ip < nJ ? sm.act[Ip] : nb, J = sm.act[Jp];
        if (I == lo && J == lo) continue;         // the next diagonal block: CTA 0 updates and factors it
        __syncthreads();
        load_block(sL, A, ld, I * CH_NB, k * CH_NB);
        load_block(sB, A, ld, J * CH_NB, k * CH_NB);
        load_block(sC, A, ld, I * CH_NB, J * CH_NB);
        __syncthreads();
        block_xyT<1>(sL, sB, sC);
        __syncthreads();
        store_block(sC, A, ld, I * CH_NB, J * CH_NB);
    }
    return true;
}

// nb = npad / NB diagonal blocks; block-row nb is the right-hand side. x_out [npad].
// Linv_ws [nb, NB*NB]: inverses of the diagonal factors (for the backward substitution).
template <bool DMMA>
__global__ void __launch_bounds__(CH_THREADS)
chol_solve_kernel(double *__restrict__ A, size_t ld, int nb, double *__restrict__ Linv_ws,
                  double *__restrict__ x_out, int *__restrict__ fail_flag,
                  unsigned long long *__restrict__ prof) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    CholSmem &sm = *reinterpret_cast<CholSmem *>(smem_raw);
    cg::grid_group grid = cg::this_grid();
    const int G = gridDim.x, cta = blockIdx.x;
    if (threadIdx.x == 0) sm.fail = 0;
    __syncthreads();
    double *const sL = sm.buf, *const sB = sm.buf + CH_NB * CH_LDS, *const sC = sm.buf + 2 * CH_NB * CH_LDS;
    // block sparsity: nz[i * nb + k] = 1 when the solved panel block L_ik (i > k) has a non-zero entry.
    // Reduced camera systems of large scenes are block-sparse (cameras that share no point give an
    // exactly zero block, and fill-in shows up as real non-zeros); zero blocks skip the panel solve,
    // the trailing tiles they would feed, and their back-substitution term.
    unsigned char *const nz = reinterpret_cast<unsigned char *>(Linv_ws + (size_t)nb * CH_NB * CH_NB);
    // optional phase profile (CTA 0, thread 0; ns): [0] diag factor [1] panel solve [2] barrier 1
    // [3] trailing update [4] barrier 2 [5] back substitution
    const bool profiling = prof != nullptr && cta == 0 && threadIdx.x == 0;
    unsigned long long t_prev = profiling ? globaltimer_ns() : 0ull;
#define CH_TICK(slot)                                              \
    do {                                                           \
        if (profiling) {                                           \
            unsigned long long t_now = globaltimer_ns();           \
            prof[slot] += t_now - t_prev;                          \
            t_prev = t_now;                                        \
        }                                                          \
    } while (0)

    // Diagonal block kb: factor it, commit L_kb,kb to A and its inverse to Linv_ws (CTA 0 only).
    // src != nullptr: the (already updated) block sits in shared memory (row stride CH_LDS, beyond the four
    // views the factor uses) instead of in A.
    auto factor_and_commit = [&](int kb, const double *src = nullptr) {
        __syncthreads();
        if (src) factor_diag_block(sm, src, (size_t)CH_LDS, prof);
        else factor_diag_block(sm, A + (size_t)kb * CH_NB * ld + kb * CH_NB, ld, prof);
        store_block(sL, A, ld, kb * CH_NB, kb * CH_NB);
        for (int e = threadIdx.x; e < CH_NB * CH_NB; e += CH_THREADS) {
            int r = e / CH_NB, c = e - r * CH_NB;
            Linv_ws[(size_t)kb * CH_NB * CH_NB + e] = sB[r * CH_LDS + c];
        }
        if (threadIdx.x == 0 && sm.fail) *fail_flag = 1;
        __syncthreads();
    };
    if (cta == 0) factor_and_commit(0);
    CH_TICK(0);
    grid.sync();
    CH_TICK(2);

    const bool split = G > 1;
    if (split) {
        // ---- critical-path form (G > 1) ----
        // The dependency chain of a blocked Cholesky is trsm(k+1,k) -> syrk(k+1,k+1) -> factor(k+1), once
        // per panel. CTA 0 runs exactly that chain and nothing else, without waiting for anybody inside a
        // panel step: L_kk^-1 is still in its shared memory from the factor it just finished, it solves
        // block (k+1,k), publishes it (`ready` = k+1, release), updates and factors the next diagonal block.
        // CTAs 1 .. G-1 solve the other block-rows of the panel, meet at a barrier of their own, wait for
        // `ready` (set long before) and run the trailing update. One grid-wide barrier per panel remains.
        unsigned *const syncw = reinterpret_cast<unsigned *>(nz + (((size_t)(nb + 1) * nb + 15) & ~(size_t)15));
        unsigned *const bar_ctr = syncw, *const ready = syncw + 1;      // zeroed by chol_solve_launch
        const unsigned P = (unsigned)(G - 1);
        for (int k = 0; k < nb; ++k) {
            if (cta == 0) {
                if (k + 1 < nb) {
                    const int i = k + 1;
                    constexpr int PER = CH_NB * CH_NB / CH_THREADS;       // 9
                    double *const sD = sm.buf + 4 * CH_NB * CH_LDS;       // fifth view: outside the factor's four
                    // both blocks the chain needs are final since the last grid barrier: one L2 round trip
                    double bv[PER], dv[PER];
#pragma unroll
                    for (int u = 0; u < PER; ++u) {
                        const int e = threadIdx.x + u * CH_THREADS, r = e / CH_NB, c = e - r * CH_NB;
                        bv[u] = A[(size_t)(i * CH_NB + r) * ld + k * CH_NB + c];
                        dv[u] = A[(size_t)(i * CH_NB + r) * ld + i * CH_NB + c];
                    }
                    int any = 0;
#pragma unroll
                    for (int u = 0; u < PER; ++u) {
                        const int e = threadIdx.x + u * CH_THREADS, r = e / CH_NB, c = e - r * CH_NB;
                        sC[r * CH_LDS + c] = bv[u];
                        sD[r * CH_LDS + c] = dv[u];
                        any |= !(bv[u] == 0.0);
                    }
                    any = __syncthreads_or(any);
                    if (threadIdx.x == 0) nz[(size_t)i * nb + k] = (unsigned char)(any != 0);
                    if (any) {
                        // sB = L_kk^-1, left there by factor_and_commit(k)
                        if (DMMA) block_xyT_mma<0>(sC, sB, sL); else block_xyT<0>(sC, sB, sL);
                        __syncthreads();
                        store_block(sL, A, ld, i * CH_NB, k * CH_NB);
                        CH_TICK(1);
                        // A_{k+1,k+1} -= L_{k+1,k} L_{k+1,k}^T, kept in shared memory
                        if (DMMA) block_xyT_mma<1>(sL, sL, sD); else block_xyT<1>(sL, sL, sD);
                    }
                    __syncthreads();
                    if (threadIdx.x == 0) {
                        __threadfence();                  // block (k+1,k) and its flag before the announcement; the
                        atomicExch(ready, (unsigned)(k + 1));   // stores drained while the product above ran
                    }
                    factor_and_commit(k + 1, sD);
                    CH_TICK(3);
                }
            } else {
                // ---- phase 1: block-rows k+2 .. nb (nb = right-hand side) over CTAs 1 .. G-1; at the last
                // panel block-row k+1 IS the right-hand side and belongs here too ----
                const int first = (k + 1 < nb ? k + 2 : k + 1) + (cta - 1);
                if (first <= nb) {
                    constexpr int PER = CH_NB * CH_NB / CH_THREADS;       // 9
                    double lv[PER], bv[PER];
                    int i = first;
#pragma unroll
                    for (int u = 0; u < PER; ++u) lv[u] = Linv_ws[(size_t)k * CH_NB * CH_NB + threadIdx.x + u * CH_THREADS];
#pragma unroll
                    for (int u = 0; u < PER; ++u) {
                        const int e = threadIdx.x + u * CH_THREADS, r = e / CH_NB, c = e - r * CH_NB;
                        bv[u] = A[(size_t)(i * CH_NB + r) * ld + k * CH_NB + c];
                    }
#pragma unroll
                    for (int u = 0; u < PER; ++u) {
                        const int e = threadIdx.x + u * CH_THREADS, r = e / CH_NB, c = e - r * CH_NB;
                        sB[r * CH_LDS + c] = lv[u];
                    }
                    for (; i <= nb; i += (int)P) {
                        int any = 0;
#pragma unroll
                        for (int u = 0; u < PER; ++u) {
                            const int e = threadIdx.x + u * CH_THREADS, r = e / CH_NB, c = e - r * CH_NB;
                            sC[r * CH_LDS + c] = bv[u];
                            any |= !(bv[u] == 0.0);
                        }
                        if (i + (int)P <= nb) {
#pragma unroll
                            for (int u = 0; u < PER; ++u) {
                                const int e = threadIdx.x + u * CH_THREADS, r = e / CH_NB, c = e - r * CH_NB;
                                bv[u] = A[(size_t)((i + (int)P) * CH_NB + r) * ld + k * CH_NB + c];
                            }
                        }
                        any = __syncthreads_or(any);
                        if (threadIdx.x == 0) nz[(size_t)i * nb + k] = (unsigned char)(any != 0);
                        if (any) {
                            if (DMMA) block_xyT_mma<0>(sC, sB, sL); else block_xyT<0>(sC, sB, sL);
                            __syncthreads();
                            store_block(sL, A, ld, i * CH_NB, k * CH_NB);
                        }
                        __syncthreads();
                    }
                }
                // barrier of CTAs 1 .. G-1 (monotone counter), then CTA 0's block (k+1,k)
                __syncthreads();
                if (threadIdx.x == 0) {
                    __threadfence();
                    atomicAdd(bar_ctr, 1u);
                    const unsigned target = P * (unsigned)(k + 1);
                    while (*(volatile unsigned *)bar_ctr < target) { }
                    if (k + 1 < nb)
                        while (*(volatile unsigned *)ready < (unsigned)(k + 1)) { }
                    __threadfence();                      // acquire; also drops this SM's stale L1 lines
                }
                __syncthreads();
                // ---- phase 2: trailing update, block (k+1,k+1) excluded (CTA 0 has it) ----
                const int m = nb - 1 - k;
                if (m >= CH_SUPER_MIN) {
                    if (!sparse_tile_phase(sm, A, ld, nz, k, nb, cta, G))
                        supertile_phase<DMMA>(sm, A, ld, nz, k, nb, cta, G, factor_and_commit);
                } else if (m > 0) {
                    const int ntiles = m * (m + 1) / 2 + m;
                    for (int t = cta; t < ntiles; t += (int)P) {          // t = 0 is block (k+1,k+1)
                        int ip, jp;
                        if (t < m * (m + 1) / 2) {
                            ip = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
                            while ((ip + 1) * (ip + 2) / 2 <= t) ++ip;
                            while (ip * (ip + 1) / 2 > t) --ip;
                            jp = t - ip * (ip + 1) / 2;
                        } else {
                            ip = m;
                            jp = t - m * (m + 1) / 2;
                        }
                        const int i = k + 1 + ip, j = k + 1 + jp;
                        __syncthreads();
                        load_block(sL, A, ld, i * CH_NB, k * CH_NB);
                        load_block(sB, A, ld, j * CH_NB, k * CH_NB);
                        load_block(sC, A, ld, i * CH_NB, j * CH_NB);
                        __syncthreads();
                        block_xyT<1>(sL, sB, sC);
                        __syncthreads();
                        store_block(sC, A, ld, i * CH_NB, j * CH_NB);
                    }
                }
            }
            grid.sync();
            CH_TICK(4);
        }
    } else {
        for (int k = 0; k < nb; ++k) {
            // ---- phase 1: panel solve A_ik <- A_ik L_kk^-T with the inverse committed one phase earlier ----
            const int rows_below = nb - k;            // block-rows k+1 .. nb (nb = RHS)
            if (cta < rows_below) {
                // the inverse and the first block of this CTA are requested together (18 independent loads per
                // thread, one L2 round trip) before anything is parked in shared memory
                constexpr int PER = CH_NB * CH_NB / CH_THREADS;       // 9
                double lv[PER], bv[PER];
                int i = k + 1 + cta;
    #pragma unroll
                for (int u = 0; u < PER; ++u) lv[u] = Linv_ws[(size_t)k * CH_NB * CH_NB + threadIdx.x + u * CH_THREADS];
    #pragma unroll
                for (int u = 0; u < PER; ++u) {
                    const int e = threadIdx.x + u * CH_THREADS, r = e / CH_NB, c = e - r * CH_NB;
                    bv[u] = A[(size_t)(i * CH_NB + r) * ld + k * CH_NB + c];
                }
    #pragma unroll
                for (int u = 0; u < PER; ++u) {
                    const int e = threadIdx.x + u * CH_THREADS, r = e / CH_NB, c = e - r * CH_NB;
                    sB[r * CH_LDS + c] = lv[u];
                }
                for (; i <= nb; i += G) {
                    int any = 0;
    #pragma unroll
                    for (int u = 0; u < PER; ++u) {
                        const int e = threadIdx.x + u * CH_THREADS, r = e / CH_NB, c = e - r * CH_NB;
                        sC[r * CH_LDS + c] = bv[u];
                        any |= !(bv[u] == 0.0);
                    }
                    if (i + G <= nb) {                // next block of this CTA (only when rows outnumber CTAs)
    #pragma unroll
                        for (int u = 0; u < PER; ++u) {
                            const int e = threadIdx.x + u * CH_THREADS, r = e / CH_NB, c = e - r * CH_NB;
                            bv[u] = A[(size_t)((i + G) * CH_NB + r) * ld + k * CH_NB + c];
                        }
                    }
                    any = __syncthreads_or(any);
                    if (threadIdx.x == 0) nz[(size_t)i * nb + k] = (unsigned char)(any != 0);
                    if (any) {                        // an exactly zero block stays zero
                        // X = A_ik L^-T : X[r][c] = sum_m A_ik[r][m] Linv[c][m]
                        block_xyT<0>(sC, sB, sL);
                        __syncthreads();
                        store_block(sL, A, ld, i * CH_NB, k * CH_NB);
                    }
                    __syncthreads();                  // sC / sL are rewritten by the next block
                }
            }
            CH_TICK(1);
            grid.sync();
            CH_TICK(2);
            // ---- phase 2: trailing update over tiles (i, j), k < j <= i <= nb, j < nb. Look-ahead: the CTA
            // that updates the tile holding the next diagonal block (tile 0 = CTA 0) factors that block
            // right away, so the factor overlaps the rest of the update instead of heading phase 1. ----
            const int m = nb - 1 - k;                 // remaining diagonal blocks
            if (m >= CH_SUPER_MIN) {
                // 96 x 96 super-tiles over the trailing block rows k+1 .. nb, aligned to even block indices
                if (G == 1 || cta == 0 || !sparse_tile_phase(sm, A, ld, nz, k, nb, cta, G))
                    supertile_phase<DMMA>(sm, A, ld, nz, k, nb, cta, G, factor_and_commit);
            } else if (m > 0) {
                // tiles: for j' in [0, m): i' in [j', m]  (i' == m is the RHS block-row)
                const int ntiles = m * (m + 1) / 2 + m;
                for (int t = st_first_tile(cta, G); t < ntiles; t = st_next_tile(t, cta, G)) {
                    int ip, jp;
                    if (t < m * (m + 1) / 2) {
                        // lower triangle of an m x m block matrix, row-major enumeration
                        ip = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
                        while ((ip + 1) * (ip + 2) / 2 <= t) ++ip;
                        while (ip * (ip + 1) / 2 > t) --ip;
                        jp = t - ip * (ip + 1) / 2;
                    } else {
                        ip = m;
                        jp = t - m * (m + 1) / 2;
                    }
                    const int i = k + 1 + ip, j = k + 1 + jp;
                    __syncthreads();
                    load_block(sL, A, ld, i * CH_NB, k * CH_NB);
                    load_block(sB, A, ld, j * CH_NB, k * CH_NB);
                    load_block(sC, A, ld, i * CH_NB, j * CH_NB);
                    __syncthreads();
                    block_xyT<1>(sL, sB, sC);
                    __syncthreads();
                    store_block(sC, A, ld, i * CH_NB, j * CH_NB);
                    if (t == 0) {
                        CH_TICK(3);
                        factor_and_commit(k + 1);
                        CH_TICK(0);
                    }
                }
            }
            CH_TICK(3);
            grid.sync();
            CH_TICK(4);
        }

    }

    // ---- backward substitution L^T x = y, y = first row of the RHS block-row ----
    const double *yrow = A + (size_t)nb * CH_NB * ld;
    if (nb <= CH_SINGLE_BACK) {
        // small systems: one CTA, block barriers only
        if (cta != 0) return;
        for (int e = threadIdx.x; e < nb * CH_NB; e += CH_THREADS) x_out[e] = yrow[e];
        for (int k = nb - 1; k >= 0; --k) {
            __syncthreads();
            if (threadIdx.x < CH_NB) sm.x[threadIdx.x] = x_out[k * CH_NB + threadIdx.x];
            __syncthreads();
            // x_k = Linv_kk^T y_k: thread (c, part) sums a quarter of the rows, combined through smem
            {
                const double *Li = Linv_ws + (size_t)k * CH_NB * CH_NB;
                const int c = threadIdx.x & 63, part = threadIdx.x >> 6;
                double s = 0.0;
                if (c < CH_NB)
#pragma unroll 4
                    for (int r = part * 12; r < part * 12 + 12; ++r)
                        if (r >= c) s += Li[r * CH_NB + c] * sm.x[r];
                sC[part * 64 + c] = s;
            }
            __syncthreads();
            if (threadIdx.x < CH_NB)
                sB[threadIdx.x] = sC[threadIdx.x] + sC[64 + threadIdx.x] + sC[128 + threadIdx.x] +
                                    sC[192 + threadIdx.x];
            __syncthreads();
            if (threadIdx.x < CH_NB) x_out[k * CH_NB + threadIdx.x] = sB[threadIdx.x];
            // y_j -= sum_r L[k*NB + r][j] x_k[r] for all columns j < k*NB
            for (int j = threadIdx.x; j < k * CH_NB; j += CH_THREADS) {
                double s = 0.0;
#pragma unroll 8
                for (int r = 0; r < CH_NB; ++r) s += A[(size_t)(k * CH_NB + r) * ld + j] * sB[r];
                x_out[j] -= s;
            }
        }
        CH_TICK(5);
        return;
    }
    // (A single-CTA left-looking substitution for block-sparse factors was tried here: one SM cannot pull
    // the ~10 non-zero blocks of a block column from L2 fast enough -- 5.7 us per step against 4.3 us for
    // the grid-wide steps below at C5 -- so the grid-wide form serves sparse and dense factors alike.)
    for (int e = cta * CH_THREADS + threadIdx.x; e < nb * CH_NB; e += G * CH_THREADS) x_out[e] = yrow[e];
    {
        // Everything a step reads except x itself is final before the loop starts: the inverse of the
        // diagonal block and this CTA's first block L_kb of the step are requested one step ahead, before
        // the grid barrier, so a step's chain is load x -> two 48 x 48 mat-vecs -> barrier.
        const int c = threadIdx.x & 63, part = threadIdx.x >> 6;
        double li[12], lb[12];
        int have_b = 0;
        auto prefetch = [&](int k) {
            const double *Li = Linv_ws + (size_t)k * CH_NB * CH_NB;
#pragma unroll
            for (int r = 0; r < 12; ++r) li[r] = (c < CH_NB) ? Li[(part * 12 + r) * CH_NB + c] : 0.0;
            have_b = (cta < k) && nz[(size_t)k * nb + cta];
            if (have_b) {
#pragma unroll
                for (int r = 0; r < 12; ++r)
                    lb[r] = (c < CH_NB) ? A[(size_t)(k * CH_NB + part * 12 + r) * ld + cta * CH_NB + c] : 0.0;
            }
        };
        prefetch(nb - 1);
        grid.sync();
        for (int k = nb - 1; k >= 0; --k) {
            // every CTA: x_k = Linv_kk^T y_k
            __syncthreads();
            if (threadIdx.x < CH_NB) sm.x[threadIdx.x] = x_out[k * CH_NB + threadIdx.x];
            __syncthreads();
            {
                double s = 0.0;
#pragma unroll
                for (int r = 0; r < 12; ++r)
                    if (part * 12 + r >= c) s += li[r] * sm.x[part * 12 + r];
                sC[part * 64 + c] = s;
            }
            __syncthreads();
            if (threadIdx.x < CH_NB)
                sB[threadIdx.x] = sC[threadIdx.x] + sC[64 + threadIdx.x] + sC[128 + threadIdx.x] +
                                    sC[192 + threadIdx.x];
            __syncthreads();
            // y_b -= L_kb^T x_k for b < k: each CTA takes whole column blocks (no atomics)
            for (int b = cta; b < k; b += G) {
                const bool mine = (b == cta);
                if (mine ? !have_b : !nz[(size_t)k * nb + b]) continue;      // zero block (CTA-uniform)
                double s = 0.0;
                if (mine) {
#pragma unroll
                    for (int r = 0; r < 12; ++r) s += lb[r] * sB[part * 12 + r];
                } else if (c < CH_NB) {
#pragma unroll 4
                    for (int r = part * 12; r < part * 12 + 12; ++r)
                        s += A[(size_t)(k * CH_NB + r) * ld + b * CH_NB + c] * sB[r];
                }
                __syncthreads();
                sC[part * 64 + c] = s;
                __syncthreads();
                if (threadIdx.x < CH_NB)
                    x_out[b * CH_NB + threadIdx.x] -= sC[threadIdx.x] + sC[64 + threadIdx.x] +
                                                      sC[128 + threadIdx.x] + sC[192 + threadIdx.x];
            }
            if (k > 0) prefetch(k - 1);
            grid.sync();
            if (cta == 0 && threadIdx.x < CH_NB) x_out[k * CH_NB + threadIdx.x] = sB[threadIdx.x];
        }
    }
    CH_TICK(5);
#undef CH_TICK
}

// diagnostic: cycles of the diagonal-block factor alone (one CTA)
template <int V>
__global__ void __launch_bounds__(CH_THREADS) chol_diag_bench_kernel(const double *A48, unsigned long long *out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    CholSmem &sm = *reinterpret_cast<CholSmem *>(smem_raw);
    factor_diag_block<V>(sm, A48, CH_NB, out);
}

static int g_chol_grid = 0;
static int g_chol_dmma = -1;

static bool chol_use_dmma() {
    if (g_chol_dmma < 0) {
        const char *e = getenv("SFM_CHOL_DFMA");          // development switch: plain DFMA trailing update
        g_chol_dmma = (e && e[0] == '1') ? 0 : 1;
    }
    return g_chol_dmma == 1;
}

int chol_grid_size() {
    if (g_chol_grid) return g_chol_grid;
    int per_sm = 0;
    const void *fn = chol_use_dmma() ? (const void *)chol_solve_kernel<true> : (const void *)chol_solve_kernel<false>;
    if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CholSmem)) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, CH_THREADS, sizeof(CholSmem)) != cudaSuccess ||
        per_sm < 1) {
        (void)cudaGetLastError();
        return 0;
    }
    g_chol_grid = sm_count() * (per_sm > 2 ? 2 : per_sm);
    return g_chol_grid;
}

// Enqueue the factor+solve. A: (npad + NB) x ld; d_Linv: nb*NB*NB doubles; d_x: npad doubles;
// d_fail: int (zeroed by the caller).
int chol_solve_launch(double *A, size_t ld, int npad, double *d_Linv, double *d_x, int *d_fail,
                      cudaStream_t st, unsigned long long *d_prof) {
    int G = chol_grid_size();
    if (G <= 0) return SFM_E_CUDA;
    int nb = npad / CH_NB;
    // no point in more CTAs than the widest phase can use
    int need = nb * (nb + 1) / 2 + nb + 1;
    if (G > need) G = need;
    {   // barrier counter + `ready` of the critical-path form, in the slack behind K11's block flags
        unsigned char *nzp = reinterpret_cast<unsigned char *>(d_Linv + (size_t)nb * CH_NB * CH_NB);
        void *syncw = nzp + (((size_t)(nb + 1) * nb + 15) & ~(size_t)15);
        if (cudaMemsetAsync(syncw, 0, 16, st) != cudaSuccess) return SFM_E_CUDA;
    }
    void *args[] = {&A, &ld, &nb, &d_Linv, &d_x, &d_fail, &d_prof};
    const void *fn = chol_use_dmma() ? (const void *)chol_solve_kernel<true> : (const void *)chol_solve_kernel<false>;
    cudaError_t e = cudaLaunchCooperativeKernel(fn, dim3(G), dim3(CH_THREADS), args, sizeof(CholSmem), st);
    count_launch();
    if (e != cudaSuccess) {
        record_cuda_error(e, "cudaLaunchCooperativeKernel(chol_solve_kernel)");
        return SFM_E_CUDA;
    }
    return SFM_OK;
}

// ---- standalone entry (tests): pack a dense SPD system into the padded layout ----------------
__global__ void chol_pack_kernel(const double *__restrict__ S, const double *__restrict__ g, int n, int npad,
                                 size_t ld, double *__restrict__ A) {
    size_t total = (size_t)(npad + CH_NB) * ld;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        int r = (int)(e / ld), c = (int)(e - (size_t)r * ld);
        double v = 0.0;
        if (r < n && c < n) v = S[(size_t)r * n + c];
        else if (r < npad && r == c) v = 1.0;
        else if (r == npad && c < n) v = g[c];
        A[e] = v;
    }
}

}  // namespace sfm

using namespace sfm;

extern "C" size_t sfm_chol_solve_workspace_bytes(int32_t n) {
    if (n < 1) return 0;
    size_t npad = (size_t)ceil_div(n, CH_NB) * CH_NB;
    size_t bytes = (npad + CH_NB) * npad * sizeof(double);              // A
    bytes += (npad / CH_NB) * (size_t)CH_NB * CH_NB * sizeof(double);    // Linv
    bytes += align_up((npad / CH_NB + 1) * (npad / CH_NB) + 32, 256);   // block non-zero flags + K11's sync words
    bytes += npad * sizeof(double) + 256;                                // x, flag
    return bytes;
}

static int chol_solve_impl(const double *d_S, const double *d_g, int32_t n, double *d_x, int32_t *d_fail,
                           void *d_ws, size_t ws_bytes, void *stream, unsigned long long *d_prof) {
    if (!d_S || !d_g || !d_x || !d_fail || !d_ws || n < 1) return SFM_E_INVALID;
    if (ws_bytes < sfm_chol_solve_workspace_bytes(n)) return SFM_E_WORKSPACE;
    cudaStream_t st = as_stream(stream);
    int npad = ceil_div(n, CH_NB) * CH_NB;
    size_t ld = (size_t)npad;
    double *A = reinterpret_cast<double *>(d_ws);
    double *Linv = A + (size_t)(npad + CH_NB) * ld;
    double *x = Linv + (size_t)(npad / CH_NB) * CH_NB * CH_NB + align_up((size_t)(npad / CH_NB + 1) * (npad / CH_NB) + 32, 256) / sizeof(double);
    int *flag = reinterpret_cast<int *>(x + npad);
    SFM_CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(int), st));
    chol_pack_kernel<<<sm_count() * 4, 256, 0, st>>>(d_S, d_g, n, npad, ld, A);
    SFM_LAUNCH_CHECK();
    int rc = chol_solve_launch(A, ld, npad, Linv, x, flag, st, d_prof);
    if (rc != SFM_OK) return rc;
    SFM_CUDA_TRY(cudaMemcpyAsync(d_x, x, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, st));
    SFM_CUDA_TRY(cudaMemcpyAsync(d_fail, flag, sizeof(int), cudaMemcpyDeviceToDevice, st));
    return SFM_OK;
}

extern "C" int sfm_chol_solve(const double *d_S, const double *d_g, int32_t n, double *d_x,
                              int32_t *d_fail, void *d_ws, size_t ws_bytes, void *stream) {
    return chol_solve_impl(d_S, d_g, n, d_x, d_fail, d_ws, ws_bytes, stream, nullptr);
}

extern "C" int sfm_chol_solve_profiled(const double *d_S, const double *d_g, int32_t n, double *d_x,
                                       int32_t *d_fail, void *d_ws, size_t ws_bytes, uint64_t *d_phase_ns,
                                       void *stream) {
    if (!d_phase_ns) return SFM_E_INVALID;
    return chol_solve_impl(d_S, d_g, n, d_x, d_fail, d_ws, ws_bytes, stream,
                           reinterpret_cast<unsigned long long *>(d_phase_ns));
}

extern "C" int sfm_chol_diag_bench(const double *d_A48, int32_t variant, uint64_t *d_out8, void *stream) {
    if (!d_A48 || !d_out8 || variant < 0 || variant > 5) return SFM_E_INVALID;
    cudaStream_t st = as_stream(stream);
    if (chol_grid_size() <= 0) return SFM_E_CUDA;
    void (*fn)(const double *, unsigned long long *) =
        variant == 0 ? chol_diag_bench_kernel<0> : variant == 1 ? chol_diag_bench_kernel<1> :
        variant == 2 ? chol_diag_bench_kernel<2> : variant == 3 ? chol_diag_bench_kernel<3> :
        variant == 4 ? chol_diag_bench_kernel<4> : chol_diag_bench_kernel<5>;
    SFM_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CholSmem)));
    fn<<<1, CH_THREADS, sizeof(CholSmem), st>>>(d_A48, reinterpret_cast<unsigned long long *>(d_out8));
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}
