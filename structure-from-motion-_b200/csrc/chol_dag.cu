// chol_dag.cu -- K11s: structure-aware factorisation and solve of the reduced camera system S dc = g_red.
//
// The dense kernel in chol.cu walks all panels of S one after the other: 125 sequential panel steps of
// ~15 us at C5 (1,000 cameras), issue slots 7 % busy -- a latency chain, while S there is a cyclic band
// (cameras share points only with their ~40 neighbours on the ring). Here the block structure of S is
// analysed once per problem (chol_plan.cpp): nested-dissection order on the 48 x 48 block graph, symbolic
// factorisation, and a LEFT-LOOKING task graph -- one task per non-zero block of L:
//
//   diag task D(j):  C = A_jj - sum_t L_jt L_jt^T  ->  L_jj = chol(C), L_jj^-1;   y_j = L_jj^-1 (g_j - sum_t L_jt y_t)
//   off  task B(i,j): C = A_ij - sum_t L_it L_jt^T ->  L_ij = C L_jj^-T
//   back task X(j):  x_j = L_jj^-T (y_j - sum_i L_ij^T x_i)
//
// statically list-scheduled over one persistent CTA per SM. Every block of L is written once, into a compact
// block array, and announced through a release flag (value = the epoch of this launch, so flags are never
// reset); a task spins (acquire) on the flags of the operands of each product in turn, so everything but
// the last product of a task is accumulated before the block it really waits for exists. Independent
// subtrees of the elimination tree run concurrently on different SMs; the dependency chain is the height
// of the elimination tree (~30 blocks at C5) instead of the number of panels (125). S is only read.
//
// In the reference this solve is inside MINPACK's dense pivoted QR (BundleAdjustment.py:205-212 ->
// scipy least_squares(method='lm')); see chol.cu for the dense kernel this one takes over from.
#include "chol_blocks.cuh"
#include "chol_plan.h"
#include <cooperative_groups.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <new>
#include <vector>

namespace cg = cooperative_groups;

namespace sfm {

constexpr int DG_VIEW = CH_NB * CH_LDS;                    // doubles per shared-memory block view
constexpr int DG_BLK = CH_NB * CH_NB;                      // doubles per compact block
constexpr unsigned long long DG_TIMEOUT_NS = 400ull * 1000 * 1000;   // a wait this long is a scheduling bug

constexpr int DG_VIEWS = 11;                               // shared-memory block views of a backward task (220 of 227 KB)

struct DagSmem {
    CholSmem c;                 // eight block views + the diagonal factor's scratch
    double more[(DG_VIEWS - 8) * DG_VIEW];   // views 8 .. of the backward tasks
    double v[64];               // running right-hand side of a diag task
    double yt[2][64];           // y_t of the staged product
    double red[2 * 4 * 64];
    double xs[DG_VIEWS * CH_NB];   // x_i of the staged blocks of a backward task
    int nready, abort, pre;
};

struct DagDev {
    int nb, nnzL, nblk, n_fwd;  // nblk = nnzL + scratch blocks of the partial tasks
    const DagTask *tasks;
    const DagEntry *entries;
    const int32_t *fwd_ptr, *fwd_list, *bwd_ptr, *bwd_list, *perm;
    double *Lc, *Linv, *y, *xe, *pvec;
    unsigned long long *trace;  // optional [n_tasks][4] ns: task start, operands complete, end, SM id (SFM_CHOL_DAG_TRACE)
    unsigned *flags;            // [nblk] off-diagonal + scratch blocks | [nb] diagonal | [nb] x | [1] abort
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(unsigned *p, unsigned v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void cp_async_16(double *smem_dst, const double *gmem_src) {
    unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}

// spin until *f == epoch (one thread); gives up on the abort word or after DG_TIMEOUT_NS
__device__ __forceinline__ bool spin_flag(const unsigned *f, unsigned epoch, unsigned *abort_w) {
    unsigned long long t0 = 0;
    int it = 0;
    while (ld_acquire_u32(f) != epoch) {
        if ((++it & 31) == 0) {
            if (*(volatile unsigned *)abort_w == epoch) return false;
            const unsigned long long now = globaltimer_ns();
            if (!t0) t0 = now;
            else if (now - t0 > DG_TIMEOUT_NS) { atomicExch(abort_w, epoch); return false; }
        }
    }
    return true;
}

// all threads: wait for up to two flags (threads 0 and 1 spin); false = aborted
__device__ __forceinline__ bool wait_flags(DagSmem &sm, const unsigned *f0, const unsigned *f1, unsigned epoch,
                                           unsigned *abort_w) {
    if (threadIdx.x < 2) {
        const unsigned *f = threadIdx.x ? f1 : f0;
        if (f && !spin_flag(f, epoch, abort_w)) sm.abort = 1;
    }
    __syncthreads();
    return !sm.abort;
}

// compact block (row-major 48 x 48) -> padded shared-memory view, 16-byte cp.async through L2
__device__ __forceinline__ void issue_block(double *s, const double *blk) {
    for (int e = threadIdx.x; e < DG_BLK / 2; e += CH_THREADS) {
        const int r = e / (CH_NB / 2), c2 = e - r * (CH_NB / 2);
        cp_async_16(&s[r * CH_LDS + 2 * c2], &blk[r * CH_NB + 2 * c2]);
    }
}

__global__ void __launch_bounds__(CH_THREADS)
chol_dag_kernel(const double *__restrict__ S, size_t ld, DagDev pv, unsigned epoch, double *__restrict__ x_out,
                int *__restrict__ fail_flag) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    DagSmem &sm = *reinterpret_cast<DagSmem *>(smem_raw);
    cg::grid_group grid = cg::this_grid();
    const int cta = blockIdx.x, tid = threadIdx.x;
    const int nb = pv.nb;
    unsigned *const f_off = pv.flags, *const f_diag = pv.flags + pv.nblk, *const f_x = f_diag + nb;
    unsigned *const abort_w = f_x + nb;
    if (tid == 0) { sm.c.fail = 0; sm.abort = 0; }
    __syncthreads();
    double *const V0 = sm.c.buf;
    double *const VC = V0 + 4 * DG_VIEW, *const VI = V0 + 5 * DG_VIEW, *const VO = V0 + 6 * DG_VIEW;
    const double *const rhs = S + (size_t)nb * CH_NB * ld;

    // ------------------------------------------------------------------ forward: factor + y
    for (int q = pv.fwd_ptr[cta]; q < pv.fwd_ptr[cta + 1] && !sm.abort; ++q) {
        const int task_id = pv.fwd_list[q];
        const DagTask t = pv.tasks[task_id];
        const int n = t.end - t.beg;
        const DagEntry *ent = pv.entries + t.beg;
        const bool diag = t.type == DAG_DIAG;
        if (pv.trace && tid == 0) pv.trace[4 * (size_t)task_id] = globaltimer_ns();
        const bool rhs_part = diag || (t.type == DAG_PART && t.i == 1);     // carries a right-hand-side piece
        __syncthreads();                                    // views of the previous task are dead
        // which leading products have their operands already? (one flag pair per thread, checked once)
        if (tid == 0) sm.nready = n;
        __syncthreads();
        for (int p = tid; p < n; p += CH_THREADS) {
            const DagEntry e = ent[p];
            const bool up = ld_acquire_u32(f_off + e.a) == epoch &&
                            (e.b < 0 || e.b == e.a || ld_acquire_u32(f_off + e.b) == epoch);
            if (!up) atomicMin(&sm.nready, p);
        }
        // source block -> VC
        if (t.src_r >= 0) {
            const double *A = S + (size_t)t.src_r * CH_NB * ld + (size_t)t.src_c * CH_NB;
            for (int e = tid; e < DG_BLK; e += CH_THREADS) {
                const int r = e / CH_NB, c = e - r * CH_NB;
                const double a = __ldcg(&A[(size_t)r * ld + c]);
                if (t.src_T) VC[c * CH_LDS + r] = a; else VC[r * CH_LDS + c] = a;
            }
        } else {
            for (int e = tid; e < DG_BLK; e += CH_THREADS) {
                const int r = e / CH_NB, c = e - r * CH_NB;
                VC[r * CH_LDS + c] = 0.0;
            }
        }
        double vr = 0.0;                                    // threads 192..239: entry r of g_j - sum L_jt y_t
        const int vrow = tid - 192;
        if (diag && vrow >= 0 && vrow < CH_NB) vr = __ldcg(&rhs[(size_t)t.src_r * CH_NB + vrow]);
        __syncthreads();
        const int nready = sm.nready;

        auto issue = [&](int p, int s) {
            const DagEntry e = ent[p];
            issue_block(V0 + (2 * s) * DG_VIEW, pv.Lc + (size_t)e.a * DG_BLK);
            if (e.b >= 0 && e.b != e.a) issue_block(V0 + (2 * s + 1) * DG_VIEW, pv.Lc + (size_t)e.b * DG_BLK);
            if (rhs_part && vrow >= 0 && vrow < CH_NB)      // y_t of a product, the vector piece of a partial sum
                sm.yt[s][vrow] = __ldcg(e.b >= 0 ? &pv.y[(size_t)e.t * CH_NB + vrow] : &pv.pvec[(size_t)e.t * CH_NB + vrow]);
            cp_async_commit();
        };
        auto flags_of = [&](int p, const unsigned *&a, const unsigned *&b) {
            const DagEntry e = ent[p];
            a = f_off + e.a;
            b = (e.b >= 0 && e.b != e.a) ? f_off + e.b : nullptr;
        };
        bool ok = true;
        if (n > 0) {
            if (nready < 1) {
                const unsigned *a, *b;
                flags_of(0, a, b);
                ok = wait_flags(sm, a, b, epoch, abort_w);
            }
            if (ok) issue(0, 0);
        }
        for (int p = 0; p < n && ok; ++p) {
            const int s = p & 1;
            bool pre = false;
            if (p + 1 < n) {
                if (p + 1 < nready) pre = true;
                else {
                    // one look at the next product's flags: prefetch it if they are up, else compute first
                    if (tid == 0) {
                        const DagEntry e = ent[p + 1];
                        sm.pre = ld_acquire_u32(f_off + e.a) == epoch &&
                                 (e.b < 0 || e.b == e.a || ld_acquire_u32(f_off + e.b) == epoch);
                    }
                    __syncthreads();
                    pre = sm.pre != 0;
                }
                if (pre) issue(p + 1, s ^ 1);
            }
            if (pre) cp_async_wait<1>(); else cp_async_wait<0>();
            __syncthreads();
            const double *X = V0 + (2 * s) * DG_VIEW;
            if (ent[p].b < 0) {                             // add the block of a partial task
                for (int e = tid; e < DG_BLK; e += CH_THREADS) {
                    const int r = e / CH_NB, c = e - r * CH_NB;
                    VC[r * CH_LDS + c] += X[r * CH_LDS + c];
                }
                if (rhs_part && vrow >= 0 && vrow < CH_NB) vr += sm.yt[s][vrow];
            } else {
                const double *Y = ent[p].b != ent[p].a ? X + DG_VIEW : X;
                block_xyT_mma<1>(X, Y, VC);                 // warps 0..5
                if (rhs_part && vrow >= 0 && vrow < CH_NB) {    // warps 6, 7: right-hand side
                    double a = 0.0;
#pragma unroll 8
                    for (int c = 0; c < CH_NB; ++c) a = fma(X[vrow * CH_LDS + c], sm.yt[s][c], a);
                    vr -= a;
                }
            }
            __syncthreads();
            if (p + 1 < n && !pre) {
                const unsigned *a, *b;
                flags_of(p + 1, a, b);
                ok = wait_flags(sm, a, b, epoch, abort_w);
                if (ok) issue(p + 1, s ^ 1);
            }
        }
        if (!ok) break;
        if (pv.trace && tid == 0) pv.trace[4 * (size_t)task_id + 1] = globaltimer_ns();
        if (diag) {
            if (vrow >= 0 && vrow < CH_NB) sm.v[vrow] = vr;
            __syncthreads();
            factor_diag_block(sm.c, VC, (size_t)CH_LDS);    // V0 = L_jj, V0 + VIEW = L_jj^-1
            const double *Li = V0 + DG_VIEW;
            double *dst = pv.Linv + (size_t)t.j * DG_BLK;
            for (int e = tid; e < DG_BLK; e += CH_THREADS) {
                const int r = e / CH_NB, c = e - r * CH_NB;
                dst[e] = Li[r * CH_LDS + c];
            }
            if (tid < CH_NB) {
                double a = 0.0;
                for (int c = 0; c <= tid; ++c) a = fma(Li[tid * CH_LDS + c], sm.v[c], a);
                pv.y[(size_t)t.j * CH_NB + tid] = a;
            }
            __syncthreads();
            if (tid == 0) {
                if (sm.c.fail) *fail_flag = 1;
                __threadfence();
                st_release_u32(f_diag + t.j, epoch);
            }
        } else if (t.type == DAG_PART) {
            double *dst = pv.Lc + (size_t)t.dst * DG_BLK;
            for (int e = tid; e < DG_BLK; e += CH_THREADS) {
                const int r = e / CH_NB, c = e - r * CH_NB;
                dst[e] = VC[r * CH_LDS + c];
            }
            if (rhs_part && vrow >= 0 && vrow < CH_NB) pv.pvec[(size_t)(t.dst - pv.nnzL) * CH_NB + vrow] = vr;
            __syncthreads();
            if (tid == 0) {
                __threadfence();
                st_release_u32(f_off + t.dst, epoch);
            }
        } else {
            if (!wait_flags(sm, f_diag + t.j, nullptr, epoch, abort_w)) break;
            const double *src = pv.Linv + (size_t)t.j * DG_BLK;
            for (int e = tid; e < DG_BLK; e += CH_THREADS) {
                const int r = e / CH_NB, c = e - r * CH_NB;
                VI[r * CH_LDS + c] = __ldcg(&src[e]);
            }
            __syncthreads();
            block_xyT_mma<0>(VC, VI, VO);                   // L_ij = C L_jj^-T
            __syncthreads();
            double *dst = pv.Lc + (size_t)t.dst * DG_BLK;
            for (int e = tid; e < DG_BLK; e += CH_THREADS) {
                const int r = e / CH_NB, c = e - r * CH_NB;
                dst[e] = VO[r * CH_LDS + c];
            }
            __syncthreads();
            if (tid == 0) {
                __threadfence();
                st_release_u32(f_off + t.dst, epoch);
            }
        }
        if (pv.trace && tid == 0) {
            pv.trace[4 * (size_t)task_id + 2] = globaltimer_ns();
            pv.trace[4 * (size_t)task_id + 3] = (unsigned long long)cta;
        }
    }
    cp_async_wait<0>();
    grid.sync();
    if (*(volatile unsigned *)abort_w == epoch) {
        if (cta == 0 && tid == 0) *fail_flag = 2;
        return;
    }

    // ------------------------------------------------------------------ backward: L^T x = y
    // x_j = L_jj^-T (y_j - sum_i L_ij^T x_i). L is final, so the blocks of a column stream into the eight
    // shared-memory views BEFORE the task waits for the x_i it needs; after the wait only the x_i themselves
    // (one L2 round trip) and shared-memory mat-vecs remain on the chain root -> leaves.
    const int c = tid & 63, part = tid >> 6;
    constexpr int BK = DG_VIEWS;                              // blocks staged per batch
    auto bview = [&](int e) { return e < 8 ? V0 + e * DG_VIEW : sm.more + (e - 8) * DG_VIEW; };
    for (int q = pv.bwd_ptr[cta]; q < pv.bwd_ptr[cta + 1]; ++q) {
        const int task_id = pv.bwd_list[q];
        const DagTask t = pv.tasks[task_id];
        const int n = t.end - t.beg;
        const DagEntry *ent = pv.entries + t.beg;
        if (pv.trace && tid == 0) pv.trace[4 * (size_t)task_id] = globaltimer_ns();
        double li[12];
        const double *Li = pv.Linv + (size_t)t.j * DG_BLK;
#pragma unroll
        for (int r = 0; r < 12; ++r) li[r] = (c < CH_NB) ? __ldcg(&Li[(part * 12 + r) * CH_NB + c]) : 0.0;
        double w = (part == 0 && c < CH_NB) ? __ldcg(&pv.y[(size_t)t.j * CH_NB + c]) : 0.0;
        // entries ascend in row position: entry 0 is the parent, whose x arrives last. Batches (columns longer
        // than the views) are taken from the far end, each waiting only for its own x_i.
        for (int e0 = ((n - 1) / BK) * BK; e0 >= 0 && n > 0; e0 -= BK) {
            const int ne = min(BK, n - e0);
            __syncthreads();                                  // views and xs of the previous batch / task are dead
            for (int e = 0; e < ne; ++e) issue_block(bview(e), pv.Lc + (size_t)ent[e0 + e].a * DG_BLK);
            cp_async_commit();
            for (int e = tid; e < ne; e += CH_THREADS)
                if (!spin_flag(f_x + ent[e0 + e].b, epoch, abort_w)) sm.abort = 1;
            __syncthreads();
            if (sm.abort) break;
            if (pv.trace && tid == 0 && e0 == 0) pv.trace[4 * (size_t)task_id + 1] = globaltimer_ns();
            for (int k = tid; k < ne * CH_NB; k += CH_THREADS) {
                const int e = k / CH_NB, r = k - e * CH_NB;
                sm.xs[k] = __ldcg(&pv.xe[(size_t)ent[e0 + e].b * CH_NB + r]);
            }
            cp_async_wait<0>();
            __syncthreads();
            if (c < CH_NB) {
                double a0 = 0.0, a1 = 0.0;
                for (int e = 0; e < ne; ++e) {
                    const double *L = bview(e) + (part * 12) * CH_LDS + c;
                    const double *xi = sm.xs + e * CH_NB + part * 12;
#pragma unroll
                    for (int r = 0; r < 12; r += 2) {
                        a0 = fma(L[r * CH_LDS], xi[r], a0);
                        a1 = fma(L[(r + 1) * CH_LDS], xi[r + 1], a1);
                    }
                }
                w -= a0 + a1;
            }
        }
        if (sm.abort) break;
        if (pv.trace && tid == 0 && n == 0) pv.trace[4 * (size_t)task_id + 1] = globaltimer_ns();
        __syncthreads();
        sm.red[part * 64 + c] = w;
        __syncthreads();
        if (tid < CH_NB) sm.v[tid] = sm.red[tid] + sm.red[64 + tid] + sm.red[128 + tid] + sm.red[192 + tid];
        __syncthreads();
        {
            double a = 0.0;
#pragma unroll
            for (int r = 0; r < 12; ++r)
                if (part * 12 + r >= c) a = fma(li[r], sm.v[part * 12 + r], a);
            sm.red[256 + part * 64 + c] = a;
        }
        __syncthreads();
        if (tid < CH_NB) {
            const double x = sm.red[256 + tid] + sm.red[256 + 64 + tid] + sm.red[256 + 128 + tid] + sm.red[256 + 192 + tid];
            pv.xe[(size_t)t.j * CH_NB + tid] = x;
            x_out[(size_t)t.src_r * CH_NB + tid] = x;
        }
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            st_release_u32(f_x + t.j, epoch);
            if (pv.trace) {
                pv.trace[4 * (size_t)task_id + 2] = globaltimer_ns();
                pv.trace[4 * (size_t)task_id + 3] = (unsigned long long)cta;
            }
        }
    }
    if (sm.abort && tid == 0) *fail_flag = 2;
}

// pattern of the lower block triangle from a dense lower-triangular S (tests / standalone entry)
__global__ void chol_dag_pattern_kernel(const double *__restrict__ S, int n, int nb, unsigned char *__restrict__ pat) {
    const int bi = blockIdx.y, bj = blockIdx.x;
    if (bj > bi) return;
    int any = 0;
    for (int e = threadIdx.x; e < DG_BLK; e += blockDim.x) {
        const int r = bi * CH_NB + e / CH_NB, cc = bj * CH_NB + e % CH_NB;
        if (r < n && cc < n && cc <= r && S[(size_t)r * n + cc] != 0.0) any = 1;
    }
    any = __syncthreads_or(any);
    if (threadIdx.x == 0) pat[(size_t)bi * nb + bj] = (unsigned char)(any || bi == bj);
}

// ---- plan object ---------------------------------------------------------------------------------
struct CholDagPlan {
    CholDagHost h;
    DagDev d{};
    void *blob = nullptr;        // one device allocation for everything
    size_t blob_bytes = 0;
    unsigned epoch = 0;
    int G = 0;
    uint32_t *d_cells = nullptr;   // (block row << 16 | block col) of every cell of the lower-triangle pattern (in blob)
    int n_cells = 0;
};

static int g_dag_grid = 0;

int chol_dag_grid_size() {
    if (g_dag_grid) return g_dag_grid;
    int per_sm = 0;
    if (cudaFuncSetAttribute(chol_dag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(DagSmem)) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_dag_kernel, CH_THREADS, sizeof(DagSmem)) != cudaSuccess ||
        per_sm < 1) {
        (void)cudaGetLastError();
        return 0;
    }
    g_dag_grid = sm_count();
    return g_dag_grid;
}

void chol_dag_destroy(CholDagPlan *p, cudaStream_t st) {
    if (!p) return;
    if (p->blob) cudaFreeAsync(p->blob, st);
    delete p;
}

// Build the plan for a block pattern (host array, nb x nb, lower triangle) and upload it. order_mode as in
// chol_dag_plan. Returns nullptr when the structure is not worth it (or on failure; *rc says which).
CholDagPlan *chol_dag_create(const uint8_t *h_pattern, int nb, int order_mode, cudaStream_t st, int *rc) {
    *rc = SFM_OK;
    const int G = chol_dag_grid_size();
    if (G <= 0) { *rc = SFM_E_CUDA; return nullptr; }
    CholDagPlan *p = new (std::nothrow) CholDagPlan();
    if (!p) { *rc = SFM_E_INVALID; return nullptr; }
    if (chol_dag_plan(h_pattern, nb, G, order_mode, p->h) != 0) { delete p; return nullptr; }
    p->G = G;
    const CholDagHost &h = p->h;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align_up(bytes, 256); return o; };
    const size_t o_tasks = take(sizeof(DagTask) * h.tasks.size());
    const size_t o_ent = take(sizeof(DagEntry) * std::max<size_t>(h.entries.size(), 1));
    const size_t o_fp = take(sizeof(int32_t) * h.fwd_ptr.size());
    const size_t o_fl = take(sizeof(int32_t) * std::max<size_t>(h.fwd_list.size(), 1));
    const size_t o_bp = take(sizeof(int32_t) * h.bwd_ptr.size());
    const size_t o_bl = take(sizeof(int32_t) * std::max<size_t>(h.bwd_list.size(), 1));
    const size_t o_perm = take(sizeof(int32_t) * nb);
    const size_t nblk = (size_t)h.nnzL + h.n_part;
    const size_t o_flags = take(sizeof(unsigned) * (nblk + 2 * nb + 4));
    const size_t o_Linv = take(sizeof(double) * (size_t)nb * DG_BLK);
    const size_t o_y = take(sizeof(double) * (size_t)nb * CH_NB);
    const size_t o_xe = take(sizeof(double) * (size_t)nb * CH_NB);
    const size_t o_pvec = take(sizeof(double) * std::max<size_t>((size_t)h.n_part, 1) * CH_NB);
    const char *trace_path = getenv("SFM_CHOL_DAG_TRACE");
    const size_t o_trace = take(trace_path ? sizeof(unsigned long long) * 4 * h.tasks.size() : 0);
    std::vector<uint32_t> cells;
    for (int i = 0; i < nb; ++i)
        for (int j = 0; j <= i; ++j)
            if (i == j || h_pattern[(size_t)i * nb + j]) cells.push_back(((uint32_t)i << 16) | (uint32_t)j);
    const size_t o_cells = take(sizeof(uint32_t) * cells.size());
    const size_t o_Lc = take(sizeof(double) * std::max<size_t>(nblk, 1) * DG_BLK);
    p->blob_bytes = off;
    if (cudaMallocAsync(&p->blob, off, st) != cudaSuccess) {
        record_cuda_error(cudaGetLastError(), "cudaMallocAsync(chol dag plan)");
        delete p;
        *rc = SFM_E_CUDA;
        return nullptr;
    }
    unsigned char *base = reinterpret_cast<unsigned char *>(p->blob);
    auto up = [&](size_t o, const void *src, size_t bytes) {
        return bytes == 0 || cudaMemcpyAsync(base + o, src, bytes, cudaMemcpyHostToDevice, st) == cudaSuccess;
    };
    bool ok = up(o_tasks, h.tasks.data(), sizeof(DagTask) * h.tasks.size()) &&
              up(o_ent, h.entries.data(), sizeof(DagEntry) * h.entries.size()) &&
              up(o_fp, h.fwd_ptr.data(), sizeof(int32_t) * h.fwd_ptr.size()) &&
              up(o_fl, h.fwd_list.data(), sizeof(int32_t) * h.fwd_list.size()) &&
              up(o_bp, h.bwd_ptr.data(), sizeof(int32_t) * h.bwd_ptr.size()) &&
              up(o_bl, h.bwd_list.data(), sizeof(int32_t) * h.bwd_list.size()) &&
              up(o_perm, h.perm.data(), sizeof(int32_t) * nb) &&
              up(o_cells, cells.data(), sizeof(uint32_t) * cells.size()) &&
              cudaMemsetAsync(base + o_flags, 0, sizeof(unsigned) * (nblk + 2 * nb + 4), st) == cudaSuccess &&
              cudaStreamSynchronize(st) == cudaSuccess;        // the host vectors must outlive the copies
    if (!ok) {
        record_cuda_error(cudaGetLastError(), "upload(chol dag plan)");
        cudaFreeAsync(p->blob, st);
        delete p;
        *rc = SFM_E_CUDA;
        return nullptr;
    }
    DagDev &d = p->d;
    d.nb = nb; d.nnzL = h.nnzL; d.nblk = (int)nblk; d.n_fwd = h.n_fwd;
    d.tasks = reinterpret_cast<const DagTask *>(base + o_tasks);
    d.entries = reinterpret_cast<const DagEntry *>(base + o_ent);
    d.fwd_ptr = reinterpret_cast<const int32_t *>(base + o_fp);
    d.fwd_list = reinterpret_cast<const int32_t *>(base + o_fl);
    d.bwd_ptr = reinterpret_cast<const int32_t *>(base + o_bp);
    d.bwd_list = reinterpret_cast<const int32_t *>(base + o_bl);
    d.perm = reinterpret_cast<const int32_t *>(base + o_perm);
    d.flags = reinterpret_cast<unsigned *>(base + o_flags);
    d.Linv = reinterpret_cast<double *>(base + o_Linv);
    d.y = reinterpret_cast<double *>(base + o_y);
    d.xe = reinterpret_cast<double *>(base + o_xe);
    d.pvec = reinterpret_cast<double *>(base + o_pvec);
    p->d_cells = reinterpret_cast<uint32_t *>(base + o_cells);
    p->n_cells = (int)cells.size();
    d.trace = trace_path ? reinterpret_cast<unsigned long long *>(base + o_trace) : nullptr;
    d.Lc = reinterpret_cast<double *>(base + o_Lc);
    return p;
}

// Enqueue factor + solve of the system held in the padded layout of chol.cu (A: (npad + NB) x ld, lower
// triangle, right-hand side in row npad). d_x [npad]; d_fail: int (zeroed by the caller; 1 = not positive
// definite, 2 = internal wait timed out).
int chol_dag_launch(CholDagPlan *p, const double *A, size_t ld, int npad, double *d_x, int *d_fail, cudaStream_t st) {
    if (!p || npad != p->h.nb * CH_NB) return SFM_E_INVALID;
    unsigned epoch = ++p->epoch;
    if (epoch == 0) epoch = ++p->epoch;                  // 0 is the reset value of the flags
    DagDev dv = p->d;
    void *args[] = {(void *)&A, (void *)&ld, (void *)&dv, (void *)&epoch, (void *)&d_x, (void *)&d_fail};
    cudaError_t e = cudaLaunchCooperativeKernel((const void *)chol_dag_kernel, dim3(p->G), dim3(CH_THREADS), args,
                                                sizeof(DagSmem), st);
    count_launch();
    if (e != cudaSuccess) {
        record_cuda_error(e, "cudaLaunchCooperativeKernel(chol_dag_kernel)");
        return SFM_E_CUDA;
    }
    return SFM_OK;
}

const CholDagHost &chol_dag_host(const CholDagPlan *p) { return p->h; }

// Zero exactly what the factorisation reads: the 48 x 48 cells of the pattern and the right-hand-side row. (A reduced
// camera system of 1,000 cameras is 295 MB as a dense array and 24 MB in its pattern: the memset of the whole array
// cost more than the Z pass of the Schur complement.)
__global__ void __launch_bounds__(256) chol_dag_zero_kernel(double *__restrict__ A, size_t ld, int nb,
                                                            const uint32_t *__restrict__ cells, int n_cells) {
    const int b = blockIdx.x;
    if (b < n_cells) {
        const uint32_t c = cells[b];
        double *blk = A + (size_t)(c >> 16) * CH_NB * ld + (size_t)(c & 0xffffu) * CH_NB;
        for (int e = threadIdx.x; e < DG_BLK / 2; e += 256) {
            const int r = e / (CH_NB / 2), c2 = e - r * (CH_NB / 2);
            reinterpret_cast<double2 *>(blk + (size_t)r * ld)[c2] = make_double2(0.0, 0.0);
        }
    } else {
        double *rhs = A + (size_t)nb * CH_NB * ld;
        for (size_t e = (size_t)(b - n_cells) * 256 + threadIdx.x; e < (size_t)nb * CH_NB; e += (size_t)(gridDim.x - n_cells) * 256)
            rhs[e] = 0.0;
    }
}

int chol_dag_zero_system(CholDagPlan *p, double *A, size_t ld, cudaStream_t st) {
    if (!p) return SFM_E_INVALID;
    chol_dag_zero_kernel<<<p->n_cells + 8, 256, 0, st>>>(A, ld, p->h.nb, p->d_cells, p->n_cells);
    SFM_LAUNCH_CHECK();
    return SFM_OK;
}

// SFM_CHOL_DAG_TRACE=<file>: per-task timestamps of the last launch, written as
// int64 n_tasks | tasks [n_tasks][12] int32 | trace [n_tasks][4] uint64 (diagnostic, tools/chol_dag_trace.py)
static void chol_dag_dump_trace(const CholDagPlan *p, cudaStream_t st) {
    const char *path = getenv("SFM_CHOL_DAG_TRACE");
    if (!path || !p->d.trace) return;
    const size_t nt = p->h.tasks.size();
    std::vector<unsigned long long> tr(4 * nt);
    if (cudaStreamSynchronize(st) != cudaSuccess ||
        cudaMemcpy(tr.data(), p->d.trace, sizeof(unsigned long long) * 4 * nt, cudaMemcpyDeviceToHost) != cudaSuccess) return;
    FILE *f = fopen(path, "wb");
    if (!f) return;
    const int64_t n64 = (int64_t)nt;
    fwrite(&n64, sizeof(n64), 1, f);
    fwrite(p->h.tasks.data(), sizeof(DagTask), nt, f);
    fwrite(tr.data(), sizeof(unsigned long long), 4 * nt, f);
    fclose(f);
}

// dense S [n,n] + g [n] -> the padded layout (identity on the padding diagonal, g in row npad)
__global__ void chol_dag_pack_kernel(const double *__restrict__ S, const double *__restrict__ g, int n, int npad,
                                     size_t ld, double *__restrict__ A) {
    const size_t total = (size_t)(npad + CH_NB) * ld;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int r = (int)(e / ld), c = (int)(e - (size_t)r * ld);
        double v = 0.0;
        if (r < n && c < n) v = S[(size_t)r * n + c];
        else if (r < npad && r == c) v = 1.0;
        else if (r == npad && c < n) v = g[c];
        A[e] = v;
    }
}

}  // namespace sfm

using namespace sfm;

// ---- host-only entry: the plan as arrays (tests emulate the task graph in numpy on a CPU-only box) ----
// h_pattern [nb*nb] lower triangle. Sizes first (all out pointers except h_sizes null), then the arrays.
// h_sizes[9] = {nnzL, n_tasks, n_fwd, n_entries, order_kind, depth, len(fwd_list), len(bwd_list), n_part}
extern "C" int sfm_chol_plan_export(const uint8_t *h_pattern, int32_t nb, int32_t n_ctas, int32_t order_mode,
                                    int64_t *h_sizes, double *h_sim_us, int32_t *h_perm, int32_t *h_colptr,
                                    int32_t *h_rowidx, int32_t *h_tasks, int32_t *h_entries, int32_t *h_fwd_ptr,
                                    int32_t *h_fwd_list, int32_t *h_bwd_ptr, int32_t *h_bwd_list) {
    if (!h_pattern || nb < 1 || n_ctas < 1 || !h_sizes) return SFM_E_INVALID;
    CholDagHost h;
    if (chol_dag_plan(h_pattern, nb, n_ctas, order_mode, h) != 0) return SFM_E_NOT_CONVERGED;
    h_sizes[0] = h.nnzL; h_sizes[1] = (int64_t)h.tasks.size(); h_sizes[2] = h.n_fwd;
    h_sizes[3] = (int64_t)h.entries.size(); h_sizes[4] = h.order_kind; h_sizes[5] = h.depth;
    h_sizes[6] = (int64_t)h.fwd_list.size(); h_sizes[7] = (int64_t)h.bwd_list.size(); h_sizes[8] = h.n_part;
    if (h_sim_us) { h_sim_us[0] = h.sim_fwd_us; h_sim_us[1] = h.sim_bwd_us; }
    if (h_perm) memcpy(h_perm, h.perm.data(), sizeof(int32_t) * nb);
    if (h_colptr) memcpy(h_colptr, h.colptr.data(), sizeof(int32_t) * (nb + 1));
    if (h_rowidx) memcpy(h_rowidx, h.rowidx.data(), sizeof(int32_t) * h.rowidx.size());
    if (h_tasks) memcpy(h_tasks, h.tasks.data(), sizeof(DagTask) * h.tasks.size());
    if (h_entries) memcpy(h_entries, h.entries.data(), sizeof(DagEntry) * h.entries.size());
    if (h_fwd_ptr) memcpy(h_fwd_ptr, h.fwd_ptr.data(), sizeof(int32_t) * h.fwd_ptr.size());
    if (h_fwd_list) memcpy(h_fwd_list, h.fwd_list.data(), sizeof(int32_t) * h.fwd_list.size());
    if (h_bwd_ptr) memcpy(h_bwd_ptr, h.bwd_ptr.data(), sizeof(int32_t) * h.bwd_ptr.size());
    if (h_bwd_list) memcpy(h_bwd_list, h.bwd_list.data(), sizeof(int32_t) * h.bwd_list.size());
    return SFM_OK;
}

// ---- standalone entry (tests, tools): dense SPD S [n,n] + g [n] -> x, through the structure-aware kernel.
// The block pattern is read off the numeric zeros of S. Workspace: sfm_chol_solve_workspace_bytes(n).
// Host-synchronous (plans on the host). order_mode: 0 auto, 1 natural, 2 nested dissection.
// h_info[6] (optional) = {order_kind, tree depth, nnzL blocks, simulated us, measured ms per launch (CUDA events), partial tasks}.
extern "C" int sfm_chol_solve_structured(const double *d_S, const double *d_g, int32_t n, int32_t order_mode,
                                         int32_t repeats, double *d_x, int32_t *d_fail, void *d_ws, size_t ws_bytes,
                                         double *h_info, void *stream) {
    if (!d_S || !d_g || !d_x || !d_fail || !d_ws || n < 1 || repeats < 1) return SFM_E_INVALID;
    if (ws_bytes < sfm_chol_solve_workspace_bytes(n)) return SFM_E_WORKSPACE;
    cudaStream_t st = as_stream(stream);
    const int npad = ceil_div(n, CH_NB) * CH_NB, nb = npad / CH_NB;
    const size_t ld = (size_t)npad;
    double *A = reinterpret_cast<double *>(d_ws);
    double *tail = A + (size_t)(npad + CH_NB) * ld;                 // Linv area of the dense kernel: reused
    unsigned char *d_pat = reinterpret_cast<unsigned char *>(tail);
    double *x = tail + (size_t)nb * CH_NB * CH_NB + align_up((size_t)(nb + 1) * nb + 32, 256) / sizeof(double);
    int *flag = reinterpret_cast<int *>(x + npad);
    SFM_CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(int), st));
    chol_dag_pattern_kernel<<<dim3(nb, nb), 128, 0, st>>>(d_S, n, nb, d_pat);
    SFM_LAUNCH_CHECK();
    std::vector<uint8_t> pat((size_t)nb * nb);
    SFM_CUDA_TRY(cudaMemcpyAsync(pat.data(), d_pat, pat.size(), cudaMemcpyDeviceToHost, st));
    SFM_CUDA_TRY(cudaStreamSynchronize(st));
    int rc = SFM_OK;
    CholDagPlan *plan = chol_dag_create(pat.data(), nb, order_mode, st, &rc);
    if (!plan) return rc != SFM_OK ? rc : SFM_E_NOT_CONVERGED;
    chol_dag_pack_kernel<<<sm_count() * 4, 256, 0, st>>>(d_S, d_g, n, npad, ld, A);
    count_launch();
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float ms = 0.f;
    if (h_info) { cudaEventCreate(&ev0); cudaEventCreate(&ev1); cudaEventRecord(ev0, st); }
    for (int r = 0; r < repeats && rc == SFM_OK; ++r) rc = chol_dag_launch(plan, A, ld, npad, x, flag, st);
    if (h_info) {
        cudaEventRecord(ev1, st);
        cudaEventSynchronize(ev1);
        cudaEventElapsedTime(&ms, ev0, ev1);
        cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    }
    if (rc == SFM_OK) {
        SFM_CUDA_TRY(cudaMemcpyAsync(d_x, x, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, st));
        SFM_CUDA_TRY(cudaMemcpyAsync(d_fail, flag, sizeof(int), cudaMemcpyDeviceToDevice, st));
        SFM_CUDA_TRY(cudaStreamSynchronize(st));
    }
    chol_dag_dump_trace(plan, st);
    if (h_info) {
        const CholDagHost &h = chol_dag_host(plan);
        h_info[0] = h.order_kind; h_info[1] = h.depth; h_info[2] = h.nnzL; h_info[3] = h.sim_fwd_us + h.sim_bwd_us;
        h_info[4] = (double)ms / repeats; h_info[5] = h.n_part;
    }
    chol_dag_destroy(plan, st);
    return rc;
}
