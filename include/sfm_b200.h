/*
 * sfm_b200.h -- C ABI of libsfm_b200.so, the B200 (sm_100a) numerical core for the
 * Phase-1 SfM hot path of pvrohin/Structure-from-Motion-.
 *
 * Conventions:
 *   - every entry point is extern "C", returns int (0 = SFM_OK, else an SFM_E_* code,
 *     never throws), takes plain pointers and sizes only;
 *   - pointers named d_* are DEVICE pointers owned by the caller; the library never
 *     allocates device memory behind the caller's back: each entry point that needs
 *     scratch has a *_workspace_bytes() query and takes the workspace explicitly;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it and the
 *     call returns without synchronising unless stated otherwise;
 *   - all real data is float64, all indices int32, row-major.
 *
 * Each entry point cites the reference interface it replaces ("Phase 1/<file>:<line>").
 * The reference-side binding (ctypes) is shown in INTEGRATION.md.
 */
#ifndef SFM_B200_H
#define SFM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SFM_OK 0
#define SFM_E_INVALID 1      /* bad argument (null pointer, negative size, ...) */
#define SFM_E_CUDA 2         /* a CUDA runtime call or launch failed; see sfm_last_cuda_error() */
#define SFM_E_WORKSPACE 3    /* workspace too small */
#define SFM_E_NCCL 4         /* NCCL missing or a collective failed */
#define SFM_E_NOT_CONVERGED 5

/* ABI version of this header; bumped on any signature change. */
int sfm_abi_version(void);
/* cudaGetErrorString of the last failing CUDA call seen by the library (static storage). */
const char *sfm_last_cuda_error(void);
/* Number of kernel launches issued by the library since load (all entry points). Used by
 * bench.py for the "gpu_launches" claim. */
int64_t sfm_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * Host-side sampler: the reference draws one `random.sample(range(n), k)` per RANSAC
 * hypothesis from CPython's global Mersenne Twister (GetInliersRANSAC.py:55, PnPRANSAC.py:49).
 * This advances a copy of that MT19937 state exactly as CPython's random.sample would for
 * `count` consecutive calls and writes the index sets. HOST pointers.
 *   mt_state : 624 words, mt_index : position (both from random.getstate(), written back)
 *   out      : [count, k] int32
 * ------------------------------------------------------------------------------------------ */
int sfm_mt_sample_stream(uint32_t *mt_state, int32_t *mt_index, int32_t n, int32_t k,
                         int64_t count, int32_t *out);

/* ------------------------------------------------------------------------------------------
 * F-matrix RANSAC (replaces GetInliersRANSAC.py:53-92 + EstimateFundamentalMatrix.py:30-78)
 * d_x1x2   : [n,4] rows (x1,y1,x2,y2)
 * d_sample : [h,8] int32 indices into 0..n-1 (host-generated reference stream)
 * ------------------------------------------------------------------------------------------ */

/* K1: Hartley-normalised 8-point estimate for h samples -> d_F [h,9] (row-major 3x3 each,
 * normalised by F[2][2] like EstimateFundamentalMatrix.py:78). */
int sfm_f8pt_batch(const double *d_x1x2, int32_t n, const int32_t *d_sample, int32_t h,
                   double *d_F, void *stream);

/* K2: inlier count of every hypothesis over all n correspondences with the reference's
 * symmetric epipolar distance (GetInliersRANSAC.py:67-82). d_counts [h] int32 (overwritten). */
int sfm_f_score_count(const double *d_x1x2, int32_t n, const double *d_F, int32_t h,
                      double threshold, int32_t *d_counts, void *stream);

/* Diagnostics for K2. sfm_f_score_count evaluates each test in FP32 with a forward error bound and
 * re-evaluates in FP64 whenever the bound does not separate the margin from zero, so its counts
 * equal those of the all-FP64 kernel exposed here; sfm_f_score_fp64_tests reports (and optionally
 * resets) how many tests took the FP64 path on this device. Host-synchronous. */
int sfm_f_score_count_fp64(const double *d_x1x2, int32_t n, const double *d_F, int32_t h,
                           double threshold, int32_t *d_counts, void *stream);
int sfm_f_score_fp64_tests(uint64_t *out, int32_t reset);

/* K3: first-maximum argmax (strict '>' rule of GetInliersRANSAC.py:85).
 * d_best[0] = index of the first hypothesis with the maximum count (global index =
 * local index + idx_offset), d_best[1] = that count. d_key (uint64, optional, may be null)
 * receives (count << 32) | (0xFFFFFFFF - global_idx) for a cross-GPU max-reduce. */
int sfm_argmax_first(const int32_t *d_counts, int32_t h, int32_t idx_offset, int32_t *d_best,
                     uint64_t *d_key, void *stream);

/* K3 for many consecutive RANSAC runs in one launch: segment b = d_counts[b * seg_len .. (b+1) * seg_len),
 * d_best[2b] = first index (within the segment) with the segment's maximum count, d_best[2b+1] = that count --
 * the result of n_seg calls of sfm_argmax_first (PnPRANSAC.py:72 once per registration). */
int sfm_argmax_first_batch(const int32_t *d_counts, int32_t n_seg, int32_t seg_len, int32_t *d_best,
                           void *stream);

/* K4: inlier mask + per-point error for ONE fundamental matrix d_Fone[9]
 * (GetInliersRANSAC.py:67-81). d_mask [n] uint8, d_err [n] float64 (may be null). */
int sfm_f_inlier_mask(const double *d_x1x2, int32_t n, const double *d_Fone, double threshold,
                      uint8_t *d_mask, double *d_err, void *stream);

/* K1..K4 chained on one stream. Outputs:
 *   d_F      [h,9]  all hypotheses        d_counts [h]
 *   d_best   [2]    (winner index, count) d_Fbest  [9]   (copy of d_F[winner])
 *   d_mask   [n]    inlier mask of the winner (all zero when the best count is 0)
 *   d_err    [n]    winner's per-point error in the reference's exact formula (may be null) */
int sfm_f_ransac(const double *d_x1x2, int32_t n, const int32_t *d_sample, int32_t h,
                 double threshold, double *d_F, int32_t *d_counts, int32_t *d_best,
                 double *d_Fbest, uint8_t *d_mask, double *d_err, void *stream);

/* Multi-GPU tail of the hypothesis-sharded RANSAC: d_key holds the max-allreduced
 * (count << 32) | (0xFFFFFFFF - global_idx). Decodes the global winner into d_best[2],
 * recomputes its F from d_sample_all [h_total,8] (the whole stream, replicated on every rank)
 * into d_Fbest[9] and writes the winner's mask / error like sfm_f_ransac. */
int sfm_f_winner_from_key(const double *d_x1x2, int32_t n, const int32_t *d_sample_all,
                          int32_t h_total, const uint64_t *d_key, double threshold,
                          int32_t *d_best, double *d_Fbest, uint8_t *d_mask, double *d_err,
                          void *stream);

/* EstimateFundamentalMatrix for ANY n >= 8 (EstimateFundamentalMatrix.py:3-83 called with all
 * points). d_ws: workspace of sfm_fundamental_npt_workspace_bytes(n) bytes. d_F [9]. */
size_t sfm_fundamental_npt_workspace_bytes(int32_t n);
int sfm_fundamental_npt(const double *d_x1, const double *d_x2, int32_t n, double *d_F,
                        void *d_ws, size_t ws_bytes, void *stream);

/* Allocation-free variant of sfm_fundamental_npt. It has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc). Same kernel, same result, but the 9 x n design matrix
 * lives in a file-scope `static __device__` array inside the library instead of a caller workspace:
 *   - limit: n <= SFM_FUNDAMENTAL_NPT_STATIC_MAX_N (16384 correspondences = 1.18 MB); every GPU that loads the
 *     library holds that array until the process exits. Larger n returns SFM_E_WORKSPACE and leaves a message in
 *     sfm_last_cuda_error(); use sfm_fundamental_npt with your own workspace then;
 *   - the call itself allocates nothing (no cudaMalloc / cudaFree / allocating library routine); the array's
 *     device address is looked up once per device on the first call with that device current, so make the first
 *     call on each device outside stream capture;
 *   - all calls on one device share the array, which keeps the last call's contents (the kernel overwrites what
 *     it reads, nothing needs clearing): calls that OVERLAP on the GPU (different streams) race -- serialise them. */
#define SFM_FUNDAMENTAL_NPT_STATIC_MAX_N 16384
int sfm_fundamental_npt_static(const double *d_x1, const double *d_x2, int32_t n, double *d_F, void *stream);

/* ------------------------------------------------------------------------------------------
 * LM status word written by the batched solvers: (status << 16) | trial_steps, status =
 * 1 gtol, 2 ftol, 3 xtol, 5 iteration cap, -1 failed (non-finite residuals; the input is
 * returned unchanged, like the reference's except branches). Tolerance `tol` applies to the
 * MINPACK ftol/xtol/gtol definitions (scipy default 1e-8; pass <= 1e-12 to converge at least
 * as tightly as the reference).
 * ------------------------------------------------------------------------------------------ */

/* N1: LinearTriangulation (LinearTriangulation.py:3-92) for n points and ONE camera pair: the DLT
 * null vector of the 4x4 system per point, de-homogenised when |X~[3]| > 1e-8. h_P1,h_P2: HOST 3x4
 * projection matrices K[R|-RC]; d_x1,d_x2 [n,2]; d_X [n,3] out. */
int sfm_lintri(const double *h_P1, const double *h_P2, const double *d_x1, const double *d_x2,
               int32_t n, double *d_X, void *stream);

/* K5: NonLinearTriangulation (NonLinearTriangulation.py:53-130) for n points and ONE camera
 * pair. h_P1,h_P2: HOST 3x4 projection matrices K[R|-RC] (row-major, 12 doubles each).
 * d_x1,d_x2 [n,2]; d_X0 [n,3] initial points; d_X [n,3] out (may alias d_X0);
 * d_cost [n] final 0.5|r|^2 (may be null); d_status [n] (may be null). */
int sfm_nltri(const double *h_P1, const double *h_P2, const double *d_x1, const double *d_x2,
              const double *d_X0, int32_t n, int32_t max_iter, double tol, double *d_X,
              double *d_cost, int32_t *d_status, void *stream);

/* K6: PnP-RANSAC consensus scoring (PnPRANSAC.py:59-69). d_P [h,12] hypothesis projection
 * matrices K[R|-RC]; d_X [n,3]; d_x [n,2]; d_counts [h] (overwritten). Pair with
 * sfm_argmax_first for the strict-'>' first-maximum rule (PnPRANSAC.py:72). */
int sfm_pnp_score_count(const double *d_P, const double *d_X, const double *d_x, int32_t n,
                        int32_t h, double threshold, int32_t *d_counts, void *stream);

/* K6 for n_regs registrations in one launch (PnPRANSAC.py:59-69 once per registration, BASELINE config 4):
 * registration b owns rows d_offsets[b] .. d_offsets[b+1]-1 of d_X [M,3] / d_x [M,2] (max_n = the longest),
 * hypotheses d_P[b * h_per_reg .. ) and counts d_counts[b * h_per_reg .. ) (overwritten). The same kernel and the
 * same arithmetic as sfm_pnp_score_count, so the counts equal those of n_regs separate calls. */
int sfm_pnp_score_count_batch(const double *d_P, const double *d_X, const double *d_x,
                              const int32_t *d_offsets, int32_t n_regs, int32_t max_n, int32_t h_per_reg,
                              double threshold, int32_t *d_counts, void *stream);

/* K7: NonLinearPnP (NonlinearPnP.py:47-123) for a batch of independent registrations.
 * h_K: HOST 3x3 intrinsics. Problem b uses rows d_offsets[b]..d_offsets[b+1]-1 of d_X [M,3] /
 * d_x [M,2]. d_rt0/d_rt [B,6] = [rotvec, t = -R C] in/out; d_cost [B], d_status [B] may be null.
 * A problem with fewer than 4 points returns its input (NonlinearPnP.py:82-84). */
int sfm_nlpnp_batch(const double *h_K, const double *d_X, const double *d_x,
                    const int32_t *d_offsets, int32_t n_problems, const double *d_rt0,
                    int32_t max_iter, double tol, double *d_rt, double *d_cost,
                    int32_t *d_status, void *stream);

/* ------------------------------------------------------------------------------------------
 * Bundle adjustment (replaces perform_bundle_adjustment's least_squares call,
 * BundleAdjustment.py:205-212, and bundle_adjustment_residuals, BundleAdjustment.py:43-110).
 * Sparse form of the reference's problem:
 *   d_cam     [n_cams,6]  [rotvec(3), t(3) = -R C]       (BundleAdjustment.py:183-192)
 *   d_pts     [n_pts,3]                                   (BundleAdjustment.py:195-197)
 *   d_obs     [n_obs,2]   observed pixel of observation o
 *   d_cam_idx [n_obs], d_pt_idx [n_obs]  int32; observations sorted by point, i.e. d_pt_idx
 *             non-decreasing (the reference's own order, BundleAdjustment.py:164-169)
 *   h_K       HOST 3x3 intrinsics (row-major)
 * ------------------------------------------------------------------------------------------ */
typedef struct sfm_ba_options {
    int32_t max_trials;  /* scipy max_nfev: residual evaluations at trial points (reference: 100) */
    int32_t verbose;     /* 1: one line per trial step on stderr */
    double ftol, xtol, gtol; /* MINPACK / scipy stopping tolerances (reference: scipy default 1e-8) */
    double lambda0;      /* initial damping of (J^T J + lambda D^2); D = Jacobian column norms */
    int32_t reuse_index; /* 1: d_ws still holds the camera-major index of a previous sfm_ba_solve call
                            on the SAME d_cam_idx / d_pt_idx contents (skips the index build; d_obs MAY have
                            changed: the library's camera-major copy of it is refreshed on every call) */
    int32_t reserved;    /* bit 0: record the phase profile of this solve (sfm_ba_last_profile) */
} sfm_ba_options;

typedef struct sfm_ba_result {
    double cost0, cost;  /* 0.5 |r|^2 before / after (scipy result.cost, BundleAdjustment.py:236) */
    double gnorm;        /* MINPACK gtol measure at the last linearisation */
    double lambda;       /* final damping */
    int32_t status;      /* scipy codes: -1 non-finite residuals (inputs returned unchanged),
                            0 max_trials reached, 1 gtol, 2 ftol, 3 xtol, 4 ftol and xtol */
    int32_t n_linearisations; /* LM iterations: Jacobian / normal-equation builds */
    int32_t n_trials;    /* trial steps solved and evaluated (>= n_linearisations - 1) */
    int32_t n_chol_fail; /* trial steps rejected because S was not numerically positive definite */
} sfm_ba_result;

/* One process per GPU: the BA entry points bind their per-device state (kernel attributes, side stream, plan
 * cache) to the device that is current at the FIRST call; calls made with another device current return
 * SFM_E_INVALID (message in sfm_last_cuda_error()). */
void sfm_ba_default_options(sfm_ba_options *opts);
size_t sfm_ba_workspace_bytes(int32_t n_cams, int32_t n_pts, int32_t n_obs);
/* Everything sfm_ba_solve* needs lives in the caller's workspace except the plan-time pair list of the
 * gather-form Schur complement (12 B per pair of observations of one point), which the library keeps in
 * stream-ordered allocations cached per workspace pointer (at most 16 entries, replaced / evicted
 * automatically). This releases all of them on `stream` (e.g. before handing the device to another
 * library); the next solve rebuilds what it needs. Returns the number of bytes released through *bytes
 * (may be NULL). */
int sfm_ba_release_cache(uint64_t *bytes, void *stream);
/* The same for ONE workspace (call before freeing or re-purposing d_ws; unknown pointers are ignored). */
int sfm_ba_release_workspace(const void *d_ws, void *stream);

/* Phase profile of the last sfm_ba_solve* call of this process that ran with opts.reserved bit 0 set (CUDA events
 * on the caller's stream; diagnostic, used by bench.py for the multi-GPU time breakdown). HOST out, milliseconds
 * summed over the solve: [0] Schur complement (K10*), [1] all-reduce of S / g_red incl. pack + unpack (0 on one
 * GPU), [2] camera solve K11, [3] rest of the trial steps (back substitution, update, cost pass, scalar
 * all-reduce), [4] linearisations (K9, incl. their all-reduce); then [5] trial steps, [6] linearisations,
 * [7] bytes all-reduced for S / g_red per trial step. */
int sfm_ba_last_profile(double *h_out8);

/* K8 alone: residual vector of bundle_adjustment_residuals (BundleAdjustment.py:43-110),
 * d_r [n_obs,2] interleaved [du,dv] = observed - projected (may be null), d_cost[0] = 0.5 |r|^2
 * (may be null). Synchronises the stream once (index validation). */
int sfm_ba_residuals(const double *d_cam, const double *d_pts, const double *d_obs,
                     const int32_t *d_cam_idx, const int32_t *d_pt_idx, int32_t n_cams,
                     int32_t n_pts, int32_t n_obs, const double *h_K, double *d_r, double *d_cost,
                     void *d_ws, size_t ws_bytes, void *stream);

/* The whole LM solve (K8-K12). d_cam / d_pts are updated in place; h_result is a HOST struct.
 * Synchronous: returns after the last step (the LM controller reads a few scalars per trial).
 * SFM_E_INVALID when an index is out of range or d_pt_idx is not sorted. */
int sfm_ba_solve(double *d_cam, double *d_pts, const double *d_obs, const int32_t *d_cam_idx,
                 const int32_t *d_pt_idx, int32_t n_cams, int32_t n_pts, int32_t n_obs,
                 const double *h_K, const sfm_ba_options *opts, sfm_ba_result *h_result,
                 void *d_ws, size_t ws_bytes, void *stream);

/* ------------------------------------------------------------------------------------------
 * Multi-GPU (one process per GPU). The library opens libnccl.so.2 with dlopen (the
 * copy already in the process, e.g. torch's) -- SFM_E_NCCL when there is none. Rank 0 obtains a
 * unique id, the host distributes its 128 bytes (torch.distributed / MPI / a file), every rank
 * creates the communicator.
 * ------------------------------------------------------------------------------------------ */
const char *sfm_nccl_last_error(void);
int sfm_nccl_unique_id(void *h_id128);
int sfm_nccl_comm_create(const void *h_id128, int32_t rank, int32_t world, void **comm_out);
int sfm_nccl_comm_destroy(void *comm);
/* in-place all-reduces on that communicator (the F-RANSAC winner key: max over u64) */
int sfm_allreduce_sum_f64(void *comm, double *d_buf, int64_t count, void *stream);
int sfm_allreduce_max_u64(void *comm, uint64_t *d_buf, int64_t count, void *stream);

/* Point-sharded bundle adjustment: this rank passes ALL cameras (replicated, identical on every
 * rank) and a contiguous range of the points with all of their observations (local point indices
 * 0..n_pts_local-1, sorted). Per linearisation U/g_c/cost are all-reduced, per trial step S/g_red
 * and six scalars; every rank solves the camera system redundantly and ends with identical
 * cameras and results. world == 1 (comm may be null) is sfm_ba_solve. */
int sfm_ba_solve_mgpu(double *d_cam, double *d_pts_local, const double *d_obs_local,
                      const int32_t *d_cam_idx_local, const int32_t *d_pt_idx_local, int32_t n_cams,
                      int32_t n_pts_local, int32_t n_obs_local, const double *h_K,
                      const sfm_ba_options *opts, sfm_ba_result *h_result, void *d_ws, size_t ws_bytes,
                      void *comm, int32_t rank, int32_t world, void *stream);

/* One linearisation pass in isolation, for bench.py / ncu (the passes of sfm_ba_solve, same kernels):
 * which = bit 0 point pass K9p (W, V, g_p, cost partials), bit 1 camera pass K9c (U, g_c), bit 2 cost
 * pass K8, bit 3 reuse the linearised-camera table of the previous call. Outputs stay in d_ws. */
int sfm_ba_linearise(const double *d_cam, const double *d_pts, const double *d_obs,
                     const int32_t *d_cam_idx, const int32_t *d_pt_idx, int32_t n_cams, int32_t n_pts,
                     int32_t n_obs, const double *h_K, int32_t which, int32_t reuse_index, void *d_ws,
                     size_t ws_bytes, void *stream);

/* K11 alone: dense SPD solve S x = g by the blocked Cholesky used inside sfm_ba_solve.
 * d_S [n,n] row-major (lower triangle read), d_g [n], d_x [n] out, d_fail[0] = 1 when a pivot
 * was not positive. */
size_t sfm_chol_solve_workspace_bytes(int32_t n);
int sfm_chol_solve(const double *d_S, const double *d_g, int32_t n, double *d_x, int32_t *d_fail,
                   void *d_ws, size_t ws_bytes, void *stream);

/* K11s: the same solve through the structure-aware kernel (csrc/chol_dag.cu): the 48 x 48 block pattern
 * is read off the numeric zeros of d_S, analysed on the host (nested dissection, symbolic factorisation,
 * static task schedule -- per problem, not per solve) and factored as a left-looking task graph whose chain
 * is the height of the elimination tree instead of the number of panels. Host-synchronous.
 *   order_mode 0 = pick the order with the shortest simulated schedule, 1 = natural, 2 = nested dissection
 *   repeats    the factor + solve is launched this many times (timing); results of the last launch
 *   d_fail[0]  1 = a pivot was not positive, 2 = an internal wait timed out (scheduling bug)
 *   h_info[6]  optional HOST out: {order kind, elimination-tree height, off-diagonal L blocks, simulated us,
 *              measured ms per launch (CUDA events over the repeats), partial-sum tasks}
 * Workspace as sfm_chol_solve. Replaces what MINPACK's dense QR does behind BundleAdjustment.py:205-212. */
int sfm_chol_solve_structured(const double *d_S, const double *d_g, int32_t n, int32_t order_mode,
                              int32_t repeats, double *d_x, int32_t *d_fail, void *d_ws, size_t ws_bytes,
                              double *h_info, void *stream);

/* The plan of K11s as plain HOST arrays (no device work): lets tests replay the task graph on a CPU-only
 * box. h_pattern [nb,nb] (lower triangle, != 0 = block present). Call once with only h_sizes to size the
 * arrays: h_sizes[9] = {off-diagonal L blocks, tasks, forward tasks, entries, order kind, tree height,
 * len(fwd_list), len(bwd_list), partial-sum scratch blocks}; h_tasks rows are 12 int32 (type, j, i, dst,
 * src_r, src_c, src_T, beg, end, pad x3), h_entries rows 4 int32 (a, b, t, pad), see csrc/chol_plan.h;
 * *_ptr [n_ctas+1]. Any array pointer may be null. */
int sfm_chol_plan_export(const uint8_t *h_pattern, int32_t nb, int32_t n_ctas, int32_t order_mode,
                         int64_t *h_sizes, double *h_sim_us, int32_t *h_perm, int32_t *h_colptr,
                         int32_t *h_rowidx, int32_t *h_tasks, int32_t *h_entries, int32_t *h_fwd_ptr,
                         int32_t *h_fwd_list, int32_t *h_bwd_ptr, int32_t *h_bwd_list);

/* Same solve with a per-phase time profile accumulated into d_phase_ns[6] (device, caller-zeroed;
 * nanoseconds seen by CTA 0: diagonal factor, panel solve, barrier 1, trailing update, barrier 2,
 * back substitution). Diagnostic entry point used by tools/chol_probe.py. */
int sfm_chol_solve_profiled(const double *d_S, const double *d_g, int32_t n, double *d_x,
                            int32_t *d_fail, void *d_ws, size_t ws_bytes, uint64_t *d_phase_ns,
                            void *stream);

/* Diagnostic: SM cycles of the 48 x 48 diagonal-block factor alone (d_out8[6] column sweep,
 * [7] whole factor, accumulated; variant must be 0). tools/diag_bench.py. */
int sfm_chol_diag_bench(const double *d_A48, int32_t variant, uint64_t *d_out8, void *stream);

/* ---- a8 / K6a: LinearPnP on the device (replaces LinearPnP.py:3-96) -----------------------------------
 * The hypothesis generator of PnPRANSAC (PnPRANSAC.py:49-55: LinearPnP on 4 sampled correspondences) and its
 * all-points fallback (:82-87). The reference takes `Vt[-1]` of numpy.linalg.svd of the 2N x 12 DLT matrix; for
 * N = 4 (and 5) that matrix is rank deficient and the row is LAPACK's pick inside the null space: the last column
 * of the product of the right Householder reflectors of dgebd2. The kernels replay that construction and the
 * Householder sign convention of the QR clean-up (csrc/linpnp_math.cuh), which reproduces the reference to 1e-8
 * wherever its output is a function of its input. It is not in the `det(R) < 0` branch of LinearPnP.py:88-90
 * (27 % of random 4-point samples), where LAPACK flips a singular vector chosen by rounding noise; those
 * hypotheses get a documented deterministic pose and bit 0 of their flag set (bit 1: non-finite pose).
 *   h_K, h_Kinv : HOST 3x3; h_Kinv = the caller's inverse of K (the reference uses numpy.linalg.inv), may be
 *                 null (adjugate formula then)
 *   d_P [.,12]  : K [R | -R C] of every hypothesis, the input of sfm_pnp_score_count (PnPRANSAC.py:60)
 * sfm_linear_pnp4_batch: d_sample [n_regs * h_per_reg, 4] int32, the samples of registration r index its own
 *   correspondences d_offsets[r] .. d_offsets[r+1] of the concatenated d_X [.,3], d_x [.,2] (d_offsets == null:
 *   one registration of n correspondences). d_C [.,3], d_R [.,9], d_flag [.] may be null. Thread per hypothesis.
 *   policy bit 0 (the PnPRANSAC drop-in sets it): hypotheses in the noise-decided branch do not vote -- their d_P
 *   row is NaN, so sfm_pnp_score_count gives them zero inliers. In the reference their pose is a rotation reflected
 *   along a direction drawn from rounding noise, which all but never wins the consensus (measured: never in 21
 *   registrations x 200-300 hypotheses), whereas any fixed rule occasionally produces a well-scoring pose and would
 *   change the winner; muting them keeps the WINNER equal to the reference's. policy 0: they vote with the
 *   documented deterministic pose.
 * sfm_linear_pnp_npt: all n >= 4 correspondences of one registration -> d_CR [12] = (C, R row-major), d_P [12]
 *   and d_flag [1] (either may be null); workspace sfm_linear_pnp_npt_workspace_bytes(n). */
int sfm_linear_pnp4_batch(const double *d_X, const double *d_x, const int32_t *d_offsets, int32_t n_regs,
                          int32_t n, const int32_t *d_sample, int32_t h_per_reg, const double *h_K,
                          const double *h_Kinv, int32_t policy, double *d_C, double *d_R, double *d_P,
                          uint8_t *d_flag, void *stream);
size_t sfm_linear_pnp_npt_workspace_bytes(int32_t n);
int sfm_linear_pnp_npt(const double *d_X, const double *d_x, int32_t n, const double *h_K,
                       const double *h_Kinv, double *d_CR, double *d_P, uint8_t *d_flag, void *d_ws,
                       size_t ws_bytes, void *stream);

/* ---- a3 / a5: the residual functions the reference exports next to its solvers -----------------------
 * sfm_tri_residuals: Loss(X, x1, x2, P1, P2) of NonLinearTriangulation.py:5-50 for n points:
 *   d_r [n,4] = [x1 - pi(P1 X~), x2 - pi(P2 X~)], pi = exact division, a view with |w| < 1e-8 gives 0.
 * sfm_pnp_residuals: NonLinearPnPLoss(X0, X, x, K) of NonlinearPnP.py:5-44: h_rt HOST [6] = (rotvec, t),
 *   d_r [2n] interleaved [du, dv] = x - a[:2] / (a[2] + 1e-8). h_P1, h_P2 [12], h_K [9]: HOST. */
int sfm_tri_residuals(const double *h_P1, const double *h_P2, const double *d_x1, const double *d_x2,
                      const double *d_X, int32_t n, double *d_r, void *stream);
int sfm_pnp_residuals(const double *h_K, const double *h_rt, const double *d_X, const double *d_x, int32_t n,
                      double *d_r, void *stream);

/* @@MORE@@ (further sections are appended above this marker as the path widens) */

/* ---- N4: two-view pose front end ------------------------------------------------
 * Replaces EssentialMatrixFromFundamentalMatrix.py:19, ExtractCameraPose.py:24-81, the four
 * LinearTriangulation calls of Wrapper_dev.py:166-168 and DisambiguateCameraPose.py:53-92, so a pair
 * stays on the device between F-RANSAC and NonLinearTriangulation.
 *   d_F      [9]      fundamental matrix (device, e.g. d_Fbest of sfm_f_ransac);  h_K HOST 3x3
 *   d_cfg48  [4,12]   (C[3], R[9]) in the reference's order (R1,t) (R1,-t) (R2,t) (R2,-t); the
 *                     assignment of the list positions depends on the SVD sign convention (LAPACK's
 *                     in the reference, documented in twoview.cu here): the SET is the same
 *   d_P60    [5,12]   K[I|0] and K R_i [I | -C_i]
 *   d_Xall   [4,n,3]  DLT points under each configuration;  d_counts4 [4] cheirality votes
 *   d_best2  [2]      (index, votes) of the first maximum;  d_pose12 = its (C, R) */
int sfm_two_view_configs(const double *d_F, const double *h_K, double *d_cfg48, double *d_P60,
                         int32_t *d_counts4, void *stream);
int sfm_cheirality_count(const double *d_cfg48, const double *d_Xall, int32_t n, int32_t *d_counts4,
                         void *stream);
int sfm_two_view_pose(const double *d_F, const double *h_K, const double *d_x1, const double *d_x2,
                      int32_t n, double *d_cfg48, double *d_P60, double *d_Xall, int32_t *d_counts4,
                      int32_t *d_best2, double *d_pose12, void *stream);

/* ---- N2: homography RANSAC ----------------------------------------------------------
 * Replaces GetHomographyInliers.py: find_homography :4-82 and the RANSAC loop :85-165, on the K1..K4
 * skeleton. d_x1x2 [n,4] rows (x1,y1,x2,y2); d_sample [h,4] from the reference's random.sample
 * stream (:126); threshold in pixels (strict '<' on the reprojection distance, :149). */
int sfm_h4pt_batch(const double *d_x1x2, int32_t n, const int32_t *d_sample, int32_t h, double *d_H,
                   void *stream);
int sfm_h_score_count(const double *d_x1x2, int32_t n, const double *d_H, int32_t h, double threshold,
                      int32_t *d_counts, void *stream);
int sfm_h_inlier_mask(const double *d_x1x2, int32_t n, const double *d_Hone, double threshold,
                      uint8_t *d_mask, double *d_err, void *stream);
/* K1..K4 chained; outputs as sfm_f_ransac: d_H [h,9], d_counts [h], d_best [2], d_Hbest [9],
 * d_mask [n] (all zero when the best count is 0), d_err [n] (may be null). */
int sfm_h_ransac(const double *d_x1x2, int32_t n, const int32_t *d_sample, int32_t h, double threshold,
                 double *d_H, int32_t *d_counts, int32_t *d_best, double *d_Hbest, uint8_t *d_mask,
                 double *d_err, void *stream);
/* find_homography for ANY n >= 4 (one block, Jacobi on the 2n x 9 design matrix). */
size_t sfm_homography_npt_workspace_bytes(int32_t n);
int sfm_homography_npt(const double *d_x1, const double *d_x2, int32_t n, double *d_H, void *d_ws,
                       size_t ws_bytes, void *stream);

/* ---- N3: bundle-adjustment problem assembly on the device ---------------------------
 * Replaces BundleAdjustment.py:148-176. Inputs are the wrapper's dense arrays on the device:
 *   d_valid [F] uint8 (filtered_world_coords == 1), d_flags [F,row_stride] uint8 (== 1 means observed),
 *   d_fx, d_fy [F,row_stride] f64, d_world [F,3] f64; cameras 0..n_cams-1 are used (:156).
 * Outputs in the reference's order (valid rows ascending, cameras ascending): d_cam_idx, d_pt_idx, d_obs
 * (capacity F*n_cams entries), d_pts [F,3] and d_valid_rows [F] (compacted, capacity F),
 * d_totals2 = (number of points, number of observations). */
size_t sfm_ba_assemble_workspace_bytes(int32_t n_features);
int sfm_ba_assemble(const uint8_t *d_valid, const uint8_t *d_flags, const double *d_fx, const double *d_fy,
                    const double *d_world, int32_t n_features, int32_t row_stride, int32_t n_cams,
                    int32_t *d_cam_idx, int32_t *d_pt_idx, double *d_obs, double *d_pts,
                    int32_t *d_valid_rows, int32_t *d_totals2, void *d_ws, size_t ws_bytes, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SFM_B200_H */
