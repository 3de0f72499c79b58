// chol_plan.h -- host-side plan of the structure-aware K11 (chol_dag.cu): elimination order, symbolic
// factorisation at 48 x 48 block granularity, the left-looking task graph and its static schedule.
//
// The reduced camera system S of a large scene is block sparse (two cameras that share no point give an
// exactly zero block), and which blocks are non-zero depends only on the index structure of the problem,
// so all of this is done ONCE per problem (plan time), never per LM iteration. In the reference the solve
// is hidden inside MINPACK's dense pivoted QR (BundleAdjustment.py:205-212); nothing here has a
// counterpart there.
#pragma once
#include <stdint.h>
#include <vector>

namespace sfm {

// task types
enum { DAG_DIAG = 0, DAG_OFF = 1, DAG_BACK = 2, DAG_PART = 3 };

struct DagTask {          // 48 bytes, read by the executing CTA
    int32_t type;         // DAG_DIAG / DAG_OFF / DAG_BACK / DAG_PART (partial sum of a long product list)
    int32_t j;            // column (elimination position)
    int32_t i;            // DAG_OFF: row (elimination position); DAG_PART: 1 = also carries the right-hand-side piece
    int32_t dst;          // DAG_OFF: compact id of L_ij; DAG_PART: compact id (>= nnzL) of its scratch block
    int32_t src_r, src_c; // storage block holding the source A block (block indices); src_r < 0: structural fill
    int32_t src_T;        // 1: the storage block is the transpose of the block this task needs
    int32_t beg, end;     // entry range
    int32_t pad[3];
};

struct DagEntry {         // 16 bytes
    int32_t a, b;         // forward product: compact ids of L_it and L_jt; forward add (b < 0): scratch block a of a
                          // partial task; backward: compact id of L_ij, row position i
    int32_t t;            // forward product: column t (for the right-hand side y_t); add: index of the partial task
    int32_t pad;
};

struct CholDagHost {
    int nb = 0, nnzL = 0, n_part = 0, G = 0;   // n_part: scratch blocks of partial tasks (compact ids nnzL ..)
    int order_kind = 0;               // 0 natural, 1.. nested dissection variants
    int depth = 0;                    // height of the elimination tree in blocks
    double sim_fwd_us = 0, sim_bwd_us = 0;
    std::vector<int32_t> perm;        // [nb] elimination position -> storage block index
    std::vector<int32_t> colptr, rowidx;   // strictly-lower structure of L by column (positions); id = index in rowidx
    std::vector<DagTask> tasks;       // forward tasks, then backward tasks
    int n_fwd = 0;
    std::vector<DagEntry> entries;
    std::vector<int32_t> fwd_ptr, fwd_list, bwd_ptr, bwd_list;   // per-CTA task lists (indices into tasks)
};

// pattern[i * nb + j] != 0 for i >= j (storage block indices): block (i, j) of the lower triangle of S can be
// non-zero. order_mode: 0 = choose by simulated makespan, 1 = natural order, 2 = nested dissection.
// Returns 0, or 1 when the task graph is too large to be worth it (caller falls back to the dense kernel).
int chol_dag_plan(const uint8_t *pattern, int nb, int G, int order_mode, CholDagHost &out);

}  // namespace sfm
