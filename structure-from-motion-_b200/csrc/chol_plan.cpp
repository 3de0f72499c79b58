// chol_plan.cpp -- plan of the structure-aware K11 (see chol_plan.h): ordering, symbolic factorisation,
// task graph, static schedule. Pure host code, runs once per problem.
#include "chol_plan.h"

#include <algorithm>
#include <functional>
#include <queue>
#include <utility>

namespace sfm {

namespace {

typedef std::vector<std::vector<int>> Graph;

// ---- ordering: nested dissection on BFS level structures ------------------------------------------
// For a connected piece of the block graph: breadth-first levels from a pseudo-peripheral node; a node
// of level l has neighbours only in levels l-1, l, l+1, so any level is a separator. The level where the
// node count crosses one half is taken, thinned (a separator node without a neighbour on one side joins
// the other side), both sides are ordered recursively and the separator goes last. Pieces of at most
// `leaf` nodes (or without three levels) are ordered by reverse Cuthill-McKee.
struct Orderer {
    const Graph &adj;
    int leaf;
    std::vector<int> mark, level, out;
    int stamp = 0;
    Orderer(const Graph &g, int leaf_) : adj(g), leaf(leaf_), mark(g.size(), 0), level(g.size(), 0) {}

    // BFS inside the node set stamped `s`; returns nodes in visiting order, fills level[]
    std::vector<int> bfs(int start, int s) {
        std::vector<int> order;
        const int visited = ++stamp;
        std::vector<int> q{start};
        level[start] = 0;
        mark[start] = visited;
        for (size_t h = 0; h < q.size(); ++h) {
            const int u = q[h];
            order.push_back(u);
            std::vector<int> nb;
            for (int v : adj[u])
                if (mark[v] == s) nb.push_back(v);
            std::sort(nb.begin(), nb.end(), [&](int a, int b) {
                return adj[a].size() != adj[b].size() ? adj[a].size() < adj[b].size() : a < b;
            });
            for (int v : nb) {
                mark[v] = visited;
                level[v] = level[u] + 1;
                q.push_back(v);
            }
        }
        // restore the set stamp on what was visited (callers keep using `s`)
        for (int u : order) mark[u] = s;
        return order;
    }

    void emit_rcm(const std::vector<int> &cm) {
        for (size_t k = cm.size(); k-- > 0;) out.push_back(cm[k]);
    }

    void order_set(std::vector<int> nodes) {
        if (nodes.empty()) return;
        const int s = ++stamp;
        for (int u : nodes) mark[u] = s;
        // connected components: independent, simply concatenated
        std::vector<std::vector<int>> comps;
        {
            std::vector<int> todo = nodes;
            std::sort(todo.begin(), todo.end());
            std::vector<char> seen(adj.size(), 0);
            for (int u : todo) {
                if (seen[u]) continue;
                std::vector<int> c = bfs(u, s);
                for (int v : c) seen[v] = 1;
                comps.push_back(c);
            }
        }
        if (comps.size() > 1) {
            for (auto &c : comps) order_set(c);
            return;
        }
        // pseudo-peripheral start: farthest node of a BFS, twice
        int start = *std::min_element(nodes.begin(), nodes.end(), [&](int a, int b) {
            return adj[a].size() != adj[b].size() ? adj[a].size() < adj[b].size() : a < b;
        });
        std::vector<int> cm;
        for (int rep = 0; rep < 3; ++rep) {
            cm = bfs(start, s);
            const int far = cm.back();
            if (rep == 2 || far == start) break;
            start = far;
        }
        const int n = (int)nodes.size();
        const int nlev = level[cm.back()] + 1;
        if (n <= leaf || nlev < 3) {
            emit_rcm(cm);
            return;
        }
        // level where the cumulative count crosses n / 2 (never the first or last level)
        std::vector<int> cnt(nlev, 0);
        for (int u : cm) ++cnt[level[u]];
        int cut = 1, acc = cnt[0];
        for (int l = 1; l < nlev - 1; ++l) {
            cut = l;
            if (2 * (acc + cnt[l]) >= n) break;
            acc += cnt[l];
        }
        // side: 0 left, 1 separator, 2 right
        std::vector<int> sep, left, right;
        std::vector<int> side(adj.size(), -1);
        for (int u : cm) side[u] = level[u] < cut ? 0 : (level[u] == cut ? 1 : 2);
        for (int u : cm) {
            if (side[u] != 1) continue;
            bool hl = false, hr = false;
            for (int v : adj[u]) {
                if (mark[v] != s) continue;
                hl = hl || side[v] == 0;
                hr = hr || side[v] == 2;
            }
            if (!hr) side[u] = 0;
            else if (!hl) side[u] = 2;
        }
        for (int u : cm) (side[u] == 0 ? left : side[u] == 1 ? sep : right).push_back(u);
        if (left.empty() || right.empty()) {
            emit_rcm(cm);
            return;
        }
        order_set(left);
        order_set(right);
        for (int u : sep) out.push_back(u);      // breadth-first order inside the separator
    }
};

struct Symbolic {
    int nb = 0;
    std::vector<uint8_t> Lp, Ap;     // [nb*nb] lower triangle in elimination positions
    std::vector<int32_t> colptr, rowidx, idmap;   // idmap[i*nb+j] = compact id or -1
    std::vector<int> parent;
    int depth = 0;
};

void symbolic(const uint8_t *pattern, int nb, const std::vector<int32_t> &perm, Symbolic &sy) {
    sy.nb = nb;
    sy.Lp.assign((size_t)nb * nb, 0);
    for (int i = 0; i < nb; ++i)
        for (int j = 0; j < i; ++j) {
            const int a = std::max(perm[i], perm[j]), b = std::min(perm[i], perm[j]);
            sy.Lp[(size_t)i * nb + j] = pattern[(size_t)a * nb + b] ? 1 : 0;
        }
    sy.Ap = sy.Lp;
    std::vector<int> rows;
    for (int j = 0; j < nb; ++j) {
        rows.clear();
        for (int i = j + 1; i < nb; ++i)
            if (sy.Lp[(size_t)i * nb + j]) rows.push_back(i);
        for (size_t x = 0; x < rows.size(); ++x)
            for (size_t y = 0; y < x; ++y) sy.Lp[(size_t)rows[x] * nb + rows[y]] = 1;
    }
    sy.colptr.assign(nb + 1, 0);
    sy.rowidx.clear();
    sy.idmap.assign((size_t)nb * nb, -1);
    sy.parent.assign(nb, -1);
    for (int j = 0; j < nb; ++j) {
        sy.colptr[j] = (int32_t)sy.rowidx.size();
        for (int i = j + 1; i < nb; ++i)
            if (sy.Lp[(size_t)i * nb + j]) {
                if (sy.parent[j] < 0) sy.parent[j] = i;
                sy.idmap[(size_t)i * nb + j] = (int32_t)sy.rowidx.size();
                sy.rowidx.push_back(i);
            }
    }
    sy.colptr[nb] = (int32_t)sy.rowidx.size();
    std::vector<int> h(nb, 1);
    sy.depth = 0;
    for (int j = 0; j < nb; ++j) {
        if (sy.parent[j] >= 0) h[sy.parent[j]] = std::max(h[sy.parent[j]], h[j] + 1);
        sy.depth = std::max(sy.depth, h[j]);
    }
}

// simulated costs [us] of the executor's steps on B200 (rough; they only rank schedules)
constexpr double C_LOAD = 0.8, C_PROD = 1.3, C_FACTOR = 8.8, C_TRSM = 3.5, C_LAT = 0.9;   // calibrated on a task trace of the C5 system
constexpr double C_BACK = 3.4, C_ADD = 0.5, C_STORE = 1.0;

struct Built {
    std::vector<DagTask> tasks;
    std::vector<DagEntry> entries;
    int n_fwd = 0, n_part = 0;
    std::vector<int> tid_D, tid_B;    // task index of D(j); of the task writing compact block id (L blocks, then partial sums)
};

// A task whose product list is long (separator blocks collect updates from whole subtrees) would accumulate
// them one after the other on one SM; lists above DAG_SPLIT_MIN products are cut into chunks of DAG_CHUNK, and
// every chunk but the last becomes a PARTIAL task that writes  -sum X Y^T  (and, for a diagonal block, the
// matching piece of the right-hand side) into a scratch block of its own; the main task adds those blocks.
constexpr int DAG_SPLIT_MIN = 10, DAG_CHUNK = 6;

bool build_tasks(const Symbolic &sy, const std::vector<int32_t> &perm, Built &b, size_t max_entries) {
    const int nb = sy.nb, nnzL = (int)sy.rowidx.size();
    b.tasks.clear();
    b.entries.clear();
    b.tid_D.assign(nb, -1);
    b.tid_B.assign(nnzL, -1);
    b.n_part = 0;
    // row lists: for each row i the columns t < i with L_it != 0, ascending
    std::vector<std::vector<int>> rowcols(nb);
    for (int t = 0; t < nb; ++t)
        for (int p = sy.colptr[t]; p < sy.colptr[t + 1]; ++p) rowcols[sy.rowidx[p]].push_back(t);
    std::vector<DagEntry> ev;
    // emits the partial tasks of `ev` (if it is long) and then the main task `m`
    auto emit = [&](DagTask m, bool diag) {
        std::vector<DagEntry> adds;
        size_t first_kept = 0;
        if ((int)ev.size() > DAG_SPLIT_MIN) {
            const int nchunks = ((int)ev.size() + DAG_CHUNK - 1) / DAG_CHUNK;
            for (int c = 0; c + 1 < nchunks; ++c) {
                DagTask pt{};
                pt.type = DAG_PART; pt.j = m.j; pt.i = diag ? 1 : 0; pt.dst = nnzL + b.n_part;
                pt.src_r = -1; pt.src_c = -1; pt.src_T = 0;
                pt.beg = (int32_t)b.entries.size();
                for (int e = c * DAG_CHUNK; e < (c + 1) * DAG_CHUNK; ++e) b.entries.push_back(ev[e]);
                pt.end = (int32_t)b.entries.size();
                b.tid_B.push_back((int)b.tasks.size());
                b.tasks.push_back(pt);
                adds.push_back(DagEntry{pt.dst, -1, b.n_part, 0});
                ++b.n_part;
            }
            first_kept = (size_t)(nchunks - 1) * DAG_CHUNK;
        }
        m.beg = (int32_t)b.entries.size();
        for (const DagEntry &a : adds) b.entries.push_back(a);
        for (size_t e = first_kept; e < ev.size(); ++e) b.entries.push_back(ev[e]);
        m.end = (int32_t)b.entries.size();
        b.tasks.push_back(m);
    };
    for (int j = 0; j < nb; ++j) {
        DagTask d{};
        d.type = DAG_DIAG; d.j = j; d.i = -1; d.dst = j;
        d.src_r = perm[j]; d.src_c = perm[j]; d.src_T = 0;
        ev.clear();
        for (int t : rowcols[j]) {
            const int32_t id = sy.idmap[(size_t)j * nb + t];
            ev.push_back(DagEntry{id, id, t, 0});
        }
        emit(d, true);
        b.tid_D[j] = (int)b.tasks.size() - 1;
        for (int p = sy.colptr[j]; p < sy.colptr[j + 1]; ++p) {
            const int i = sy.rowidx[p];
            DagTask o{};
            o.type = DAG_OFF; o.j = j; o.i = i; o.dst = p;
            if (sy.Ap[(size_t)i * nb + j]) {
                if (perm[i] > perm[j]) { o.src_r = perm[i]; o.src_c = perm[j]; o.src_T = 0; }
                else { o.src_r = perm[j]; o.src_c = perm[i]; o.src_T = 1; }
            } else {
                o.src_r = -1; o.src_c = -1; o.src_T = 0;
            }
            ev.clear();
            // common columns of rows i and j (both ascending)
            const std::vector<int> &ri = rowcols[i], &rj = rowcols[j];
            size_t x = 0, y = 0;
            while (x < ri.size() && y < rj.size()) {
                if (ri[x] < rj[y]) ++x;
                else if (ri[x] > rj[y]) ++y;
                else {
                    const int t = ri[x];
                    ev.push_back(DagEntry{sy.idmap[(size_t)i * nb + t], sy.idmap[(size_t)j * nb + t], t, 0});
                    ++x; ++y;
                }
            }
            emit(o, false);
            b.tid_B[p] = (int)b.tasks.size() - 1;
            if (b.entries.size() > max_entries) return false;
        }
    }
    b.n_fwd = (int)b.tasks.size();
    // backward tasks, last column first
    for (int j = nb - 1; j >= 0; --j) {
        DagTask k{};
        k.type = DAG_BACK; k.j = j; k.i = -1; k.dst = j; k.src_r = perm[j]; k.src_c = -1;
        k.beg = (int32_t)b.entries.size();
        for (int p = sy.colptr[j]; p < sy.colptr[j + 1]; ++p) b.entries.push_back(DagEntry{p, sy.rowidx[p], 0, 0});
        k.end = (int32_t)b.entries.size();
        b.tasks.push_back(k);
    }
    return true;
}

struct Sched {
    std::vector<int32_t> ptr, list;
    double makespan = 0;
};

// Event-driven list scheduling of the forward tasks over G CTAs. A CTA runs its list in order and spins on
// the flags of each product's operands, so a task may be started before its last operands exist; it is
// handed out once everything except the operands of its LAST column (and, for an off-diagonal task, the
// inverse of its diagonal block) is there. Every task is assigned after all tasks it waits for, so the
// assignment order is a topological order and the per-CTA lists, being subsequences of it, cannot deadlock.
Sched schedule_forward(const Built &b, int G) {
    const int nt = b.n_fwd;
    std::vector<std::vector<int>> deps(nt), users(nt);
    for (int k = 0; k < nt; ++k) {
        const DagTask &t = b.tasks[k];
        for (int e = t.beg; e < t.end; ++e) {
            deps[k].push_back(b.tid_B[b.entries[e].a]);
            if (b.entries[e].b >= 0 && b.entries[e].b != b.entries[e].a) deps[k].push_back(b.tid_B[b.entries[e].b]);
        }
        if (t.type == DAG_OFF) deps[k].push_back(b.tid_D[t.j]);
        for (int d : deps[k]) users[d].push_back(k);
    }
    // priority: longest path to the end with nominal durations
    std::vector<double> prio(nt, 0.0);
    for (int k = nt - 1; k >= 0; --k) {
        const DagTask &t = b.tasks[k];
        double dur = C_LOAD + C_PROD * (t.end - t.beg) + (t.type == DAG_DIAG ? C_FACTOR : t.type == DAG_OFF ? C_TRSM : C_STORE);
        double best = 0.0;
        for (int u : users[k]) best = std::max(best, prio[u]);
        prio[k] = dur + best + C_LAT;
    }
    std::vector<double> fin(nt, -1.0), start(nt, -1.0);
    std::vector<int> waiting(nt, 0);
    auto ready_at = [&](const DagEntry &e) {
        double r = fin[b.tid_B[e.a]];
        if (e.b >= 0) r = std::max(r, fin[b.tid_B[e.b]]);
        return r + C_LAT;
    };
    typedef std::pair<double, int> PD;
    std::priority_queue<PD, std::vector<PD>, std::greater<PD>> pending;      // (eligible time, task)
    std::priority_queue<PD> ready;                                            // (priority, task)
    auto eligible_time = [&](int k) {
        const DagTask &t = b.tasks[k];
        double el = 0.0;
        if (t.end > t.beg) {
            const int tlast = b.entries[t.end - 1].t;
            for (int e = t.beg; e < t.end; ++e) {
                if (b.entries[e].b >= 0 && b.entries[e].t == tlast) continue;
                el = std::max(el, ready_at(b.entries[e]));
            }
        }
        if (t.type == DAG_OFF) el = std::max(el, start[b.tid_D[t.j]]);
        return el;
    };
    for (int k = 0; k < nt; ++k) {
        waiting[k] = (int)deps[k].size();
        if (!waiting[k]) pending.push(PD(0.0, k));
    }
    std::priority_queue<PD, std::vector<PD>, std::greater<PD>> ctas;          // (free time, cta)
    for (int c = 0; c < G; ++c) ctas.push(PD(0.0, c));
    std::vector<std::vector<int32_t>> lists(G);
    int assigned = 0;
    double makespan = 0.0;
    while (assigned < nt) {
        PD c = ctas.top();
        ctas.pop();
        double T = c.first;
        while (!pending.empty() && pending.top().first <= T) {
            ready.push(PD(prio[pending.top().second], pending.top().second));
            pending.pop();
        }
        if (ready.empty()) {
            if (pending.empty()) break;                      // cannot happen for a consistent graph
            ctas.push(PD(pending.top().first, c.second));
            continue;
        }
        const int k = ready.top().second;
        ready.pop();
        const DagTask &t = b.tasks[k];
        start[k] = T;
        double tm = T + C_LOAD;
        for (int e = t.beg; e < t.end; ++e)
            tm = std::max(tm, ready_at(b.entries[e])) + (b.entries[e].b >= 0 ? C_PROD : C_ADD);
        if (t.type == DAG_DIAG) tm += C_FACTOR;
        else if (t.type == DAG_OFF) tm = std::max(tm, fin[b.tid_D[t.j]] + C_LAT) + C_TRSM;
        else tm += C_STORE;
        fin[k] = tm;
        makespan = std::max(makespan, tm);
        lists[c.second].push_back(k);
        ++assigned;
        ctas.push(PD(tm, c.second));
        for (int u : users[k])
            if (--waiting[u] == 0) pending.push(PD(eligible_time(u), u));
    }
    Sched s;
    s.makespan = assigned == nt ? makespan : -1.0;
    s.ptr.assign(G + 1, 0);
    for (int c = 0; c < G; ++c) {
        s.ptr[c] = (int32_t)s.list.size();
        s.list.insert(s.list.end(), lists[c].begin(), lists[c].end());
    }
    s.ptr[G] = (int32_t)s.list.size();
    return s;
}

// backward tasks: x_j needs x_i of every row i of column j. A task streams the blocks of its column into shared
// memory BEFORE it waits for those x_i, so it should be started long before its turn: the columns are sorted by
// their level below the roots of the elimination tree (a topological order) and dealt round-robin -- with
// nb <= G every task has a CTA of its own and the chain pays one flag latency + one small mat-vec per level.
Sched schedule_backward(const Built &b, const Symbolic &sy, int G) {
    const int nb = sy.nb, base = b.n_fwd;
    auto tid = [&](int j) { return base + (nb - 1 - j); };
    std::vector<int> lev(nb, 0), order(nb);
    int maxlev = 0;
    for (int j = nb - 1; j >= 0; --j) {
        for (int p = sy.colptr[j]; p < sy.colptr[j + 1]; ++p) lev[j] = std::max(lev[j], lev[sy.rowidx[p]] + 1);
        maxlev = std::max(maxlev, lev[j]);
    }
    for (int j = 0; j < nb; ++j) order[j] = j;
    std::sort(order.begin(), order.end(), [&](int x, int y) { return lev[x] != lev[y] ? lev[x] < lev[y] : x > y; });
    std::vector<std::vector<int32_t>> lists(G);
    for (int k = 0; k < nb; ++k) lists[k % G].push_back(tid(order[k]));
    Sched s;
    s.makespan = (maxlev + 1) * (C_BACK + C_LAT) * std::max(1.0, (double)nb / (4.0 * G));
    s.ptr.assign(G + 1, 0);
    for (int c = 0; c < G; ++c) {
        s.ptr[c] = (int32_t)s.list.size();
        s.list.insert(s.list.end(), lists[c].begin(), lists[c].end());
    }
    s.ptr[G] = (int32_t)s.list.size();
    return s;
}

bool try_order(const uint8_t *pattern, int nb, int G, const std::vector<int32_t> &perm, int kind, CholDagHost &best,
               bool &have) {
    Symbolic sy;
    symbolic(pattern, nb, perm, sy);
    Built b;
    if (!build_tasks(sy, perm, b, (size_t)3 << 20)) return false;
    Sched f = schedule_forward(b, G);
    if (f.makespan < 0) return false;
    Sched k = schedule_backward(b, sy, G);
    if (k.makespan < 0) return false;
    if (have && f.makespan + k.makespan >= best.sim_fwd_us + best.sim_bwd_us) return true;
    have = true;
    best.nb = nb;
    best.G = G;
    best.nnzL = (int)sy.rowidx.size();
    best.n_part = b.n_part;
    best.order_kind = kind;
    best.depth = sy.depth;
    best.sim_fwd_us = f.makespan;
    best.sim_bwd_us = k.makespan;
    best.perm = perm;
    best.colptr = sy.colptr;
    best.rowidx = sy.rowidx;
    best.tasks = b.tasks;
    best.n_fwd = b.n_fwd;
    best.entries = b.entries;
    best.fwd_ptr = f.ptr; best.fwd_list = f.list;
    best.bwd_ptr = k.ptr; best.bwd_list = k.list;
    return true;
}

}  // namespace

int chol_dag_plan(const uint8_t *pattern, int nb, int G, int order_mode, CholDagHost &out) {
    if (!pattern || nb < 1 || G < 1) return 1;
    bool have = false;
    std::vector<int32_t> nat(nb);
    for (int i = 0; i < nb; ++i) nat[i] = i;
    if (order_mode == 0 || order_mode == 1 || nb < 4) try_order(pattern, nb, G, nat, 0, out, have);
    if ((order_mode == 0 || order_mode == 2) && nb >= 4) {
        Graph adj(nb);
        size_t nnz = 0;
        for (int i = 0; i < nb; ++i)
            for (int j = 0; j < i; ++j)
                if (pattern[(size_t)i * nb + j]) { adj[i].push_back(j); adj[j].push_back(i); ++nnz; }
        const bool dense = nnz * 2 >= (size_t)nb * (nb - 1) * 9 / 10;      // nothing to dissect
        if (!dense || order_mode == 2) {
            const int leaves[3] = {4, 8, 16};
            for (int v = 0; v < 3; ++v) {
                Orderer o(adj, leaves[v]);
                std::vector<int> all(nb);
                for (int i = 0; i < nb; ++i) all[i] = i;
                o.order_set(all);
                if ((int)o.out.size() != nb) continue;
                std::vector<int32_t> perm(o.out.begin(), o.out.end());
                try_order(pattern, nb, G, perm, 1 + v, out, have);
            }
        }
    }
    return have ? 0 : 1;
}

}  // namespace sfm
