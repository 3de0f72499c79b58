"""CPU-only: the host plan of the structure-aware K11 (csrc/chol_plan.cpp) replayed in numpy.

The plan (elimination order, symbolic factorisation, left-looking task graph, per-CTA schedule) is pure
host code; sfm_chol_plan_export hands it out as arrays. Here every task is executed in numpy exactly as
the kernel would (same operands, same order of products), with the per-CTA lists run by a cooperative
round-robin that only lets a task proceed when its flags are up -- so a schedule that could deadlock on
the GPU deadlocks here too -- and the result is checked against LAPACK. What is left for the GPU tests
is the device code itself (tests/test_gpu_ba.py::test_chol_structured_*)."""
import ctypes as C
import zlib

import numpy as np
import pytest

from sfm_b200 import _lib

NB = 48


def export_plan(pattern, n_ctas=148, order_mode=0):
    lib = _lib.load()
    nb = pattern.shape[0]
    pat = np.ascontiguousarray(pattern, np.uint8)
    sizes = np.zeros(9, np.int64)
    sim = np.zeros(2)
    p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
    rc = lib.sfm_chol_plan_export(p(pat), nb, n_ctas, order_mode, p(sizes), p(sim), *([None] * 9))
    assert rc == 0, rc
    nnz, ntask, nfwd, nent, kind, depth, lf, lb, npart = (int(v) for v in sizes)
    perm = np.zeros(nb, np.int32)
    colptr = np.zeros(nb + 1, np.int32)
    rowidx = np.zeros(max(nnz, 1), np.int32)
    tasks = np.zeros((ntask, 12), np.int32)
    entries = np.zeros((max(nent, 1), 4), np.int32)
    fwd_ptr, bwd_ptr = np.zeros(n_ctas + 1, np.int32), np.zeros(n_ctas + 1, np.int32)
    fwd_list, bwd_list = np.zeros(max(lf, 1), np.int32), np.zeros(max(lb, 1), np.int32)
    rc = lib.sfm_chol_plan_export(p(pat), nb, n_ctas, order_mode, p(sizes), p(sim), p(perm), p(colptr), p(rowidx),
                                  p(tasks), p(entries), p(fwd_ptr), p(fwd_list), p(bwd_ptr), p(bwd_list))
    assert rc == 0
    return dict(nb=nb, nnz=nnz, npart=npart, nfwd=nfwd, kind=kind, depth=depth, sim=sim, perm=perm, colptr=colptr,
                rowidx=rowidx[:nnz], tasks=tasks, entries=entries[:nent], fwd_ptr=fwd_ptr, fwd_list=fwd_list[:lf],
                bwd_ptr=bwd_ptr, bwd_list=bwd_list[:lb])


def replay(plan, S, g):
    """Run the task lists like the kernel: lower triangle of S (padded to nb*48) is the only input."""
    nb, perm = plan["nb"], plan["perm"]
    tasks, ent = plan["tasks"], plan["entries"]
    blk = lambda r, c: S[r * NB:(r + 1) * NB, c * NB:(c + 1) * NB]
    Lc = [None] * (plan["nnz"] + plan["npart"])
    pvec = [None] * plan["npart"]
    Linv, y, x = [None] * nb, [None] * nb, [None] * nb
    xout = np.zeros(nb * NB)

    def run_fwd(k):
        """returns False when an operand is not there yet"""
        ty, j, i, dst, sr, sc, sT, beg, end = (int(v) for v in tasks[k, :9])
        for e in range(beg, end):
            if Lc[ent[e, 0]] is None or (ent[e, 1] >= 0 and Lc[ent[e, 1]] is None):
                return False
        if ty in (0, 3):
            rhs = ty == 0 or i == 1
            if ty == 0:
                Cm = np.tril(blk(sr, sc)).copy()
                Cm = Cm + np.tril(Cm, -1).T
                v = g[sr * NB:(sr + 1) * NB].copy()
            else:
                assert sr < 0 and dst >= plan["nnz"]
                Cm, v = np.zeros((NB, NB)), np.zeros(NB)
            for e in range(beg, end):
                X = Lc[ent[e, 0]]
                if ent[e, 1] < 0:                      # block of a partial task
                    Cm += X
                    if rhs:
                        v += pvec[ent[e, 2]]
                    continue
                Cm -= X @ Lc[ent[e, 1]].T
                if rhs:
                    assert ent[e, 0] == ent[e, 1]
                    v -= X @ y[ent[e, 2]]
            if ty == 3:
                Lc[dst] = Cm
                pvec[dst - plan["nnz"]] = v if rhs else None
                return True
            L = np.linalg.cholesky(Cm)
            Linv[j] = np.linalg.inv(L)
            y[j] = Linv[j] @ v
        else:
            if Linv[j] is None:
                return False
            if sr >= 0:
                assert sr > sc
                Cm = blk(sr, sc).T.copy() if sT else blk(sr, sc).copy()
            else:
                Cm = np.zeros((NB, NB))
            for e in range(beg, end):
                Cm = Cm + Lc[ent[e, 0]] if ent[e, 1] < 0 else Cm - Lc[ent[e, 0]] @ Lc[ent[e, 1]].T
            Lc[dst] = Cm @ Linv[j].T
        return True

    def run_bwd(k):
        ty, j, i, dst, sr, sc, sT, beg, end = (int(v) for v in tasks[k, :9])
        assert ty == 2
        for e in range(beg, end):
            if x[ent[e, 1]] is None:
                return False
        w = y[j].copy()
        for e in range(beg, end):
            w -= Lc[ent[e, 0]].T @ x[ent[e, 1]]
        x[j] = Linv[j].T @ w
        xout[sr * NB:(sr + 1) * NB] = x[j]
        return True

    for ptr, lst, run in ((plan["fwd_ptr"], plan["fwd_list"], run_fwd), (plan["bwd_ptr"], plan["bwd_list"], run_bwd)):
        pos = ptr[:-1].copy()
        remaining = int(ptr[-1] - ptr[0])
        while remaining:
            progressed = False
            for c in range(len(pos)):
                while pos[c] < ptr[c + 1] and run(int(lst[pos[c]])):
                    pos[c] += 1
                    remaining -= 1
                    progressed = True
            assert progressed, "the per-CTA task lists deadlock"
    return xout


def spd_from_pattern(pattern, rng, n=None):
    """SPD matrix whose 48-blocks are non-zero exactly where `pattern` says (diagonally dominant)."""
    nb = pattern.shape[0]
    N = nb * NB
    S = np.zeros((N, N))
    for i in range(nb):
        for j in range(i):
            if pattern[i, j]:
                S[i * NB:(i + 1) * NB, j * NB:(j + 1) * NB] = rng.normal(0, 1, (NB, NB))
    S = S + S.T
    S += np.diag(np.abs(S).sum(1) + rng.uniform(1, 2, N))
    if n is not None and n < N:                       # identity padding like the BA system
        S[n:, :] = 0
        S[:, n:] = 0
        S[np.arange(n, N), np.arange(n, N)] = 1.0
    return S


def ring_pattern(nb, bw, cyclic=True):
    pat = np.zeros((nb, nb), np.uint8)
    for i in range(nb):
        for j in range(i + 1):
            d = i - j
            if d <= bw or (cyclic and nb - d <= bw):
                pat[i, j] = 1
    return pat


def random_pattern(nb, density, rng):
    pat = np.tril((rng.random((nb, nb)) < density).astype(np.uint8))
    pat[np.arange(nb), np.arange(nb)] = 1
    return pat


def arrow_pattern(nb, bw, head):
    pat = ring_pattern(nb, bw, cyclic=False)
    pat[nb - head:, :] = 1
    return np.tril(pat)


CASES = [
    ("dense7", lambda rng: np.tril(np.ones((7, 7), np.uint8))),
    ("band12", lambda rng: ring_pattern(12, 2, cyclic=False)),
    ("ring40", lambda rng: ring_pattern(40, 3)),
    ("ring125", lambda rng: ring_pattern(125, 5)),
    ("arrow30", lambda rng: arrow_pattern(30, 2, 3)),
    ("random25", lambda rng: random_pattern(25, 0.12, rng)),
    ("two_components", lambda rng: np.kron(np.eye(2, dtype=np.uint8), ring_pattern(9, 2, cyclic=False))),
    ("diag5", lambda rng: np.eye(5, dtype=np.uint8)),
    ("single", lambda rng: np.ones((1, 1), np.uint8)),
]


@pytest.mark.parametrize("name,make", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("order_mode", [0, 1, 2])
def test_plan_replay_matches_lapack(name, make, order_mode):
    rng = np.random.default_rng(zlib.crc32(name.encode()))
    pat = make(rng)
    nb = pat.shape[0]
    small = nb <= 40
    plan = export_plan(pat, n_ctas=148 if not small else 5, order_mode=order_mode)
    assert sorted(plan["perm"].tolist()) == list(range(nb))
    if not small and order_mode != 0:
        return                                         # the numeric replay of the big case runs once (auto order)
    S = spd_from_pattern(pat, rng, n=nb * NB - 7 if nb > 1 else None)
    g = rng.normal(0, 1, nb * NB)
    x = replay(plan, np.tril(S), g)
    ref = np.linalg.solve(S, g)
    assert np.linalg.norm(x - ref) <= 1e-10 * np.linalg.norm(ref)


def test_ring_plan_is_shallow():
    """C5's reduced camera system: 125 blocks, cyclic band of 5. Natural order chains all 125 panels; the plan
    must cut the elimination tree to a fraction of that and predict a much shorter schedule."""
    pat = ring_pattern(125, 5)
    nat = export_plan(pat, order_mode=1)
    auto = export_plan(pat, order_mode=0)
    assert nat["depth"] == 125
    assert auto["kind"] >= 1 and auto["depth"] <= 45, (auto["kind"], auto["depth"])
    assert auto["sim"].sum() < 0.45 * nat["sim"].sum(), (auto["sim"], nat["sim"])
    # every forward task appears exactly once over the CTA lists, backward likewise
    assert sorted(auto["fwd_list"].tolist()) == list(range(auto["nfwd"]))
    assert sorted(auto["bwd_list"].tolist()) == list(range(auto["nfwd"], auto["nfwd"] + 125))


def test_dense_plan_keeps_natural_order():
    pat = np.tril(np.ones((20, 20), np.uint8))
    plan = export_plan(pat)
    assert plan["kind"] == 0 and plan["depth"] == 20 and plan["nnz"] == 190
