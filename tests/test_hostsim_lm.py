"""Kernel arithmetic of K5/K7 (host build of the SFM_HD code) against the oracle / reference goldens.
Tolerances: north_star 1e-6 relative on points/poses; asserted tighter against the CONVERGED oracle."""
import ctypes as C

import numpy as np
import pytest
from scipy.spatial.transform import Rotation

import hostsim
from conftest import golden, rel_err
from oracle import sfm_oracle as so


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def hs_nltri(hs, P1, P2, x1, x2, X0, max_iter=50, tol=1e-8):
    n = len(x1)
    X, cost, st = np.zeros((n, 3)), np.zeros(n), np.zeros(n, np.int32)
    hs.hs_nltri(_p(np.ascontiguousarray(P1)), _p(np.ascontiguousarray(P2)), _p(np.ascontiguousarray(x1)),
                _p(np.ascontiguousarray(x2)), _p(np.ascontiguousarray(X0)), n, max_iter, C.c_double(tol),
                _p(X), _p(cost), _p(st))
    return X, cost, st


def test_rotvec_matches_scipy():
    hs = hostsim.load()
    rng = np.random.default_rng(0)
    for r in [np.zeros(3), np.array([1e-9, 0, 0]), np.array([0.003, -0.004, 0.001]), rng.normal(0, 1, 3),
              np.array([3.0, 0.5, -0.2]), np.array([0.0, np.pi - 1e-6, 0.0])]:
        R = np.zeros(9)
        hs.hs_rotvec_to_R(_p(np.ascontiguousarray(r, dtype=float)), _p(R))
        assert np.max(np.abs(R.reshape(3, 3) - Rotation.from_rotvec(r).as_matrix())) < 2e-15


def test_pnp_jacobian_matches_central_differences_of_reference_loss():
    hs = hostsim.load()
    g = golden("pnp.npz")
    X, x, K = g["X_a"][:40].copy(), g["x_a"][:40].copy(), g["K"]
    for rv in [np.zeros(3), np.array([2.4e-5, 0, 0]), np.array([0.3, -0.5, 0.2]), np.array([2.0, 1.8, -1.4])]:
        rt = np.hstack([rv, [0.1, -0.2, 0.3]])
        r, J = np.zeros(80), np.zeros((80, 6))
        hs.hs_pnp_residual_jac(_p(K), _p(rt), _p(X), _p(x), 40, _p(r), _p(J))
        assert np.max(np.abs(r - so.pnp_loss(rt, X, x, K))) < 1e-9
        Jn = np.zeros((80, 6))
        for k in range(6):
            h = 1e-6
            e = np.zeros(6)
            e[k] = h
            Jn[:, k] = (so.pnp_loss(rt + e, X, x, K) - so.pnp_loss(rt - e, X, x, K)) / (2 * h)
        assert np.max(np.abs(J - Jn)) / np.max(np.abs(Jn)) < 1e-7


def test_nltri_matches_reference():
    hs = hostsim.load()
    g = golden("nltri.npz")
    P1 = so.projection_matrix(g["K"], g["C1"], g["R1"])
    P2 = so.projection_matrix(g["K"], g["C2"], g["R2"])
    for x0, Xref in ((g["x0"], g["X"]), (g["x0b"], g["Xb"])):
        X, cost, st = hs_nltri(hs, P1, P2, g["x1"], g["x2"], x0)
        rel = np.linalg.norm(X - Xref, axis=1) / np.linalg.norm(Xref, axis=1)
        assert rel.max() < 1e-6, rel.max()                  # north_star tolerance
        assert np.median(rel) < 1e-9, np.median(rel)        # same trajectory as MINPACK
        status = st >> 16
        assert np.all(np.isin(status, [1, 2, 3, 4])), np.unique(status)
        # never worse than the reference's own stopping point
        cref = so.tri_cost(g["K"], g["C1"], g["R1"], g["C2"], g["R2"], g["x1"], g["x2"], Xref)
        assert np.all(cost <= cref * (1 + 1e-6) + 1e-18)
    # against the converged oracle the agreement is much tighter
    sl = slice(0, 40)
    Xt = so.nonlinear_triangulation(g["K"], g["C1"], g["R1"], g["C2"], g["R2"], g["x1"][sl], g["x2"][sl],
                                    g["x0"][sl], tight=True)
    X, c, _ = hs_nltri(hs, P1, P2, g["x1"][sl], g["x2"][sl], g["x0"][sl], max_iter=2000, tol=1e-15)
    # (the "converged" scipy run still differentiates by forward differences, so its own minimum
    # is only good to ~1e-8; ours uses the analytic Jacobian and must not have a larger cost)
    assert (np.linalg.norm(X - Xt, axis=1) / np.linalg.norm(Xt, axis=1)).max() < 1e-7
    ct = so.tri_cost(g["K"], g["C1"], g["R1"], g["C2"], g["R2"], g["x1"][sl], g["x2"][sl], Xt)
    assert np.all(c <= ct * (1 + 1e-9) + 1e-20)      # equal up to evaluation rounding (~1e-11)


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_nlpnp_matches_reference(tag):
    hs = hostsim.load()
    g = golden("pnp.npz")
    X, x, K = np.ascontiguousarray(g[f"X_{tag}"]), np.ascontiguousarray(g[f"x_{tag}"]), g["K"]
    rt = so.pose_to_params(g[f"C0_{tag}"], g[f"R0_{tag}"]).copy()
    cost, iters = C.c_double(), C.c_int()
    st = hs.hs_nlpnp(_p(K), _p(X), _p(x), len(X), _p(rt), 100, C.c_double(1e-8), C.byref(cost), C.byref(iters))
    assert st in (1, 2, 3, 4)
    Cn, Rn = so.params_to_pose(rt)
    assert rel_err(Cn, g[f"C_{tag}"]) < 1e-6 and rel_err(Rn, g[f"R_{tag}"]) < 1e-6
    cref = 0.5 * np.sum(so.pnp_loss(so.pose_to_params(g[f"C_{tag}"], g[f"R_{tag}"]), X, x, K) ** 2)
    assert cost.value <= cref * (1 + 1e-6)


def test_lintri_matches_reference_golden():
    """N1: the DLT null vector by one-sided Jacobi equals the reference's LinearTriangulation
    (LAPACK dgesdd) on the frozen reference output, x0 of tests/golden/nltri.npz."""
    import ctypes as C
    g = golden("nltri.npz")
    K = g["K"]
    P1 = np.ascontiguousarray(K @ np.hstack([g["R1"], -g["R1"] @ g["C1"].reshape(3, 1)]))
    P2 = np.ascontiguousarray(K @ np.hstack([g["R2"], -g["R2"] @ g["C2"].reshape(3, 1)]))
    x1, x2 = np.ascontiguousarray(g["x1"]), np.ascontiguousarray(g["x2"])
    n = len(x1)
    X = np.zeros((n, 3))
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    hostsim.load().hs_lintri(vp(P1), vp(P2), vp(x1), vp(x2), C.c_int(n), vp(X))
    err = np.linalg.norm(X - g["x0"], axis=1) / np.linalg.norm(g["x0"], axis=1)
    assert err.max() < 1e-9, err.max()
    # and against the oracle restatement on fresh inputs, far and near points included
    rng = np.random.default_rng(1)
    Xw = np.stack([rng.uniform(-3, 3, 200), rng.uniform(-2, 2, 200), rng.uniform(2, 60, 200)], axis=1)
    xa = (P1 @ np.hstack([Xw, np.ones((200, 1))]).T).T
    xb = (P2 @ np.hstack([Xw, np.ones((200, 1))]).T).T
    xa, xb = np.ascontiguousarray(xa[:, :2] / xa[:, 2:]), np.ascontiguousarray(xb[:, :2] / xb[:, 2:])
    ref = so.linear_triangulation(K, g["C1"], g["R1"], g["C2"], g["R2"], xa, xb)
    Y = np.zeros((200, 3))
    hostsim.load().hs_lintri(vp(P1), vp(P2), vp(xa), vp(xb), C.c_int(200), vp(Y))
    assert rel_err(Y, ref) < 1e-9 and rel_err(Y, Xw) < 1e-6
