"""Two-view pose front end: oracle against the frozen reference outputs, the device
arithmetic compiled for the host (tests/hostsim) against the same, and -- marked gpu -- the kernels
through the C ABI. Configurations are compared as a SET (their list order is LAPACK's sign choice in
the reference); votes, selected pose and points exactly / to 1e-9."""
import ctypes as C
import io
from contextlib import redirect_stdout

import numpy as np
import pytest

import hostsim

from conftest import golden, rel_err
from oracle import sfm_oracle as so


def _match_configs(C_got, R_got, C_ref, R_ref, tol=1e-9):
    """permutation p with got[i] == ref[p[i]]; asserts it is a bijection"""
    perm = []
    for i in range(4):
        d = [max(np.abs(C_got[i] - C_ref[j]).max(), np.abs(R_got[i] - R_ref[j]).max()) for j in range(4)]
        j = int(np.argmin(d))
        assert d[j] < tol, (i, d)
        perm.append(j)
    assert sorted(perm) == [0, 1, 2, 3]
    return perm


@pytest.mark.parametrize("tag", ["a", "b"])
def test_oracle_matches_reference(tag):
    g = golden("twoview.npz")
    E = so.essential_from_fundamental(g[f"F_{tag}"], g[f"K_{tag}"])
    assert np.array_equal(E, g[f"E_{tag}"])
    Cs, Rs = so.extract_camera_pose(E)
    assert np.array_equal(Cs, g[f"Cset_{tag}"]) and np.array_equal(Rs, g[f"Rset_{tag}"])
    Xs = [so.linear_triangulation(g[f"K_{tag}"], np.zeros(3), np.eye(3), c, r, g[f"x1_{tag}"], g[f"x2_{tag}"])
          for c, r in zip(Cs, Rs)]
    assert rel_err(np.array(Xs), g[f"Xset_{tag}"]) < 1e-12
    Cb, Rb, Xb, _, _ = so.disambiguate_camera_pose(Cs, Rs, list(g[f"Xset_{tag}"]))
    assert np.array_equal(Cb, g[f"C_{tag}"]) and np.array_equal(Rb, g[f"R_{tag}"]) and np.array_equal(Xb, g[f"X_{tag}"])


@pytest.mark.parametrize("tag", ["a", "b"])
def test_hostsim_configs_match_reference_as_a_set(tag):
    g = golden("twoview.npz")
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    F, K = np.ascontiguousarray(g[f"F_{tag}"]), np.ascontiguousarray(g[f"K_{tag}"])
    cfg, P = np.zeros((4, 12)), np.zeros((5, 12))
    hostsim.load().hs_two_view_configs(vp(F), vp(K), vp(cfg), vp(P))
    perm = _match_configs(cfg[:, :3], cfg[:, 3:].reshape(4, 3, 3), g[f"Cset_{tag}"], g[f"Rset_{tag}"])
    # structure of the list is the reference's: (R1,t) (R1,-t) (R2,t) (R2,-t)
    assert np.array_equal(cfg[0, 3:], cfg[1, 3:]) and np.array_equal(cfg[2, 3:], cfg[3, 3:])
    assert np.allclose(cfg[0, :3], -cfg[1, :3]) and np.allclose(cfg[2, :3], -cfg[3, :3])
    # projection matrices K R [I | -C]
    for i in range(4):
        R, Cc = cfg[i, 3:].reshape(3, 3), cfg[i, :3]
        assert rel_err(P[i + 1].reshape(3, 4), K @ np.hstack([R, -R @ Cc.reshape(3, 1)])) < 1e-13
    assert rel_err(P[0].reshape(3, 4), K @ np.hstack([np.eye(3), np.zeros((3, 1))])) < 1e-15
    assert len(set(perm)) == 4


def test_hostsim_svd3():
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    rng = np.random.default_rng(0)
    for k in range(50):
        M = rng.normal(size=(3, 3))
        if k % 3 == 0:                               # rank 2, like an essential matrix
            u, s, vt = np.linalg.svd(M)
            M = u @ np.diag([s[0], s[1], 0.0]) @ vt
        M = np.ascontiguousarray(M)
        U, s, V = np.zeros((3, 3)), np.zeros(3), np.zeros((3, 3))
        hostsim.load().hs_svd3(vp(M), vp(U), vp(s), vp(V))
        assert np.allclose(s, np.linalg.svd(M, compute_uv=False), atol=1e-13)
        assert np.allclose(U @ np.diag(s) @ V.T, M, atol=1e-13)
        assert np.allclose(U.T @ U, np.eye(3), atol=1e-13) and np.allclose(V.T @ V, np.eye(3), atol=1e-13)


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["a", "b"])
def test_gpu_two_view_pose_matches_reference(cuda_device, tag):
    from sfm_b200 import two_view
    g = golden("twoview.npz")
    out = two_view.two_view_pose(g[f"F_{tag}"], g[f"K_{tag}"], g[f"x1_{tag}"], g[f"x2_{tag}"])
    perm = _match_configs(out["Cset"], out["Rset"], g[f"Cset_{tag}"], g[f"Rset_{tag}"])
    ref_counts = so.cheirality_counts(g[f"Cset_{tag}"], g[f"Rset_{tag}"], g[f"Xset_{tag}"])
    for i in range(4):
        ref_X = g[f"Xset_{tag}"][perm[i]]
        err = np.linalg.norm(out["Xset"][i] - ref_X, axis=1) / np.maximum(np.linalg.norm(ref_X, axis=1), 1e-12)
        assert np.median(err) < 1e-9 and np.mean(err < 1e-6) > 0.99        # DLT of outliers is ill-conditioned
        assert abs(int(out["counts"][i]) - int(ref_counts[perm[i]])) <= max(1, len(ref_X) // 200)
    assert perm[out["best"]] == int(np.argmax(ref_counts))
    assert rel_err(out["C"], g[f"C_{tag}"]) < 1e-9 and rel_err(out["R"], g[f"R_{tag}"]) < 1e-9
    sel = np.linalg.norm(out["X"] - g[f"X_{tag}"], axis=1) / np.linalg.norm(g[f"X_{tag}"], axis=1)
    assert np.median(sel) < 1e-9 and np.mean(sel < 1e-6) > 0.99


@pytest.mark.gpu
def test_gpu_dropins(cuda_device):
    from sfm_b200 import two_view
    g = golden("twoview.npz")
    E = two_view.EssentialMatrixFromFundamentalMatrix(g["F_a"], g["K_a"])
    assert np.array_equal(E, g["E_a"])
    Cs, Rs = two_view.ExtractCameraPose(E)
    _match_configs(Cs, Rs, g["Cset_a"], g["Rset_a"])
    buf = io.StringIO()
    with redirect_stdout(buf):
        Cb, Rb, Xb = two_view.DisambiguateCameraPose(g["Cset_a"], g["Rset_a"], list(g["Xset_a"]))
    assert np.array_equal(Cb, g["C_a"]) and np.array_equal(Rb, g["R_a"]) and np.array_equal(Xb, g["X_a"])
    assert buf.getvalue().startswith("Selected camera pose configuration 1 with 400 points")
    # array input and the empty / ragged conventions of DisambiguateCameraPose.py:38-62,84-88
    with redirect_stdout(io.StringIO()):
        Cb2, _, _ = two_view.DisambiguateCameraPose(g["Cset_b"], g["Rset_b"], g["Xset_b"])
        assert np.array_equal(Cb2, g["C_b"])
        ragged = [g["Xset_b"][0][:0], g["Xset_b"][1][:10], g["Xset_b"][2][:50], g["Xset_b"][3]]
        ref = so.disambiguate_camera_pose(g["Cset_b"], g["Rset_b"], ragged)
        got = two_view.DisambiguateCameraPose(g["Cset_b"], g["Rset_b"], ragged)
        assert np.array_equal(got[0], ref[0]) and np.array_equal(got[2], ref[2])
    buf = io.StringIO()
    with redirect_stdout(buf):
        got = two_view.DisambiguateCameraPose(g["Cset_a"], g["Rset_a"], [np.zeros((0, 3))] * 4)
    assert "Warning: No valid camera pose found" in buf.getvalue() and np.array_equal(got[0], g["Cset_a"][0])


@pytest.mark.gpu
def test_gpu_two_view_full_size_properties(cuda_device):
    """1M correspondences: the selected configuration puts (almost) every inlier in front of both cameras,
    the other three almost none; votes equal a recount from the returned points."""
    import torch
    from sfm_b200 import two_view, device
    from oracle import scenes
    sc = scenes.two_view(n=1_000_000, outlier_frac=0.0, seed=77)
    F = so.estimate_fundamental_matrix(sc["x1"][:2000], sc["x2"][:2000]) if hasattr(so, "estimate_fundamental_matrix") else None
    if F is None:
        pytest.skip("oracle has no n-point F")
    r = two_view.two_view_pose_device(device.to_device(np.ascontiguousarray(F).reshape(9)), sc["K"],
                                      device.to_device(sc["x1"]), device.to_device(sc["x2"]))
    counts, best = r.counts.cpu().numpy(), r.best.cpu().numpy()
    assert best[1] == counts.max() and counts[best[0]] > 0.999 * 1_000_000
    cfg = r.cfg.cpu().numpy()
    X = r.Xall[int(best[0])].cpu().numpy()
    Cc, R = cfg[best[0], :3], cfg[best[0], 3:].reshape(3, 3)
    z2 = ((X - Cc) @ R.T)[:, 2]
    assert int(np.sum((X[:, 2] > 0) & (z2 > 0))) == int(best[1])
