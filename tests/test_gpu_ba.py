"""GPU parity tests of the bundle-adjustment chain (K8-K12) through the C ABI, against the
reference's own outputs frozen in tests/golden/ba_*.npz and the CPU oracle (oracle/ba_oracle.py).
Tolerances are north_star's: final cost within 1e-6 relative, reprojection RMSE within 1e-3 px."""
import io
from contextlib import redirect_stdout

import numpy as np
import pytest

from conftest import golden, rel_err
from oracle import ba_oracle as bo

pytestmark = pytest.mark.gpu

COST_RTOL = 1e-6
RMSE_ATOL = 1e-3


@pytest.fixture(params=["k10_atomic", "k10g_pairs"])
def schur_path(request, monkeypatch):
    """Both Schur-complement kernels on the same small problems: the L2-atomic K10 (default below 16384
    observations) and the plan-time pair-list gather K10g (default above), forced through the library's
    development switches, which are read at every index build."""
    if request.param == "k10g_pairs":
        monkeypatch.setenv("SFM_SCHUR_PAIRS_MIN", "0")
    else:
        monkeypatch.setenv("SFM_SCHUR_PAIRS", "0")
    return request.param


def _x0(sc):
    from scipy.spatial.transform import Rotation
    cams = np.array([np.hstack([Rotation.from_matrix(R).as_rotvec(), -R @ C]) for R, C in zip(sc["R0"], sc["C0"])])
    return cams, np.array(sc["pts0"], float)


def _rmse(cost, n_obs):
    return np.sqrt(2.0 * cost / n_obs)


@pytest.mark.parametrize("n", [5, 48, 49, 300, 306, 768, 1000, 1105, 2400, 3001])
def test_chol_solve_matches_lapack(cuda_device, n):
    import torch
    from sfm_b200 import _lib, device
    rng = np.random.default_rng(n)
    A = rng.normal(size=(n, n + 5))
    S = A @ A.T + 0.5 * np.eye(n)
    g = rng.normal(size=n)
    lib = _lib.load()
    nbytes = int(lib.sfm_chol_solve_workspace_bytes(n))
    ws = device.empty((nbytes,), torch.uint8)
    d_x = device.empty((n,), torch.float64)
    d_fail = device.empty((1,), torch.int32)
    Sl = np.tril(S) + np.triu(rng.normal(size=(n, n)), 1)       # garbage above the diagonal is never read
    d_S, d_g = device.to_device(Sl), device.to_device(g)          # keep the tensors alive across the call
    _lib.check(lib.sfm_chol_solve(device.ptr(d_S), device.ptr(d_g), n,
                                  device.ptr(d_x), device.ptr(d_fail), device.ptr(ws), nbytes, device.stream()),
               "sfm_chol_solve")
    x = d_x.cpu().numpy()
    assert int(d_fail.cpu()[0]) == 0
    ref = np.linalg.solve(S, g)
    assert rel_err(x, ref) < 1e-9 * max(1.0, np.linalg.cond(S) * 1e-6), rel_err(x, ref)
    # an indefinite matrix is reported, not silently "solved"
    S2 = S.copy()
    S2[n // 2, n // 2] = -1.0
    d_S2 = device.to_device(S2)
    _lib.check(lib.sfm_chol_solve(device.ptr(d_S2), device.ptr(d_g), n,
                                  device.ptr(d_x), device.ptr(d_fail), device.ptr(ws), nbytes, device.stream()),
               "sfm_chol_solve")
    assert int(d_fail.cpu()[0]) == 1


def _structured_case(kind, rng):
    """SPD systems with the block structures the structure-aware K11 meets: n, dense lower-triangular S."""
    from test_chol_plan import NB, arrow_pattern, random_pattern, ring_pattern, spd_from_pattern
    if kind.startswith("dense"):
        n = int(kind[5:])
        A = rng.normal(size=(n, n + 5))
        return n, A @ A.T + 0.5 * np.eye(n)
    pat = {"band": lambda: ring_pattern(14, 2, cyclic=False), "ring": lambda: ring_pattern(40, 3),
           "ring125": lambda: ring_pattern(125, 5), "arrow": lambda: arrow_pattern(30, 2, 3),
           "random": lambda: random_pattern(25, 0.12, rng),
           "blocks": lambda: np.kron(np.eye(3, dtype=np.uint8), ring_pattern(6, 2, cyclic=False))}[kind]()
    n = pat.shape[0] * NB - 5                                      # not a multiple of 48: identity padding
    S = spd_from_pattern(pat, rng)[:n, :n]
    return n, S


@pytest.mark.parametrize("kind", ["dense5", "dense48", "dense300", "dense1105", "band", "ring", "ring125", "arrow",
                                  "random", "blocks"])
@pytest.mark.parametrize("order_mode", [0, 1, 2])
def test_chol_structured_matches_lapack(cuda_device, kind, order_mode):
    """K11s (csrc/chol_dag.cu) against LAPACK on dense, banded, cyclic-banded (the C5 shape), arrowhead, random
    sparse and block-diagonal systems, in natural and nested-dissection order; launched three times on the same
    plan (the flags are epochs, never reset)."""
    import torch
    from sfm_b200 import _lib, device
    if kind == "ring125" and order_mode == 1:
        pytest.skip("covered by the auto order")
    rng = np.random.default_rng(len(kind) * 7 + order_mode)
    n, S = _structured_case(kind, rng)
    g = rng.normal(size=n)
    lib = _lib.load()
    nbytes = int(lib.sfm_chol_solve_workspace_bytes(n))
    ws = device.empty((nbytes,), torch.uint8)
    d_x = device.empty((n,), torch.float64)
    d_fail = device.empty((1,), torch.int32)
    d_S, d_g = device.to_device(np.tril(S)), device.to_device(g)
    info = np.zeros(6)
    _lib.check(lib.sfm_chol_solve_structured(device.ptr(d_S), device.ptr(d_g), n, order_mode, 3, device.ptr(d_x),
                                             device.ptr(d_fail), device.ptr(ws), nbytes, info.ctypes.data,
                                             device.stream()), "sfm_chol_solve_structured")
    assert int(d_fail.cpu()[0]) == 0
    ref = np.linalg.solve(S, g)
    err = rel_err(d_x.cpu().numpy(), ref)
    assert err < 1e-9 * max(1.0, np.linalg.cond(S) * 1e-6), err
    if kind == "ring125" and order_mode != 1:
        assert info[1] <= 45, info                                 # elimination tree cut from 125 blocks
    # an indefinite matrix is reported
    S2 = np.tril(S).copy()
    S2[n // 2, n // 2] = -1.0
    d_S2 = device.to_device(S2)
    _lib.check(lib.sfm_chol_solve_structured(device.ptr(d_S2), device.ptr(d_g), n, order_mode, 1, device.ptr(d_x),
                                             device.ptr(d_fail), device.ptr(ws), nbytes, None, device.stream()),
               "sfm_chol_solve_structured")
    assert int(d_fail.cpu()[0]) == 1


@pytest.mark.parametrize("tag", ["a", "b"])
def test_ba_residuals_match_reference(cuda_device, tag):
    from sfm_b200 import bundle
    g = golden(f"ba_{tag}.npz")
    nc, npts = len(g["R0"]), len(g["pts0"])
    r0 = bundle.bundle_adjustment_residuals(g["x0"], nc, npts, g["cam_idx"], g["pt_idx"], g["obs"], g["K"])
    assert r0.shape == g["r0"].shape
    assert np.max(np.abs(r0 - g["r0"])) < 1e-9
    # caller's observation order is preserved when it is not point-major
    perm = np.random.default_rng(0).permutation(len(g["cam_idx"]))
    r1 = bundle.bundle_adjustment_residuals(g["x0"], nc, npts, g["cam_idx"][perm], g["pt_idx"][perm],
                                            g["obs"][perm], g["K"])
    assert np.max(np.abs(r1.reshape(-1, 2) - g["r0"].reshape(-1, 2)[perm])) < 1e-9


@pytest.mark.parametrize("tag", ["a", "b"])
def test_perform_bundle_adjustment_matches_reference(cuda_device, schur_path, tag):
    from sfm_b200 import bundle
    g = golden(f"ba_{tag}.npz")
    nc, npts, nobs = len(g["R0"]), len(g["pts0"]), len(g["cam_idx"])
    R_set, C_set = [r for r in g["R0"]], [c for c in g["C0"]]
    buf, info = io.StringIO(), {}
    with redirect_stdout(buf):
        R_opt, C_opt, world = bundle.perform_bundle_adjustment(g["world"], g["valid"], g["fx"], g["fy"], g["flags"],
                                                               R_set, C_set, g["K"], nc - 1, info=info)
    log = buf.getvalue()
    assert f"Bundle adjustment: {nc} cameras, {npts} points, {nobs} observations" in log
    assert "Bundle adjustment completed. Final cost:" in log
    ref_cost = 0.5 * float(g["rf"] @ g["rf"])
    assert abs(ref_cost - float(g["printed_cost"])) < 1e-5
    # cost recomputed by the ORACLE residual on the returned parameters (SURVEY 8d parity rule)
    from scipy.spatial.transform import Rotation
    xf = np.concatenate([np.concatenate([np.hstack([Rotation.from_matrix(R).as_rotvec(), -R @ C])
                                         for R, C in zip(R_opt, C_opt)]),
                         world[g["valid"].flatten() == 1].reshape(-1)])
    rf = bo.ba_residuals(xf, nc, npts, g["cam_idx"], g["pt_idx"], g["obs"], g["K"])
    cost = 0.5 * float(rf @ rf)
    assert abs(cost - ref_cost) / ref_cost < COST_RTOL, (cost, ref_cost)
    assert abs(cost - info["cost"]) / cost < 1e-9
    assert abs(_rmse(cost, nobs) - _rmse(ref_cost, nobs)) < RMSE_ATOL
    assert info["status"] in (1, 2, 3, 4)
    # rows that are not valid points are untouched, inputs are not mutated
    inval = g["valid"].flatten() != 1
    assert np.array_equal(world[inval], g["world"][inval])
    assert len(R_opt) == nc and len(C_opt) == nc and world.shape == g["world"].shape


@pytest.mark.parametrize("nc,npts,k,seed", [(8, 150, 5, 7), (12, 300, 6, 9), (20, 800, 7, 11)])
def test_ba_sparse_matches_oracle_lm(cuda_device, schur_path, nc, npts, k, seed):
    from sfm_b200 import bundle
    from oracle import scenes
    sc = scenes.ba_scene(n_cams=nc, n_pts=npts, obs_per_pt=k, seed=seed)
    cams0, pts0 = _x0(sc)
    x0 = np.concatenate([cams0.reshape(-1), pts0.reshape(-1)])
    method = "lm" if npts <= 300 else "trf"
    ref = bo.ba_solve(x0, nc, npts, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"], method=method)
    cams, pts, info = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    x = np.concatenate([cams.reshape(-1), pts.reshape(-1)])
    r = bo.ba_residuals(x, nc, npts, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    cost = 0.5 * float(r @ r)
    assert abs(cost - ref.cost) / ref.cost < COST_RTOL, (cost, ref.cost, info)
    assert abs(_rmse(cost, len(sc["obs"])) - _rmse(ref.cost, len(sc["obs"]))) < RMSE_ATOL
    assert info["status"] in (1, 2, 3, 4) and info["n_linearisations"] <= 30


def test_ba_long_tracks_and_duplicate_cameras(cuda_device, schur_path):
    """Points seen by more than 32 cameras take the global-memory Schur path; an observation list
    that names the same camera twice on one point must still match the oracle."""
    from sfm_b200 import bundle
    from oracle import scenes
    sc = scenes.ba_scene(n_cams=40, n_pts=120, obs_per_pt=36, seed=21)
    cam_idx, pt_idx, obs = sc["cam_idx"].copy(), sc["pt_idx"].copy(), sc["obs"].copy()
    # duplicate a few observations (same camera, slightly different pixel) on short-track points
    sc2 = scenes.ba_scene(n_cams=40, n_pts=120, obs_per_pt=5, seed=22)
    dup = np.arange(0, len(sc2["cam_idx"]), 17)
    c2 = np.concatenate([sc2["cam_idx"], sc2["cam_idx"][dup]])
    p2 = np.concatenate([sc2["pt_idx"], sc2["pt_idx"][dup]])
    o2 = np.concatenate([sc2["obs"], sc2["obs"][dup] + 0.3])
    order = np.argsort(p2, kind="stable")
    for (sc_, ci, pi, ob) in ((sc, cam_idx, pt_idx, obs), (sc2, c2[order], p2[order], o2[order])):
        cams0, pts0 = _x0(sc_)
        x0 = np.concatenate([cams0.reshape(-1), pts0.reshape(-1)])
        ref = bo.ba_solve(x0, 40, 120, ci, pi, ob, sc_["K"], method="trf")
        cams, pts, info = bundle.ba_solve_sparse(cams0, pts0, ci, pi, ob, sc_["K"])
        x = np.concatenate([cams.reshape(-1), pts.reshape(-1)])
        r = bo.ba_residuals(x, 40, 120, ci, pi, ob, sc_["K"])
        cost = 0.5 * float(r @ r)
        assert abs(cost - ref.cost) / ref.cost < COST_RTOL, (cost, ref.cost, info)


def test_ba_error_conventions(cuda_device):
    from sfm_b200 import _lib, bundle
    from oracle import scenes
    g = {k: v for k, v in golden("ba_a.npz").items()}       # materialise once: identity checks below
    nc = len(g["R0"])
    R_set, C_set = [r for r in g["R0"]], [c for c in g["C0"]]
    # no valid points -> the inputs come back, nothing printed (BundleAdjustment.py:152-153)
    buf = io.StringIO()
    with redirect_stdout(buf):
        out = bundle.perform_bundle_adjustment(g["world"], np.zeros_like(g["valid"]), g["fx"], g["fy"], g["flags"],
                                               R_set, C_set, g["K"], nc - 1)
    assert out[0] is R_set and out[1] is C_set and out[2] is g["world"] and buf.getvalue() == ""
    # no observations (:171-172)
    with redirect_stdout(buf):
        out = bundle.perform_bundle_adjustment(g["world"], g["valid"], g["fx"], g["fy"], np.zeros_like(g["flags"]),
                                               R_set, C_set, g["K"], nc - 1)
    assert out[0] is R_set and out[2] is g["world"]
    # fewer residuals than variables: scipy's 'lm' raises, the reference prints and returns inputs
    flags = g["flags"].copy()
    flags[:, 1:] = 0
    buf = io.StringIO()
    with redirect_stdout(buf):
        out = bundle.perform_bundle_adjustment(g["world"], g["valid"], g["fx"], g["fy"], flags, R_set, C_set,
                                               g["K"], nc - 1)
    assert out[0] is R_set and out[2] is g["world"] and "Bundle adjustment failed:" in buf.getvalue()
    # non-finite initial point -> same convention
    wbad = g["world"].copy()
    wbad[np.where(g["valid"].flatten() == 1)[0][0]] = np.nan
    buf = io.StringIO()
    with redirect_stdout(buf):
        out = bundle.perform_bundle_adjustment(wbad, g["valid"], g["fx"], g["fy"], g["flags"], R_set, C_set,
                                               g["K"], nc - 1)
    assert out[2] is wbad and "Bundle adjustment failed:" in buf.getvalue()
    # out-of-range indices are rejected by the library
    sc = scenes.ba_scene(n_cams=4, n_pts=30, obs_per_pt=3, seed=3)
    cams0, pts0 = _x0(sc)
    bad = sc["cam_idx"].copy()
    bad[5] = 99
    with pytest.raises(_lib.SfmError):
        bundle.ba_solve_sparse(cams0, pts0, bad, sc["pt_idx"], sc["obs"], sc["K"])


def test_ba_c3_scale_properties(cuda_device):
    """BASELINE config 3 (50 cameras, 20k points, 200k observations): the reference cannot run it
    (SURVEY 6), so check size-independent properties: monotone decrease to a stationary point whose
    RMSE is at the noise level, determinism, and agreement of K8 with the oracle residual."""
    from sfm_b200 import bundle, device
    from oracle import scenes
    sc = scenes.ba_scene()
    cams0, pts0 = _x0(sc)
    nobs = len(sc["obs"])
    cams, pts, info = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    assert info["status"] in (1, 2, 3, 4), info
    assert info["cost"] < info["cost0"]
    x = np.concatenate([cams.reshape(-1), pts.reshape(-1)])
    r = bo.ba_residuals(x, 50, 20000, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    cost = 0.5 * float(r @ r)
    assert abs(cost - info["cost"]) / cost < 1e-10
    rmse = _rmse(cost, nobs)
    # noise 0.5 px per coordinate; (m - n) / m degrees of freedom remain after the fit
    expect = 0.5 * np.sqrt(2.0) * np.sqrt((2 * nobs - 60300) / (2 * nobs))
    assert abs(rmse - expect) < 0.01, (rmse, expect)
    # tighter tolerances barely move the cost: the default stop is at the minimum to << 1e-6
    cams_t, pts_t, info_t = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"],
                                                   tol=1e-13, max_nfev=60)
    assert abs(info_t["cost"] - info["cost"]) / info["cost"] < COST_RTOL
    # bit-reproducible
    cams2, pts2, info2 = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    assert info2["cost"] == info["cost"] or abs(info2["cost"] - info["cost"]) / info["cost"] < 1e-12


def test_schur_pair_list_equals_atomic_path(cuda_device, monkeypatch):
    """K10g against K10 on one damped trial system: same parameters after a fixed number of trial steps
    (the two kernels sum the same products in different orders, so agreement is to rounding, far inside
    the 1e-6 cost tolerance), on a scene with long tracks, duplicated cameras and a camera seen once."""
    from sfm_b200 import bundle
    from oracle import scenes
    sc = scenes.ba_scene(n_cams=30, n_pts=2500, obs_per_pt=9, seed=5)
    cam_idx, pt_idx, obs = sc["cam_idx"].copy(), sc["pt_idx"].copy(), sc["obs"].copy()
    dup = np.arange(3, len(cam_idx), 101)
    c2 = np.concatenate([cam_idx, cam_idx[dup]])
    p2 = np.concatenate([pt_idx, pt_idx[dup]])
    o2 = np.concatenate([obs, obs[dup] + 0.2])
    order = np.argsort(p2, kind="stable")
    c2, p2, o2 = c2[order], p2[order], o2[order]
    cams0, pts0 = _x0(sc)
    out = {}
    for name, env in (("atomic", {"SFM_SCHUR_PAIRS": "0"}), ("pairs", {"SFM_SCHUR_PAIRS_MIN": "0"})):
        monkeypatch.delenv("SFM_SCHUR_PAIRS", raising=False)
        monkeypatch.delenv("SFM_SCHUR_PAIRS_MIN", raising=False)
        for k_, v_ in env.items():
            monkeypatch.setenv(k_, v_)
        cams, pts, info = bundle.ba_solve_sparse(cams0, pts0, c2, p2, o2, sc["K"])
        out[name] = (cams, pts, info)
    (ca, pa, ia), (cb, pb, ib) = out["atomic"], out["pairs"]
    assert ia["n_trials"] == ib["n_trials"] and ia["status"] == ib["status"]
    assert abs(ia["cost"] - ib["cost"]) / ia["cost"] < 1e-10
    assert rel_err(cb, ca) < 1e-7 and rel_err(pb, pa) < 1e-7


def test_release_cache_frees_pair_lists_and_solves_again(cuda_device, monkeypatch):
    """sfm_ba_release_cache drops the cached pair lists (reported in bytes); the next solve rebuilds them and
    returns the same result."""
    from sfm_b200 import bundle
    from oracle import scenes
    monkeypatch.setenv("SFM_SCHUR_PAIRS_MIN", "0")
    sc = scenes.ba_scene(n_cams=10, n_pts=400, obs_per_pt=5, seed=3)
    cams0, pts0 = _x0(sc)
    bundle.release_cache()
    c1, p1, i1 = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    freed = bundle.release_cache()
    assert freed >= 12 * 400 * 15                      # 15 pairs per point, 12 B each, before padding
    assert bundle.release_cache() == 0
    c2, p2, i2 = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    assert i1["n_trials"] == i2["n_trials"] and abs(i1["cost"] - i2["cost"]) <= 1e-12 * i1["cost"]


@pytest.mark.parametrize("name", ["ring100", "C3"])
def test_ba_benchmark_sizes_against_the_sparse_oracle(cuda_device, name):
    """Parity AT THE BENCHMARK SIZE: BASELINE configs[2] itself (C3: 50 cameras, 20k points, 200k observations,
    60,300 parameters) and a 100-camera ring scene of the C5 kind (banded reduced camera system, 160k observations)
    against the scalable secondary oracle SURVEY 8(c) names -- scipy least_squares(method='trf', x_scale='jac') with
    the block-sparse Jacobian of the reference's residual (oracle/ba_oracle.py; the reference's own 'lm' needs a dense
    193 GB Jacobian here). north_star's criteria: final cost within 1e-6 relative, reprojection RMSE within 1e-3 px;
    the cost is recomputed by the oracle's restatement of bundle_adjustment_residuals on the returned parameters."""
    import bench_config
    from oracle import scenes
    from sfm_b200 import bundle
    cfg = bench_config.C3 if name == "C3" else dict(n_cams=100, n_pts=20_000, obs_per_pt=8, seed=5, layout="ring", window=20)
    sc = scenes.ba_scene(**cfg)
    cams0, pts0 = _x0(sc)
    nc, npts, nobs = cfg["n_cams"], cfg["n_pts"], len(sc["obs"])
    cams, pts, info = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    x = np.concatenate([cams.reshape(-1), pts.reshape(-1)])
    r = bo.ba_residuals(x, nc, npts, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    cost_gpu = 0.5 * float(r @ r)
    assert abs(cost_gpu - info["cost"]) <= 1e-9 * cost_gpu                 # the library's own figure is the oracle's
    x0 = np.concatenate([cams0.reshape(-1), pts0.reshape(-1)])
    ref = bo.ba_solve(x0, nc, npts, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"], method="trf")
    assert ref.status > 0
    assert abs(cost_gpu - ref.cost) / ref.cost < COST_RTOL, (cost_gpu, ref.cost)
    assert abs(_rmse(cost_gpu, nobs) - _rmse(ref.cost, nobs)) < RMSE_ATOL
    assert info["status"] in (1, 2, 3, 4) and cost_gpu <= ref.cost * (1 + COST_RTOL)


def test_host_api_reuses_the_plan_of_an_unchanged_index_structure(cuda_device):
    """ba_solve_sparse caches the problem (workspace, camera-major index, pair list, K11s plan) by the CONTENT of
    (cam_idx, pt_idx): a second call with the same tracks and new measurements must give exactly what a fresh problem
    gives, the cache stays bounded, and close() / clear_problem_cache() release the library's side."""
    from oracle import scenes
    from sfm_b200 import bundle
    bundle.clear_problem_cache()
    sc = scenes.ba_scene(n_cams=12, n_pts=3000, obs_per_pt=6, seed=4)          # > 16384 observations: K10g + K11s plans
    cams0, pts0 = _x0(sc)
    a = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    assert len(bundle._PROBLEMS) == 1
    obs2 = sc["obs"] + np.random.default_rng(0).normal(0, 0.3, sc["obs"].shape)
    b = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"].copy(), sc["pt_idx"].copy(), obs2, sc["K"])   # hit
    assert len(bundle._PROBLEMS) == 1
    fresh = bundle.BAProblem(12, 3000, sc["cam_idx"], sc["pt_idx"], obs2, sc["K"])
    from sfm_b200 import device
    d_c, d_p = device.to_device(cams0), device.to_device(pts0)
    info = fresh.solve_device(d_c, d_p)
    assert abs(info["cost"] - b[2]["cost"]) <= 1e-9 * info["cost"] and b[2]["cost"] != a[2]["cost"]
    assert np.allclose(d_c.cpu().numpy(), b[0], atol=1e-7)
    fresh.close()
    # an unsorted caller order is restored on a hit too
    perm = np.random.default_rng(1).permutation(len(sc["obs"]))
    c = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"][perm], sc["pt_idx"][perm], obs2[perm], sc["K"])
    c2 = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"][perm], sc["pt_idx"][perm], obs2[perm], sc["K"])
    assert abs(c[2]["cost"] - b[2]["cost"]) <= 1e-9 * b[2]["cost"] and abs(c2[2]["cost"] - c[2]["cost"]) <= 1e-9 * c[2]["cost"]
    # the identity fast path must not trust objects whose CONTENT changed in place
    ci, pi = sc["cam_idx"].copy(), sc["pt_idx"].copy()
    d1 = bundle.ba_solve_sparse(cams0, pts0, ci, pi, obs2, sc["K"])
    d2 = bundle.ba_solve_sparse(cams0, pts0, ci, pi, obs2, sc["K"])             # identity hit
    assert abs(d2[2]["cost"] - d1[2]["cost"]) <= 1e-9 * d1[2]["cost"]
    keep = pi < 2999                                             # drop the last point's observations, same objects reused
    ci2, pi2, ob2 = ci[keep], pi[keep], obs2[keep]
    ci[:len(ci2)] = ci2                                          # (write into the old objects too: must be noticed)
    d3 = bundle.ba_solve_sparse(cams0, pts0[:2999], ci2, pi2, ob2, sc["K"])
    f3 = bundle.BAProblem(12, 2999, ci2, pi2, ob2, sc["K"])
    d_c3, d_p3 = device.to_device(cams0), device.to_device(pts0[:2999])
    i3 = f3.solve_device(d_c3, d_p3)
    assert abs(i3["cost"] - d3[2]["cost"]) <= 1e-9 * i3["cost"]
    f3.close()
    for s in range(6):                                           # different structures: the cache evicts
        sc_s = scenes.ba_scene(n_cams=5, n_pts=60 + s, obs_per_pt=4, seed=s)
        cs, ps = _x0(sc_s)
        bundle.ba_solve_sparse(cs, ps, sc_s["cam_idx"], sc_s["pt_idx"], sc_s["obs"], sc_s["K"])
    assert len(bundle._PROBLEMS) <= bundle.PROBLEM_CACHE_ENTRIES
    bundle.clear_problem_cache()
    assert len(bundle._PROBLEMS) == 0


@pytest.mark.parametrize("tag", ["a", "b"])
def test_budget_limited_runs_quantify_the_controller_difference(cuda_device, tag, capsys):
    """The reference's LM is MINPACK lmder (trust region through lmpar); the Schur-complement solver here updates the
    damping directly (Nielsen), because J has a 7-dimensional gauge null space that lmpar's secular equation walks
    through on rounding noise. The two agree WHERE THEY CONVERGE -- which is what every other
    test pins -- and differ along the way. This test pins what a caller can rely on when the evaluation budget
    (max_nfev, BundleAdjustment.py:211: 100) cuts a run short, and prints the measured gap per budget:
      * every budget: cost decreases from the start and never undershoots the common minimum;
      * with the reference's own budget of 100 both are converged and equal (1e-6);
      * per budget the GPU's cost is within a factor of the reference's at the SAME budget (measured ~1, asserted 10)."""
    from sfm_b200 import bundle
    g = golden(f"ba_{tag}.npz")
    prob = bo.ba_build_problem(g["world"], g["valid"], g["fx"], g["fy"], g["flags"], list(g["R0"]), list(g["C0"]))
    nc, npts = prob["n_cameras"], prob["n_points"]
    args = (nc, npts, prob["cam_idx"], prob["pt_idx"], prob["pts2d"], g["K"])
    x0 = prob["x0"]
    cams0, pts0 = x0[:6 * nc].reshape(nc, 6), x0[6 * nc:].reshape(npts, 3)
    cost0 = 0.5 * float(np.sum(bo.ba_residuals(x0, *args) ** 2))
    ref_conv = bo.ba_solve(x0, *args, method="lm").cost
    rows = []
    for budget in (1, 2, 3, 5, 8, 100):
        ref = bo.ba_solve(x0, *args, method="lm", max_nfev=budget)
        _, _, info = bundle.ba_solve_sparse(cams0, pts0, prob["cam_idx"], prob["pt_idx"], prob["pts2d"], g["K"],
                                            max_nfev=budget)
        rows.append((budget, ref.cost, ref.status, info["cost"], info["status"], info["n_trials"]))
        assert info["n_trials"] <= budget
        assert ref_conv * (1 - 1e-6) <= info["cost"] <= cost0
        assert info["cost"] <= 10.0 * ref.cost
    assert abs(rows[-1][3] - rows[-1][1]) <= COST_RTOL * rows[-1][1] and rows[-1][4] in (1, 2, 3, 4)
    costs = [r[3] for r in rows]
    assert all(b <= a * (1 + 1e-12) for a, b in zip(costs, costs[1:]))          # a larger budget never ends higher
    with capsys.disabled():
        print(f"\nBA {tag}: cost0 {cost0:.4f}, converged {ref_conv:.6f}; budget: reference cost (status) | GPU cost (status, trials)")
        for b, rc, rs, gc, gs, gt in rows:
            print(f"  max_nfev {b:3d}: {rc:14.6f} ({rs}) | {gc:14.6f} ({gs}, {gt})")
