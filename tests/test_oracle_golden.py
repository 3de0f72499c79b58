"""The oracle (oracle/*.py) against the frozen outputs of the unmodified reference
(tests/golden/*.npz, made by oracle/make_golden.py). CPU only."""
import io
import random
from contextlib import redirect_stdout

import numpy as np
import pytest

from conftest import golden, rel_err
from oracle import ba_oracle, sfm_oracle as so


def test_f8pt_matches_reference():
    g = golden("f8pt.npz")
    for s, Fref in zip(g["idx"], g["F"]):
        F = so.estimate_fundamental_matrix(g["x1"][s], g["x2"][s])
        assert rel_err(F.ravel(), Fref) < 1e-12
    good = g["good"]
    assert rel_err(so.estimate_fundamental_matrix(g["x1"][good], g["x2"][good]), g["F_all_good"]) < 1e-12
    assert rel_err(so.estimate_fundamental_matrix(g["x1"], g["x2"]), g["F_all"]) < 1e-12


@pytest.mark.parametrize("tag", ["a", "b", "c", "d"])
def test_fransac_matches_reference(tag):
    g = golden(f"fransac_{tag}.npz")
    random.seed(int(g["seed"]))
    inl, outl, F = so.get_inliers_ransac_full(g["x1"], g["x2"], g["index"], float(g["thr"]), int(g["n_max"]))
    assert np.array_equal(inl, g["inlier_index"])
    assert np.array_equal(outl, g["outlier_indices"])
    assert rel_err(F, g["F"]) < 1e-12
    # the recorded sample stream is the one the reference consumed
    random.seed(int(g["seed"]))
    assert np.array_equal(so.draw_samples(len(g["x1"]), 8, int(g["n_max"])), g["samples"])
    F2, idx = so.get_inliers_ransac(g["x1"], g["x2"], g["index"], float(g["thr"]), samples=g["samples"])
    assert np.array_equal(idx, g["alias_inlier_idx"]) and rel_err(F2, g["F"]) < 1e-12


@pytest.mark.parametrize("pair", ["12", "23", "14"])
def test_p3data_fransac_matches_reference(pair):
    g = golden("p3data_fransac.npz")
    F, idx = so.get_inliers_ransac(g[f"p1_{pair}"], g[f"p2_{pair}"], g[f"hidx_{pair}"], 0.06,
                                   samples=g[f"samples_{pair}"])
    assert np.array_equal(idx, g[f"fidx_{pair}"])
    assert rel_err(F, g[f"F_{pair}"]) < 1e-12


def test_nltri_matches_reference():
    g = golden("nltri.npz")
    P1 = so.projection_matrix(g["K"], g["C1"], g["R1"])
    P2 = so.projection_matrix(g["K"], g["C2"], g["R2"])
    for i in range(0, 300, 7):
        assert np.allclose(so.tri_loss(g["x0"][i], g["x1"][i], g["x2"][i], P1, P2), g["loss0"][i],
                           rtol=0, atol=1e-12)
    sl = slice(0, 60)
    X = so.nonlinear_triangulation(g["K"], g["C1"], g["R1"], g["C2"], g["R2"], g["x1"][sl], g["x2"][sl], g["x0"][sl])
    assert np.max(np.abs(X - g["X"][sl])) < 1e-12
    Xb = so.nonlinear_triangulation(g["K"], g["C1"], g["R1"], g["C2"], g["R2"], g["x1"][sl], g["x2"][sl], g["x0b"][sl])
    assert np.max(np.abs(Xb - g["Xb"][sl])) < 1e-12


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_pnp_matches_reference(tag):
    g = golden("pnp.npz")
    X, x, K = g[f"X_{tag}"], g[f"x_{tag}"], g["K"]
    p0 = so.pose_to_params(g[f"C0_{tag}"], g[f"R0_{tag}"])
    assert np.allclose(so.pnp_loss(p0, X, x, K), g[f"loss0_{tag}"], rtol=0, atol=1e-11)
    C, R = so.nonlinear_pnp(X, x, K, g[f"C0_{tag}"], g[f"R0_{tag}"])
    assert rel_err(C, g[f"C_{tag}"]) < 1e-12 and rel_err(R, g[f"R_{tag}"]) < 1e-12
    Cl, Rl = so.linear_pnp(X, x, K)
    assert rel_err(Cl, g[f"Clin_{tag}"]) < 1e-9 and rel_err(Rl, g[f"Rlin_{tag}"]) < 1e-9
    with redirect_stdout(io.StringIO()):
        Cr, Rr = so.pnp_ransac(X, x, K, 200, samples=g[f"samples_{tag}"])
    assert rel_err(Cr, g[f"Cransac_{tag}"]) < 1e-9 and rel_err(Rr, g[f"Rransac_{tag}"]) < 1e-9


@pytest.mark.parametrize("tag", ["a", "b"])
def test_ba_matches_reference(tag):
    g = golden(f"ba_{tag}.npz")
    nc, npts = len(g["R0"]), len(g["pts0"])
    r0 = ba_oracle.ba_residuals(g["x0"], nc, npts, g["cam_idx"], g["pt_idx"], g["obs"], g["K"])
    assert np.max(np.abs(r0 - g["r0"])) < 1e-11
    prob = ba_oracle.ba_build_problem(g["world"], g["valid"], g["fx"], g["fy"], g["flags"],
                                      list(g["R0"]), list(g["C0"]))
    assert np.array_equal(prob["cam_idx"], g["cam_idx"]) and np.array_equal(prob["pt_idx"], g["pt_idx"])
    assert np.array_equal(prob["pts2d"], g["obs"]) and np.allclose(prob["x0"], g["x0"], rtol=0, atol=1e-15)
    info = {}
    with redirect_stdout(io.StringIO()):
        R_opt, C_opt, world = ba_oracle.perform_bundle_adjustment(
            g["world"], g["valid"], g["fx"], g["fy"], g["flags"], list(g["R0"]), list(g["C0"]), g["K"],
            nc - 1, info=info)
    cost_ref = 0.5 * float(np.sum(g["rf"] ** 2))
    assert abs(info["cost"] - cost_ref) / cost_ref < 1e-9
    assert abs(float(g["printed_cost"]) - cost_ref) < 1e-5
    # rows that are not valid points come back untouched
    inval = g["valid"].flatten() != 1
    assert np.array_equal(world[inval], g["world"][inval])


def test_linear_triangulation_matches_reference():
    g = golden("nltri.npz")
    X = so.linear_triangulation(g["K"], g["C1"], g["R1"], g["C2"], g["R2"], g["x1"], g["x2"])
    assert np.max(np.abs(X - g["x0"])) < 1e-9
