"""The rest of the hot path for __graft_entry__.smoke(): one small invocation each of triangulation,
two-view pose, homography RANSAC, PnP RANSAC + NonlinearPnP and bundle adjustment on cuda:0, checked against the oracle.
(Test infrastructure hook: it imports oracle/, so nothing in the product path may import it.)"""
from __future__ import annotations

import io
import random
from contextlib import redirect_stdout

import numpy as np


def run() -> str:
    from oracle import ba_oracle as bo
    from oracle import sfm_oracle as so
    from sfm_b200 import _lib, bundle, homography, refine, two_view
    from oracle import scenes

    before = _lib.launch_count()
    # N1 + K5: linear then non-linear triangulation
    sc = scenes.two_view(n=256, outlier_frac=0.0, seed=11)
    x0 = refine.LinearTriangulation(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"], sc["x1"], sc["x2"])
    x0_ref = so.linear_triangulation(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"], sc["x1"], sc["x2"])
    assert np.max(np.linalg.norm(x0 - x0_ref, axis=1) / np.linalg.norm(x0_ref, axis=1)) < 1e-8
    X = refine.NonLinearTriangulation(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"], sc["x1"], sc["x2"], x0)
    c_gpu = so.tri_cost(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"], sc["x1"], sc["x2"], X)
    c_ref = so.tri_cost(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"], sc["x1"], sc["x2"],
                        so.nonlinear_triangulation(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"], sc["x1"], sc["x2"], x0_ref))
    assert np.all(np.abs(c_gpu - c_ref) <= 1e-6 * np.maximum(c_ref, 1e-12) + 1e-12)
    # N4: two-view pose from the F of all correspondences
    F = so.estimate_fundamental_matrix(sc["x1"], sc["x2"])
    tv = two_view.two_view_pose(F, sc["K"], sc["x1"], sc["x2"])
    Cs, Rs = so.extract_camera_pose(so.essential_from_fundamental(F, sc["K"]))
    Xs = [so.linear_triangulation(sc["K"], np.zeros(3), np.eye(3), c, r, sc["x1"], sc["x2"]) for c, r in zip(Cs, Rs)]
    Cb, Rb, _, _, nb = so.disambiguate_camera_pose(Cs, Rs, Xs)
    assert tv["best_count"] == nb and np.allclose(tv["C"], Cb, atol=1e-8) and np.allclose(tv["R"], Rb, atol=1e-8)
    # N2: homography RANSAC
    rng = np.random.default_rng(2)
    Ht = np.array([[0.9, 0.05, 30.0], [-0.04, 1.1, -12.0], [1e-5, -2e-5, 1.0]])
    p1 = rng.uniform(0, 1000, (300, 2))
    t = (Ht @ np.hstack([p1, np.ones((300, 1))]).T).T
    p2 = t[:, :2] / t[:, 2:3] + rng.normal(0, 0.5, (300, 2))
    p2[:60] = rng.uniform(0, 1000, (60, 2))
    random.seed(3)
    hs = so.draw_samples(300, 4, 128)
    hg = homography.h_ransac_host(p1, p2, hs, 5.0)
    hr = so.h_ransac_detailed(p1, p2, 5.0, hs)
    assert hg["best_idx"] == hr["best_idx"] and hg["best_count"] == hr["best_count"]
    # K6a + K6 + K7: PnP RANSAC (LinearPnP hypotheses on the device) from the reference's RNG stream, then NonlinearPnP
    from sfm_b200 import pnp_ransac
    ps = scenes.pnp_scene(n=300, seed=17)
    random.seed(5)
    Cp, Rp = pnp_ransac.PnPRANSAC(ps["X"], ps["x"], scenes.K_REF, 200, 200)
    random.seed(5)
    Co, Ro = so.pnp_ransac(ps["X"], ps["x"], scenes.K_REF, 200, 200)
    assert np.linalg.norm(Rp - Ro) < 1e-6 * np.linalg.norm(Ro) and np.linalg.norm(Cp - Co) <= 1e-6 * max(1.0, np.linalg.norm(Co))
    Cn, Rn = refine.nonlinear_PnP(scenes.K_REF, Cp, Rp, ps["x"], ps["X"])
    Cr, Rr = so.nonlinear_pnp(ps["X"], ps["x"], scenes.K_REF, Co, Ro)
    pnp_cost = lambda C, R: 0.5 * float(np.sum(so.pnp_loss(so.pose_to_params(C, R), ps["X"], ps["x"], scenes.K_REF) ** 2))
    assert abs(pnp_cost(Cn, Rn) - pnp_cost(Cr, Rr)) <= 1e-6 * pnp_cost(Cr, Rr)
    # K8..K12: a small bundle adjustment against scipy's 'lm' on the same residual
    from scipy.spatial.transform import Rotation
    bs = scenes.ba_scene(n_cams=5, n_pts=60, obs_per_pt=4, seed=13)
    cams0 = np.hstack([Rotation.from_matrix(bs["R0"]).as_rotvec(), -np.einsum("nij,nj->ni", bs["R0"], bs["C0"])])
    with redirect_stdout(io.StringIO()):
        cams, pts, info = bundle.ba_solve_sparse(cams0, bs["pts0"], bs["cam_idx"], bs["pt_idx"], bs["obs"], bs["K"])
    x0 = np.concatenate([cams0.reshape(-1), bs["pts0"].reshape(-1)])
    ref = bo.ba_solve(x0, 5, 60, bs["cam_idx"], bs["pt_idx"], bs["obs"], bs["K"], method="lm")
    x = np.concatenate([cams.reshape(-1), pts.reshape(-1)])
    r = bo.ba_residuals(x, 5, 60, bs["cam_idx"], bs["pt_idx"], bs["obs"], bs["K"])
    cost = 0.5 * float(r @ r)
    rel = abs(cost - ref.cost) / ref.cost
    assert rel < 1e-6, (cost, ref.cost)
    launches = _lib.launch_count() - before
    return (f"; triangulation, two-view pose ({tv['best_count']} in front), homography RANSAC "
            f"({hg['best_count']} inliers), PnP RANSAC + NonlinearPnP (cost {pnp_cost(Cn, Rn):.5g}), BA cost {cost:.6g} (rel. diff to scipy lm {rel:.1e}), "
            f"{launches} more launches")
