"""End to end: the reference wrapper's control flow (tests/wrapper_flow.py) driven once by the oracle
and once by the GPU drop-in modules bound exactly as Wrapper_dev.py:5-14 binds them (bare module names
from structure-from-motion-_b200/dropin on sys.path), on the same seeded feature arrays and the same
`random` stream. Every stage must agree: F, the selected two-view pose, triangulated points, the
registered cameras, the bundle-adjusted structure."""
import importlib
import io
import os
import random
import sys
import types
from contextlib import redirect_stdout

import numpy as np
import pytest

import wrapper_flow
from conftest import rel_err
from oracle import ba_oracle as bo
from oracle import sfm_oracle as so

pytestmark = pytest.mark.gpu

DROPIN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "structure-from-motion-_b200", "dropin")


def _ba_cost(world, have, fx, fy, flags, R_set, C_set, K):
    prob = bo.ba_build_problem(world, have, fx, fy, flags, R_set, C_set)
    r = bo.ba_residuals(prob["x0"], prob["n_cameras"], prob["n_points"], prob["cam_idx"], prob["pt_idx"], prob["pts2d"], K)
    return 0.5 * float(r @ r)


def _oracle_ns():
    ns = types.SimpleNamespace()
    ns.get_homography_inliers = so.get_homography_inliers
    ns.get_inliers_ransac = lambda p1, p2, idx, threshold, n_max: so.get_inliers_ransac(p1, p2, idx, threshold, n_max)
    ns.EssentialMatrixFromFundamentalMatrix = so.essential_from_fundamental
    ns.ExtractCameraPose = so.extract_camera_pose
    ns.linear_triangulation = so.linear_triangulation
    ns.DisambiguateCameraPose = lambda C, R, X: so.disambiguate_camera_pose(C, R, X)[:3]
    ns.nonlinear_triangulation = so.nonlinear_triangulation
    ns.PnPRANSAC = lambda X, x, K, threshold, n_max: so.pnp_ransac(X, x, K, threshold, n_max)
    ns.nonlinear_PnP = lambda K, C, R, x, X: so.nonlinear_pnp(X, x, K, C, R)
    ns.perform_bundle_adjustment = bo.perform_bundle_adjustment
    ns.ba_cost = _ba_cost
    return ns


def _dropin_ns():
    """Bind the way the wrapper does: bare names, dropin/ first on sys.path."""
    sys.path.insert(0, DROPIN)
    try:
        mods = {}
        for name in ("GetHomographyInliers", "GetInliersRANSAC", "EssentialMatrixFromFundamentalMatrix",
                     "ExtractCameraPose", "DisambiguateCameraPose", "LinearTriangulation", "NonLinearTriangulation",
                     "NonlinearPnP", "PnPRANSAC", "BundleAdjustment"):
            sys.modules.pop(name, None)
            mods[name] = importlib.import_module(name)
            assert os.path.dirname(os.path.abspath(mods[name].__file__)) == DROPIN
    finally:
        sys.path.remove(DROPIN)
    ns = types.SimpleNamespace()
    ns.get_homography_inliers = mods["GetHomographyInliers"].get_homography_inliers
    ns.get_inliers_ransac = mods["GetInliersRANSAC"].get_inliers_ransac
    ns.EssentialMatrixFromFundamentalMatrix = mods["EssentialMatrixFromFundamentalMatrix"].EssentialMatrixFromFundamentalMatrix
    ns.ExtractCameraPose = mods["ExtractCameraPose"].ExtractCameraPose
    ns.DisambiguateCameraPose = mods["DisambiguateCameraPose"].DisambiguateCameraPose
    ns.linear_triangulation = mods["LinearTriangulation"].linear_triangulation
    ns.nonlinear_triangulation = mods["NonLinearTriangulation"].nonlinear_triangulation
    ns.PnPRANSAC = mods["PnPRANSAC"].PnPRANSAC
    ns.nonlinear_PnP = mods["NonlinearPnP"].nonlinear_PnP
    ns.perform_bundle_adjustment = mods["BundleAdjustment"].perform_bundle_adjustment
    ns.ba_cost = _ba_cost
    return ns


def _lockstep(ns, K, log):
    """Every bundle adjustment of the drop-in flow is ALSO run by the oracle on exactly the inputs the GPU call
    received, and the two results are compared there. (Comparing the costs of two independent flows instead made
    this test depend on luck: the scene keeps a few gross outliers in the tracks, so its bundle adjustments have
    flat directions and more than one minimum, the GPU's atomics make the parameters of one converged call differ
    by ~1e-5 along those directions from run to run, and once in several runs the NEXT call then started in the
    basin of the other minimum -- measured: tools/pipeline_stress.py, costs 399.66 / 324.85.)"""
    gpu_ba = ns.perform_bundle_adjustment

    def both(world, have, fx, fy, flags, R_set, C_set, K_, cam_index):
        args = (np.array(world), np.array(have), fx, fy, np.array(flags), [np.array(r) for r in R_set],
                [np.array(c) for c in C_set], K_, cam_index)
        Rg, Cg, Wg = gpu_ba(world, have, fx, fy, flags, R_set, C_set, K_, cam_index)
        Rr, Cr, Wr = bo.perform_bundle_adjustment(*args)
        sel = np.asarray(have).reshape(-1) == 1
        reproj = 0.0
        for i in range(len(Rg)):
            pg = so.projection_matrix(K_, Cg[i], Rg[i]) @ np.hstack([np.asarray(Wg)[sel], np.ones((sel.sum(), 1))]).T
            pr = so.projection_matrix(K_, Cr[i], Rr[i]) @ np.hstack([np.asarray(Wr)[sel], np.ones((sel.sum(), 1))]).T
            reproj = max(reproj, float(np.max(np.abs(pg[:2] / pg[2] - pr[:2] / pr[2]))))
        prob = bo.ba_build_problem(world, have, fx, fy, flags, R_set, C_set)
        log.append(dict(gpu=_ba_cost(Wg, have, fx, fy, flags, Rg, Cg, K_), ref=_ba_cost(Wr, have, fx, fy, flags, Rr, Cr, K_),
                        start=_ba_cost(world, have, fx, fy, flags, R_set, C_set, K_), reproj=reproj,
                        n_obs=len(prob["cam_idx"])))
        return Rg, Cg, Wg
    ns.perform_bundle_adjustment = both
    return ns


def test_wrapper_flow_dropins_match_oracle(cuda_device):
    fx, fy, flag, K = wrapper_flow.synthetic_features()
    ba_log = []
    with redirect_stdout(io.StringIO()), np.errstate(all="ignore"):
        random.seed(11)
        ref = wrapper_flow.run(_oracle_ns(), fx, fy, flag, K)
        state_ref = random.getstate()
        random.seed(11)
        got = wrapper_flow.run(_lockstep(_dropin_ns(), K, ba_log), fx, fy, flag, K)
        state_got = random.getstate()
    assert state_got == state_ref                               # the same draws, in the same order
    assert np.array_equal(got["filtered"], ref["filtered"])     # identical RANSAC inlier sets for all pairs
    assert rel_err(got["F_12"], ref["F_12"]) < 1e-8
    assert got["n12"] == ref["n12"] and got["n12"] >= 20
    assert rel_err(got["C2"], ref["C2"]) < 1e-6 and rel_err(got["R2"], ref["R2"]) < 1e-6
    assert rel_err(got["X12"], ref["X12"]) < 1e-6
    assert len(got["R_set"]) == len(ref["R_set"]) == fx.shape[1]          # every image was registered
    assert len(got["ba_costs"]) == len(ref["ba_costs"]) == len(ba_log) >= 3
    assert np.array_equal(got["have"], ref["have"])
    # bundle adjustment, call by call on identical inputs, by north_star's criteria: final cost within 1e-6
    # relative, reprojection RMSE within 1e-3 px. The 7-dimensional gauge is free (BundleAdjustment.py:180-199) and
    # the outlier-laden calls have flat directions along which scipy's own ftol = 1e-8 stop leaves ~0.02 px of slack
    # (costs equal to 1e-9), so the individual reprojections only get a sanity bound
    for k, e in enumerate(ba_log):
        assert e["gpu"] < e["start"], (k, e)
        assert abs(e["gpu"] - e["ref"]) / e["ref"] < 1e-6, (k, ba_log)
        rmse = lambda c: np.sqrt(2.0 * c / e["n_obs"])
        assert abs(rmse(e["gpu"]) - rmse(e["ref"])) < 1e-3, (k, ba_log)
        assert e["reproj"] < 0.1, (k, ba_log)
    # the first call starts from bit-comparable inputs in both flows, so there the two FLOWS must agree as well
    assert abs(got["ba_costs"][0] - ref["ba_costs"][0]) / ref["ba_costs"][0] < 1e-6
