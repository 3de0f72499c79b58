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


def test_wrapper_flow_dropins_match_oracle(cuda_device):
    fx, fy, flag, K = wrapper_flow.synthetic_features()
    with redirect_stdout(io.StringIO()), np.errstate(all="ignore"):
        random.seed(11)
        ref = wrapper_flow.run(_oracle_ns(), fx, fy, flag, K)
        state_ref = random.getstate()
        random.seed(11)
        got = wrapper_flow.run(_dropin_ns(), fx, fy, flag, K)
        state_got = random.getstate()
    assert state_got == state_ref                               # the same draws, in the same order
    assert np.array_equal(got["filtered"], ref["filtered"])     # identical RANSAC inlier sets for all pairs
    assert rel_err(got["F_12"], ref["F_12"]) < 1e-8
    assert got["n12"] == ref["n12"] and got["n12"] >= 20
    assert rel_err(got["C2"], ref["C2"]) < 1e-6 and rel_err(got["R2"], ref["R2"]) < 1e-6
    assert rel_err(got["X12"], ref["X12"]) < 1e-6
    assert len(got["R_set"]) == len(ref["R_set"]) == fx.shape[1]          # every image was registered
    assert len(got["ba_costs"]) == len(ref["ba_costs"]) >= 3
    for a, b in zip(got["ba_costs"], ref["ba_costs"]):
        assert abs(a - b) / b < 1e-6, (got["ba_costs"], ref["ba_costs"])
    assert np.array_equal(got["have"], ref["have"])
    # bundle adjustment leaves the 7-dimensional gauge free (BundleAdjustment.py:180-199): compare what is
    # gauge invariant -- the final cost above and the reprojections
    for i in range(fx.shape[1]):
        sel = got["have"][:, 0] == 1
        pg = so.projection_matrix(K, got["C_set"][i], got["R_set"][i]) @ np.hstack([got["world"][sel], np.ones((sel.sum(), 1))]).T
        pr = so.projection_matrix(K, ref["C_set"][i], ref["R_set"][i]) @ np.hstack([ref["world"][sel], np.ones((sel.sum(), 1))]).T
        assert np.max(np.abs(pg[:2] / pg[2] - pr[:2] / pr[2])) < 1e-2      # pixels
