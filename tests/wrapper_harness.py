"""Runs the UNMODIFIED reference driver -- Wrapper_dev.sfm_wrapper (Wrapper_dev.py:20) on P3Data -- either as shipped
or with the drop-in directory ahead of the reference on sys.path, which is the whole integration: the wrapper binds its
hot functions by module name at import (Wrapper_dev.py:5-14). TEST INFRASTRUCTURE.

The reference tree is staged under oracle/_ref by oracle/stage_ref.py (git-ignored, travels to the GPU box). Drawing
functions are stubbed (no matplotlib on the boxes; cv2 drawing is irrelevant), every hot call is recorded by wrapping
the NAMES THE WRAPPER BOUND (module attributes of Wrapper_dev), so the wrapper's own control flow decides what runs."""
import io
import os
import random
import sys
import tempfile
from contextlib import redirect_stdout

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "structure-from-motion-_b200", "dropin")
sys.path.insert(0, ROOT) if ROOT not in sys.path else None

from oracle import stage_ref  # noqa: E402

# names Wrapper_dev binds from the hot path (Wrapper_dev.py:5-14; get_homography_inliers via `import *`)
HOT = ("get_homography_inliers", "get_inliers_ransac", "EssentialMatrixFromFundamentalMatrix", "ExtractCameraPose",
       "linear_triangulation", "nonlinear_triangulation", "PnPRANSAC", "nonlinear_PnP", "perform_bundle_adjustment")
# where each of them lives in the product package
HOME = {"get_homography_inliers": "sfm_b200.homography", "get_inliers_ransac": "sfm_b200.ransac",
        "EssentialMatrixFromFundamentalMatrix": "sfm_b200.two_view", "ExtractCameraPose": "sfm_b200.two_view",
        "linear_triangulation": "sfm_b200.refine", "nonlinear_triangulation": "sfm_b200.refine",
        "PnPRANSAC": "sfm_b200.pnp_ransac", "nonlinear_PnP": "sfm_b200.refine",
        "perform_bundle_adjustment": "sfm_b200.bundle"}
DRAW = ("draw_feature_matches", "plot_world_coords", "draw_reprojections", "plot_camera_pose")


def available():
    stage_ref.stage()
    return stage_ref.available()


def import_wrapper(dropin: bool):
    """The unmodified Wrapper_dev module, freshly imported with sys.path = [dropin/ (optional), oracle/_ref/phase1]."""
    mods = stage_ref.load_modules(("Wrapper_dev",), first_on_path=[DROPIN] if dropin else [])
    return mods["Wrapper_dev"]


def _copy(v):
    if isinstance(v, np.ndarray):
        return v.copy()
    if isinstance(v, (list, tuple)):
        return [_copy(u) for u in v]
    return v


def run(dropin: bool, seed: int):
    """-> (log, W): log[name] = list of (args, result) per call, log['get_data'] = the feature arrays."""
    W = import_wrapper(dropin)
    log = {n: [] for n in HOT + ("get_correct_camera_pose", "get_data")}

    def spy(name):
        fn = getattr(W, name)

        def wrapped(*a, **k):
            out = fn(*a, **k)
            log[name].append((_copy(list(a)), dict(k), _copy(out)))
            return out
        return wrapped
    for n in HOT + ("get_correct_camera_pose", "get_data"):
        setattr(W, n, spy(n))
    for n in DRAW:
        setattr(W, n, lambda *a, **k: None)
    random.seed(seed)
    with tempfile.TemporaryDirectory() as tmp, redirect_stdout(io.StringIO()), np.errstate(all="ignore"):
        W.sfm_wrapper(stage_ref.data_path(), os.path.join(tmp, "results"))
    log["rng_after"] = random.random()
    return log, W
