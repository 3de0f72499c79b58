"""BASELINE config 1 ("P3Data 6-image Wrapper_dev.py pipeline, reference runs on CPU") as a bench
object: the wrapper's control flow (tests/wrapper_flow.py, Wrapper_dev.py:68-312 without drawing) on
seeded in-memory feature arrays -- P3Data itself lives in the reference tree and does not travel to the
GPU box -- driven once by the GPU drop-in modules and once by the CPU oracle port of the reference, on
the same `random` stream. Only bench.py's cpu_baseline leg reaches the oracle."""
from __future__ import annotations

import importlib
import io
import os
import random
import sys
import time
import types
from contextlib import redirect_stdout

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
DROPIN = os.path.join(ROOT, "structure-from-motion-_b200", "dropin")
N_IMG, N_PTS, N_MAX, F_THR, SEED = 5, 120, 1000, 0.5, 11


def _flow():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        import wrapper_flow
    finally:
        sys.path.pop(0)
    return wrapper_flow


def _no_cost(*a, **k):
    return 0.0


def dropin_namespace():
    sys.path.insert(0, DROPIN)
    try:
        m = {}
        for name in ("GetHomographyInliers", "GetInliersRANSAC", "EssentialMatrixFromFundamentalMatrix",
                     "ExtractCameraPose", "DisambiguateCameraPose", "LinearTriangulation", "NonLinearTriangulation",
                     "NonlinearPnP", "PnPRANSAC", "BundleAdjustment"):
            sys.modules.pop(name, None)
            m[name] = importlib.import_module(name)
    finally:
        sys.path.remove(DROPIN)
    return types.SimpleNamespace(
        get_homography_inliers=m["GetHomographyInliers"].get_homography_inliers,
        get_inliers_ransac=m["GetInliersRANSAC"].get_inliers_ransac,
        EssentialMatrixFromFundamentalMatrix=m["EssentialMatrixFromFundamentalMatrix"].EssentialMatrixFromFundamentalMatrix,
        ExtractCameraPose=m["ExtractCameraPose"].ExtractCameraPose,
        DisambiguateCameraPose=m["DisambiguateCameraPose"].DisambiguateCameraPose,
        linear_triangulation=m["LinearTriangulation"].linear_triangulation,
        nonlinear_triangulation=m["NonLinearTriangulation"].nonlinear_triangulation,
        PnPRANSAC=m["PnPRANSAC"].PnPRANSAC, nonlinear_PnP=m["NonlinearPnP"].nonlinear_PnP,
        perform_bundle_adjustment=m["BundleAdjustment"].perform_bundle_adjustment, ba_cost=_no_cost)


def oracle_namespace():
    from oracle import ba_oracle as bo
    from oracle import sfm_oracle as so
    return types.SimpleNamespace(
        get_homography_inliers=so.get_homography_inliers,
        get_inliers_ransac=lambda p1, p2, idx, threshold, n_max: so.get_inliers_ransac(p1, p2, idx, threshold, n_max),
        EssentialMatrixFromFundamentalMatrix=so.essential_from_fundamental, ExtractCameraPose=so.extract_camera_pose,
        linear_triangulation=so.linear_triangulation,
        DisambiguateCameraPose=lambda C, R, X: so.disambiguate_camera_pose(C, R, X)[:3],
        nonlinear_triangulation=so.nonlinear_triangulation,
        PnPRANSAC=lambda X, x, K, threshold, n_max: so.pnp_ransac(X, x, K, threshold, n_max),
        nonlinear_PnP=lambda K, C, R, x, X: so.nonlinear_pnp(X, x, K, C, R),
        perform_bundle_adjustment=bo.perform_bundle_adjustment, ba_cost=_no_cost)


def _timed(ns, wf, fx, fy, flag, K):
    random.seed(SEED)
    t = time.perf_counter()
    with redirect_stdout(io.StringIO()), np.errstate(all="ignore"):
        log = wf.run(ns, fx, fy, flag, K, n_max=N_MAX, f_threshold=F_THR)
    return time.perf_counter() - t, log


def run_p3data(seed: int = 0, runs: int = 3) -> dict | None:
    """BASELINE configs[0] itself: the UNMODIFIED Wrapper_dev.sfm_wrapper on P3Data at the wrapper's own settings
    (F threshold 0.06 px, 1000 hypotheses), once as shipped (reference CPU code from oracle/_ref) and once with the
    drop-in directory ahead of it on sys.path. Drawing is stubbed in both; Utils.get_data (text parsing, reference
    code) is inside both timings. None when oracle/_ref did not travel."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        import wrapper_harness as wh
    finally:
        sys.path.pop(0)
    if not wh.available():
        return None
    wh.run(True, seed)                                              # warm
    torch.cuda.synchronize()
    ts = []
    for _ in range(runs):
        t = time.perf_counter()
        got, _ = wh.run(True, seed)
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t)
    t = time.perf_counter()
    ref, _ = wh.run(False, seed)
    t_ref = time.perf_counter() - t
    same = all(np.array_equal(np.asarray(a[2][1]), np.asarray(b[2][1]))
               for a, b in zip(got["get_inliers_ransac"], ref["get_inliers_ransac"]))
    Xg, Xr = got["nonlinear_triangulation"][0][2], ref["nonlinear_triangulation"][0][2]
    return {"workload": "C1: unmodified Wrapper_dev.sfm_wrapper on P3Data (5 images, 10 pairs: homography RANSAC 30 px / "
                        "1000 + F RANSAC 0.06 px / 1000 per pair, two-view pose, linear + non-linear triangulation of "
                        f"{len(Xr)} points; PnP and bundle adjustment are not reached on this data), seed {seed}",
            "api": "dropin/ ahead of the reference on sys.path; nothing else changes",
            "seconds": float(np.mean(ts)), "seconds_min": float(min(ts)), "runs": runs,
            "cpu_baseline": {"seconds": t_ref, "cores": 1, "kind": "reference",
                             "sample": "the same call as shipped, one run"},
            "speedup_vs_reference": t_ref / float(np.mean(ts)),
            "parity": {"f_inlier_sets_equal_all_pairs": bool(same), "rng_state_equal": got["rng_after"] == ref["rng_after"],
                       "refined_points_rel_err": float(np.linalg.norm(Xg - Xr) / np.linalg.norm(Xr))}}


def run(with_cpu: bool = True) -> dict:
    import torch
    wf = _flow()
    fx, fy, flag, K = wf.synthetic_features(n_img=N_IMG, n_pts=N_PTS, seed=5)
    ns = dropin_namespace()
    _timed(ns, wf, fx, fy, flag, K)                                  # warm (library load, allocator)
    torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        dt, got = _timed(ns, wf, fx, fy, flag, K)
        ts.append(dt)
    out = {"workload": f"C1-shaped incremental SfM: {N_IMG} images, {N_PTS} tracks, wrapper flow "
                       f"(homography + F RANSAC per pair at {N_MAX} hypotheses, two-view pose, PnP registration, "
                       f"triangulation, bundle adjustment after every pair); F threshold {F_THR} px "
                       "(the wrapper's 0.06 px leaves too few inliers to register an image)",
           "api": "the drop-in modules bound by bare name, host numpy in/out at every call",
           "seconds": float(np.mean(ts)), "seconds_min": float(min(ts)), "runs": len(ts),
           "registered_images": len(got["R_set"]), "ba_calls": len(got["ba_costs"])}
    if with_cpu:
        dt, ref = _timed(oracle_namespace(), wf, fx, fy, flag, K)
        same = bool(np.array_equal(ref["filtered"], got["filtered"]) and len(ref["R_set"]) == len(got["R_set"]))
        out["cpu_baseline"] = {"seconds": dt, "cores": 1, "kind": "port",
                               "sample": "the same flow, one run, oracle port of the reference (numpy/scipy)",
                               "same_inlier_sets_and_registrations": same}
        out["speedup_vs_cpu_port"] = dt / out["seconds"]
        p3 = run_p3data()
        if p3 is not None:
            out["p3data"] = p3
    return out


if __name__ == "__main__":
    import json
    print(json.dumps(run()))
