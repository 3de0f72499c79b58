"""CPU oracle for bundle adjustment -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see sfm_oracle.py).

Restates BundleAdjustment.py (/root/reference/Phase 1/) with a vectorised residual so that it
also runs at sizes where the reference's per-observation Python loop (BundleAdjustment.py:95-108)
and O(Nobs*Np) observation-list build (:164-169) cannot. Pinned by tests/golden/ba_*.npz, which
hold the unmodified reference's residual vectors and perform_bundle_adjustment outputs.
"""
from __future__ import annotations

import numpy as np
from scipy.optimize import least_squares
from scipy.sparse import lil_matrix
from scipy.spatial.transform import Rotation

EPS = 1e-8


def project_points(K, C, R, X):
    """BundleAdjustment.py:8-40."""
    X = np.asarray(X, float)
    P = K @ np.hstack([R, -R @ np.asarray(C, float).reshape(3, 1)])
    xh = (P @ np.hstack([X, np.ones((X.shape[0], 1))]).T).T
    return xh[:, :2] / (xh[:, 2:3] + EPS)


def ba_residuals(params, n_cameras, n_points, camera_indices, point_indices, points_2d, K,
                 n_cam_params=6):
    """bundle_adjustment_residuals, BundleAdjustment.py:43-110, vectorised over observations.
    Layout (:73-76): [rotvec(3), t(3)] x n_cameras, then [X,Y,Z] x n_points; output interleaved
    [du, dv] per observation, observed - projected (:107-108)."""
    params = np.asarray(params, float)
    cams = params[:n_cameras * n_cam_params].reshape(n_cameras, n_cam_params)
    pts = params[n_cameras * n_cam_params:].reshape(n_points, 3)
    R = Rotation.from_rotvec(cams[:, :3]).as_matrix().reshape(n_cameras, 3, 3)
    C = -np.einsum("nji,nj->ni", R, cams[:, 3:6])          # C = -R^T t   (:88)
    t = -np.einsum("nij,nj->ni", R, C)                     # -R C         (project_points :32)
    P = np.einsum("ij,njk->nik", K, np.concatenate([R, t[:, :, None]], axis=2))
    Xh = np.hstack([pts[point_indices], np.ones((len(point_indices), 1))])
    xh = np.einsum("oij,oj->oi", P[camera_indices], Xh)
    proj = xh[:, :2] / (xh[:, 2:3] + EPS)
    return (np.asarray(points_2d, float) - proj).reshape(-1)


def ba_build_problem(all_world_coords, filtered_world_coords, feature_x, feature_y,
                     filtered_feature_flags, R_set, C_set):
    """Observation list and x0 of perform_bundle_adjustment (BundleAdjustment.py:148-199):
    valid points ascending, observations point-major then camera ascending, camera block
    [as_rotvec(R), -R C]. Returns None when there is nothing to optimise (:152-153, :171-172)."""
    valid = np.where(np.asarray(filtered_world_coords).flatten() == 1)[0]
    if len(valid) == 0:
        return None
    n_cameras = len(R_set)                                  # cam_index is unused (:155-156)
    flags = np.asarray(filtered_feature_flags)[valid][:, :n_cameras] == 1
    pt_local, cam = np.nonzero(flags)                       # row-major = point-major (:164-169)
    if len(cam) == 0:
        return None
    fx, fy = np.asarray(feature_x), np.asarray(feature_y)
    pts2d = np.stack([fx[valid[pt_local], cam], fy[valid[pt_local], cam]], axis=1).astype(float)
    x0 = []
    for i in range(n_cameras):
        R = np.array(R_set[i])
        Cc = np.array(C_set[i])
        x0.extend(Rotation.from_matrix(R).as_rotvec())      # :188 (outside the try)
        x0.extend(-R @ Cc)                                  # :191
    x0 = np.concatenate([np.array(x0, float), np.asarray(all_world_coords, float)[valid].reshape(-1)])
    return dict(valid=valid, n_cameras=n_cameras, n_points=len(valid), cam_idx=cam.astype(np.int64),
                pt_idx=pt_local.astype(np.int64), pts2d=pts2d, x0=x0)


def ba_sparsity(n_cameras, n_points, cam_idx, pt_idx):
    m, n = 2 * len(cam_idx), 6 * n_cameras + 3 * n_points
    A = lil_matrix((m, n), dtype=int)
    o = np.arange(len(cam_idx))
    for s in range(6):
        A[2 * o, cam_idx * 6 + s] = 1
        A[2 * o + 1, cam_idx * 6 + s] = 1
    for s in range(3):
        A[2 * o, 6 * n_cameras + pt_idx * 3 + s] = 1
        A[2 * o + 1, 6 * n_cameras + pt_idx * 3 + s] = 1
    return A


def ba_solve(x0, n_cameras, n_points, cam_idx, pt_idx, pts2d, K, method="lm", tight=False,
             max_nfev=None):
    """least_squares exactly as BundleAdjustment.py:205-212 (method='lm', max_nfev=100), or the
    scalable secondary oracle (method='trf' with the block sparsity) for medium sizes."""
    args = (n_cameras, n_points, cam_idx, pt_idx, pts2d, K)
    kw = {}
    if tight:
        kw.update(ftol=1e-15, xtol=1e-15, gtol=1e-15)
    if method == "lm":
        kw["max_nfev"] = max_nfev if max_nfev is not None else (100000 if tight else 100)
        return least_squares(ba_residuals, x0, args=args, method="lm", verbose=0, **kw)
    kw["max_nfev"] = max_nfev if max_nfev is not None else 200
    return least_squares(ba_residuals, x0, args=args, method="trf", x_scale="jac",
                         jac_sparsity=ba_sparsity(n_cameras, n_points, cam_idx, pt_idx),
                         verbose=0, **kw)


def ba_unpack(x, n_cameras):
    R_set, C_set = [], []
    for i in range(n_cameras):
        R = Rotation.from_rotvec(x[6 * i:6 * i + 3]).as_matrix()
        R_set.append(R)
        C_set.append(-R.T @ x[6 * i + 3:6 * i + 6])
    return R_set, C_set, x[6 * n_cameras:].reshape(-1, 3)


def perform_bundle_adjustment(all_world_coords, filtered_world_coords, feature_x, feature_y,
                              filtered_feature_flags, R_set, C_set, K, cam_index, info=None):
    """BundleAdjustment.py:113-242 including its 'return the inputs' failure convention."""
    prob = ba_build_problem(all_world_coords, filtered_world_coords, feature_x, feature_y,
                            filtered_feature_flags, R_set, C_set)
    if prob is None:
        return R_set, C_set, all_world_coords
    print(f"  Bundle adjustment: {prob['n_cameras']} cameras, {prob['n_points']} points, "
          f"{len(prob['cam_idx'])} observations")
    try:
        res = ba_solve(prob["x0"], prob["n_cameras"], prob["n_points"], prob["cam_idx"],
                       prob["pt_idx"], prob["pts2d"], np.asarray(K, float))
        R_opt, C_opt, pts = ba_unpack(res.x, prob["n_cameras"])
        out = np.array(all_world_coords, copy=True)
        out[prob["valid"]] = pts
        print(f"  Bundle adjustment completed. Final cost: {res.cost:.6f}")
        if info is not None:
            info.update(cost=res.cost, nfev=res.nfev, status=res.status, x=res.x)
        return R_opt, C_opt, out
    except Exception as e:
        print(f"  Bundle adjustment failed: {e}")
        return R_set, C_set, all_world_coords
