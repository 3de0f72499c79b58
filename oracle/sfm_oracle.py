"""CPU oracle for the Phase-1 SfM hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A numpy/scipy restatement of the reference algorithms on the path BASELINE.json names.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; nothing under structure-from-motion-_b200/ does.

Parity pinning: the reference ships no golden vectors or asserting tests,
so this oracle is pinned against OUTPUTS OF THE REFERENCE ITSELF: oracle/make_golden.py imports
the unmodified modules from /root/reference/Phase 1 in the build container, runs them on
seeded inputs and freezes inputs+outputs under tests/golden/*.npz; tests/test_oracle_golden.py
checks every function below against those files (on the GPU box too, where the reference
does not exist). The LM arithmetic itself lives in scipy 1.18.1 (MINPACK lmder through
scipy.optimize.least_squares) and LAPACK (numpy 2.3.5 / OpenBLAS); neither is vendored by the
reference (README.md:145 pins no versions), so the oracle calls the same library entry points
the reference calls.

All citations are into /root/reference/Phase 1/.
"""
from __future__ import annotations

import random

import numpy as np
from scipy.optimize import least_squares
from scipy.spatial.transform import Rotation

EPS = 1e-8


# ----------------------------------------------------------------------------------------------
# Fundamental matrix: EstimateFundamentalMatrix.py:3-83
# ----------------------------------------------------------------------------------------------
def _hartley(points):
    """Similarity T with centroid at 0 and mean distance sqrt(2) (EstimateFundamentalMatrix.py:30-47)."""
    mean = points.mean(axis=0)
    dist = np.sqrt(((points - mean) ** 2).sum(axis=1))
    s = np.sqrt(2) / (dist.mean() + EPS)
    return np.array([[s, 0, -s * mean[0]], [0, s, -s * mean[1]], [0, 0, 1]])


def estimate_fundamental_matrix(points1, points2):
    points1 = np.array(points1)
    points2 = np.array(points2)
    if points1.shape[1] != 2:  # EstimateFundamentalMatrix.py:80-81
        raise ValueError("Points must be in 2D format (N x 2)")
    n = points1.shape[0]
    T1, T2 = _hartley(points1), _hartley(points2)
    h1 = np.hstack([points1, np.ones((n, 1))])
    h2 = np.hstack([points2, np.ones((n, 1))])
    n1 = (T1 @ h1.T).T
    n2 = (T2 @ h2.T).T
    x1, y1, x2, y2 = n1[:, 0], n1[:, 1], n2[:, 0], n2[:, 1]
    # row layout of EstimateFundamentalMatrix.py:62
    A = np.stack([x1 * x2, x1 * y2, x1, y1 * x2, y1 * y2, y1, x2, y2, np.ones(n)], axis=1)
    _, _, Vt = np.linalg.svd(A)                 # :65
    Fh = Vt[-1].reshape(3, 3)                   # :66-67
    U, S, Vt = np.linalg.svd(Fh)                # :70
    S[2] = 0
    Fh = U @ np.diag(S) @ Vt                    # :72
    F = T2.T @ Fh @ T1                          # :75  (quirk Q1: kept as is)
    return F / F[2, 2]                          # :78


def f_errors(F, points1, points2):
    """Mean of the two point-to-epipolar-line distances, GetInliersRANSAC.py:67-78."""
    n = len(points1)
    h1 = np.hstack([points1, np.ones((n, 1))])
    h2 = np.hstack([points2, np.ones((n, 1))])
    Fx1 = (F @ h1.T).T
    Ftx2 = (F.T @ h2.T).T
    e = np.sum(h2 * Fx1, axis=1)
    d1 = np.abs(e) / (np.sqrt(Fx1[:, 0] ** 2 + Fx1[:, 1] ** 2) + EPS)
    d2 = np.abs(e) / (np.sqrt(Ftx2[:, 0] ** 2 + Ftx2[:, 1] ** 2) + EPS)
    return (d1 + d2) / 2


def draw_samples(n_points, k, n_max):
    """The reference's RNG consumption: one random.sample per hypothesis on the GLOBAL
    generator (GetInliersRANSAC.py:55, PnPRANSAC.py:49)."""
    return np.array([random.sample(range(n_points), min(k, n_points)) for _ in range(n_max)],
                    dtype=np.int32).reshape(n_max, min(k, n_points))


def f_ransac_detailed(points1, points2, threshold, samples):
    """All hypotheses of GetInliersRANSAC.py:53-92 for given sample index sets.
    Returns dict(F[h,9], counts[h], best_idx, best_count, mask[n], err[n])."""
    points1 = np.asarray(points1, dtype=np.float64)
    points2 = np.asarray(points2, dtype=np.float64)
    h = len(samples)
    Fs = np.full((h, 9), np.nan)
    counts = np.zeros(h, dtype=np.int64)
    best_count, best_idx, best_mask, best_err = 0, -1, None, None
    with np.errstate(all="ignore"):
        for i, s in enumerate(samples):
            try:
                F = estimate_fundamental_matrix(points1[s], points2[s])
                err = f_errors(F, points1, points2)
                mask = err < threshold                     # :81 strict
                c = int(mask.sum())
            except Exception:                              # :90-92
                continue
            Fs[i] = F.ravel()
            counts[i] = c
            if c > best_count:                             # :85 strict -> first maximum
                best_count, best_idx, best_mask, best_err = c, i, mask, err
    return dict(F=Fs, counts=counts, best_idx=best_idx, best_count=best_count,
                mask=best_mask, err=best_err)


def get_inliers_ransac_full(points1, points2, index, threshold=0.06, n_max=1000, samples=None):
    """GetInliersRANSAC.py:5-106 -> (inlier_index, outlier_indices, F_best)."""
    points1 = np.array(points1)
    points2 = np.array(points2)
    index = np.array(index)
    n = len(points1)
    if n < 8:                                              # :38-40
        return np.array([]), index, None
    if samples is None:
        samples = draw_samples(n, 8, n_max)
    d = f_ransac_detailed(points1, points2, threshold, samples)
    if d["best_count"] == 0:                               # :95-96
        return np.array([]), index, None
    inl = np.where(d["mask"])[0]
    return inl, index[~d["mask"]], d["F"][d["best_idx"]].reshape(3, 3)


def get_inliers_ransac(points1, points2, index, threshold=0.06, n_max=1000, samples=None):
    """Alias with the wrapper's return order, GetInliersRANSAC.py:109-121."""
    index = np.array(index)
    inl, _, F = get_inliers_ransac_full(points1, points2, index, threshold, n_max, samples)
    return F, (index[inl] if len(inl) > 0 else np.array([]))


# ----------------------------------------------------------------------------------------------
# Non-linear triangulation: NonLinearTriangulation.py:5-130
# ----------------------------------------------------------------------------------------------
def projection_matrix(K, C, R):
    """P = K [R | -R C]  (NonLinearTriangulation.py:96-100, BundleAdjustment.py:32)."""
    K, C, R = np.asarray(K, float), np.asarray(C, float), np.asarray(R, float)
    return K @ np.hstack([R, -R @ C.reshape(3, 1)])


def tri_loss(X, x1, x2, P1, P2):
    """Loss(), NonLinearTriangulation.py:5-50: exact division, view residual 0 when |w|<1e-8."""
    Xh = np.array([X[0], X[1], X[2], 1.0])
    out = np.empty(4)
    for k, (P, x) in enumerate(((P1, x1), (P2, x2))):
        v = P @ Xh
        proj = np.array([x[0], x[1]]) if abs(v[2]) < 1e-8 else v[:2] / v[2]
        out[2 * k:2 * k + 2] = x - proj
    return out


def nonlinear_triangulation(K, C1, R1, C2, R2, x1, x2, x0, tight=False):
    """NonLinearTriangulation.py:53-130; tight=True is the 'converged oracle' of SURVEY 8(c)."""
    x1, x2, x0 = np.array(x1, float), np.array(x2, float), np.array(x0, float)
    P1, P2 = projection_matrix(K, C1, R1), projection_matrix(K, C2, R2)
    kw = dict(ftol=1e-15, xtol=1e-15, gtol=1e-15, max_nfev=2000) if tight else dict(max_nfev=50)
    out = []
    for i in range(len(x1)):
        try:
            r = least_squares(lambda X: tri_loss(X, x1[i], x2[i], P1, P2), x0[i], method="lm",
                              verbose=0, **kw)
            out.append(r.x)
        except Exception:                                  # :126-128
            out.append(x0[i])
    return np.array(out)


def tri_cost(K, C1, R1, C2, R2, x1, x2, X):
    """0.5*|Loss|^2 per point (the quantity LM minimises)."""
    P1, P2 = projection_matrix(K, C1, R1), projection_matrix(K, C2, R2)
    return np.array([0.5 * np.sum(tri_loss(X[i], x1[i], x2[i], P1, P2) ** 2) for i in range(len(X))])


# ----------------------------------------------------------------------------------------------
# Non-linear PnP: NonlinearPnP.py:5-123
# ----------------------------------------------------------------------------------------------
def pnp_loss(X0, X, x, K):
    """NonLinearPnPLoss, NonlinearPnP.py:5-44: interleaved [du,dv]; +1e-8 in the denominator."""
    R = Rotation.from_rotvec(X0[:3]).as_matrix()
    C = -R.T @ X0[3:6]
    P = K @ np.hstack([R, -R @ C.reshape(3, 1)])
    ph = (P @ np.hstack([X, np.ones((X.shape[0], 1))]).T).T
    return (x - ph[:, :2] / (ph[:, 2:3] + EPS)).flatten()


def pose_to_params(C, R):
    """x0 = [rotvec, -R C]  (NonlinearPnP.py:87-93, BundleAdjustment.py:188-193)."""
    R = np.asarray(R, float)
    return np.hstack([Rotation.from_matrix(R).as_rotvec(), -R @ np.asarray(C, float)])


def params_to_pose(p):
    R = Rotation.from_rotvec(p[:3]).as_matrix()
    return -R.T @ p[3:6], R


def nonlinear_pnp(X, x, K, C, R, tight=False):
    """NonLinearPnP, NonlinearPnP.py:47-123 -> (Cnew, Rnew)."""
    X, x, K, C, R = (np.array(a, float) for a in (X, x, K, C, R))
    if len(X) < 4:                                         # :82-84
        return C, R
    p0 = pose_to_params(C, R)                              # outside the try, like :87
    kw = dict(ftol=1e-15, xtol=1e-15, gtol=1e-15, max_nfev=5000) if tight else dict(max_nfev=100)
    try:
        r = least_squares(pnp_loss, p0, args=(X, x, K), method="lm", verbose=0, **kw)
        return params_to_pose(r.x)
    except Exception as e:                                 # :120-123
        print(f"Non-linear PnP optimization failed: {e}")
        return C, R


# ----------------------------------------------------------------------------------------------
# PnP RANSAC + LinearPnP: PnPRANSAC.py:6-89, LinearPnP.py:3-96
# ----------------------------------------------------------------------------------------------
def linear_pnp(Xset, xset, K):
    Xset, xset, K = np.array(Xset, float), np.array(xset, float), np.array(K, float)
    n = len(Xset)
    if n < 4:
        raise ValueError("At least 4 point correspondences are required for PnP")
    xn = (np.linalg.inv(K) @ np.hstack([xset, np.ones((n, 1))]).T).T        # :36-38
    A = np.zeros((2 * n, 12))
    Xh = np.hstack([Xset, np.ones((n, 1))])
    A[0::2, 0:4] = Xh
    A[0::2, 8:12] = -xn[:, 0:1] * Xh
    A[1::2, 4:8] = Xh
    A[1::2, 8:12] = -xn[:, 1:2] * Xh                                         # :45-53
    _, _, Vt = np.linalg.svd(A)
    P = Vt[-1].reshape(3, 4)
    M, t = P[:, :3], P[:, 3]
    if np.linalg.det(M) < 0:                                                 # :65-67
        M, t = -M, -t
    Q, Rqr = np.linalg.qr(M.T)                                               # :73
    R = Q.T
    scale = np.mean(np.diag(Rqr.T))
    if scale < 0:
        R, scale = -R, -scale
    U, _, Vt = np.linalg.svd(R)                                              # :84
    R = U @ Vt
    if np.linalg.det(R) < 0:
        U[:, -1] *= -1
        R = U @ Vt
    return -R.T @ (t / scale), R                                             # :93-94


def pnp_errors(P, X, x):
    """Reprojection distance of PnPRANSAC.py:59-65 for one 3x4 P."""
    ph = (P @ np.hstack([X, np.ones((len(X), 1))]).T).T
    proj = ph[:, :2] / (ph[:, 2:3] + EPS)
    return np.sqrt(np.sum((x - proj) ** 2, axis=1))


def pnp_ransac_detailed(X, x, K, threshold, samples):
    """All hypotheses of PnPRANSAC.py:47-79: poses from linear_pnp on each 4-sample, counts."""
    X, x, K = np.array(X, float), np.array(x, float), np.array(K, float)
    h = len(samples)
    Ps = np.full((h, 12), np.nan)
    counts = np.zeros(h, dtype=np.int64)
    poses = [None] * h
    best_count, best_idx = 0, -1
    with np.errstate(all="ignore"):
        for i, s in enumerate(samples):
            try:
                C, R = linear_pnp(X[s], x[s], K)
                P = K @ np.hstack([R, -R @ C.reshape(3, 1)])
                c = int((pnp_errors(P, X, x) < threshold).sum())
            except Exception:
                continue
            Ps[i], counts[i], poses[i] = P.ravel(), c, (C, R)
            if c > best_count:
                best_count, best_idx = c, i
    return dict(P=Ps, counts=counts, poses=poses, best_idx=best_idx, best_count=best_count)


def pnp_ransac(X, x, K, threshold=200, n_max=1000, samples=None):
    X, x, K = np.array(X, float), np.array(x, float), np.array(K, float)
    n = len(X)
    if n < 4:                                                                # :38-39
        raise ValueError("At least 4 point correspondences are required for PnP")
    if samples is None:
        samples = draw_samples(n, 4, n_max)
    d = pnp_ransac_detailed(X, x, K, threshold, samples)
    if d["best_idx"] < 0 or d["best_count"] < 4:                             # :82-87
        try:
            C, R = linear_pnp(X, x, K)
            print("Warning: PnP RANSAC failed, using linear PnP on all points")
            return C, R
        except Exception as e:
            raise ValueError(f"PnP estimation failed: {e}")
    return d["poses"][d["best_idx"]]


# ----------------------------------------------------------------------------------------------
# Linear triangulation: LinearTriangulation.py:3-92 (pinned by tests/golden/nltri.npz "x0")
# ----------------------------------------------------------------------------------------------
def linear_triangulation(K, C1, R1, C2, R2, x1, x2):
    K, C1, R1, C2, R2 = (np.array(a, float) for a in (K, C1, R1, C2, R2))
    x1, x2 = np.array(x1, float), np.array(x2, float)
    P1 = K @ np.hstack([R1, -R1 @ C1.reshape(3, 1)])               # :44-50
    P2 = K @ np.hstack([R2, -R2 @ C2.reshape(3, 1)])
    out = []
    for (u1, v1), (u2, v2) in zip(x1, x2):
        A = np.stack([v1 * P1[2] - P1[1], P1[0] - u1 * P1[2],      # :76-81
                      v2 * P2[2] - P2[1], P2[0] - u2 * P2[2]])
        Xh = np.linalg.svd(A)[2][-1]                               # :84-85
        out.append(Xh[:3] / Xh[3] if abs(Xh[3]) > 1e-8 else Xh[:3])   # :88-92
    return np.array(out)


# ------------------------------------------------------------------ two-view pose front end (N4)
def essential_from_fundamental(F, K):
    """EssentialMatrixFromFundamentalMatrix.py:19."""
    return np.asarray(K).T @ np.asarray(F) @ np.asarray(K)


def extract_camera_pose(E):
    """ExtractCameraPose.py:24-81: four (C, R) configurations from numpy's SVD of E. The ORDER of the
    list follows LAPACK's sign choice for the singular vectors; the set does not."""
    U, _, Vt = np.linalg.svd(np.array(E))
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        U[:, -1] *= -1
    W = np.array([[0, -1, 0], [1, 0, 0], [0, 0, 1]])
    R1, R2 = U @ W @ Vt, U @ W.T @ Vt
    t = U[:, 2] / np.linalg.norm(U[:, 2])
    Cset = np.array([-R1.T @ t, -R1.T @ (-t), -R2.T @ t, -R2.T @ (-t)])
    Rset = np.array([R1, R1, R2, R2])
    return Cset, Rset


def cheirality_counts(Cset, Rset, Xset):
    """DisambiguateCameraPose.py:66-76: points with positive depth in both cameras, per configuration."""
    out = []
    for C, R, X in zip(Cset, Rset, Xset):
        X = np.asarray(X).reshape(-1, 3)
        z2 = ((R @ (X - np.asarray(C).reshape(1, 3)).T).T)[:, 2] if len(X) else np.zeros(0)
        out.append(int(np.sum((X[:, 2] > 0) & (z2 > 0))) if len(X) else 0)
    return np.array(out)


def disambiguate_camera_pose(Cset, Rset, Xset):
    """DisambiguateCameraPose.py:53-92: first maximum of the vote (strict '>' from -1, empty sets skipped)."""
    counts = cheirality_counts(Cset, Rset, Xset)
    best, best_count = None, -1
    for i in range(len(Cset)):
        if len(Xset[i]) == 0:
            continue
        if counts[i] > best_count:
            best, best_count = i, counts[i]
    if best is None:
        return np.asarray(Cset[0]), np.asarray(Rset[0]), np.asarray(Xset[0]), 0, -1
    return np.array(Cset[best]), np.array(Rset[best]), np.array(Xset[best]), best, int(best_count)


# ------------------------------------------------------------------ homography RANSAC (N2)
def find_homography(image1_coords, image2_coords):
    """GetHomographyInliers.py:4-82."""
    p1, p2 = np.array(image1_coords, float), np.array(image2_coords, float)
    n = len(p1)
    if n < 4:
        raise ValueError("At least 4 point correspondences are required for homography estimation")
    m1, m2 = np.mean(p1, axis=0), np.mean(p2, axis=0)
    c1, c2 = p1 - m1, p2 - m2
    s1 = np.sqrt(2) / (np.mean(np.sqrt(np.sum(c1 ** 2, axis=1))) + 1e-8)
    s2 = np.sqrt(2) / (np.mean(np.sqrt(np.sum(c2 ** 2, axis=1))) + 1e-8)
    T1 = np.array([[s1, 0, -s1 * m1[0]], [0, s1, -s1 * m1[1]], [0, 0, 1]])
    T2 = np.array([[s2, 0, -s2 * m2[0]], [0, s2, -s2 * m2[1]], [0, 0, 1]])
    ones = np.ones((n, 1))
    n1 = (T1 @ np.hstack([p1, ones]).T).T
    n2 = (T2 @ np.hstack([p2, ones]).T).T
    A = np.zeros((2 * n, 9))
    x1, y1, x2, y2 = n1[:, 0], n1[:, 1], n2[:, 0], n2[:, 1]
    A[0::2, 3], A[0::2, 4], A[0::2, 5] = -x1, -y1, -1
    A[0::2, 6], A[0::2, 7], A[0::2, 8] = y2 * x1, y2 * y1, y2
    A[1::2, 0], A[1::2, 1], A[1::2, 2] = x1, y1, 1
    A[1::2, 6], A[1::2, 7], A[1::2, 8] = -x2 * x1, -x2 * y1, -x2
    _, _, Vt = np.linalg.svd(A)
    H = np.linalg.inv(T2) @ Vt[-1, :].reshape(3, 3) @ T1
    return H / H[2, 2]


def h_errors(H, points1, points2):
    """GetHomographyInliers.py:139-146."""
    n = len(points1)
    t = (H @ np.hstack([points1, np.ones((n, 1))]).T).T
    t2 = t[:, :2] / (t[:, 2:3] + 1e-8)
    return np.sqrt(np.sum((t2 - points2) ** 2, axis=1))


def h_ransac_detailed(points1, points2, threshold, samples):
    """The loop of GetHomographyInliers.py:124-161 on given 4-index sample sets."""
    p1, p2 = np.array(points1, float), np.array(points2, float)
    h = len(samples)
    Hs, counts = np.full((h, 9), np.nan), np.zeros(h, dtype=np.int64)
    best_count, best_idx, mask_best, err_best = 0, -1, np.zeros(len(p1), bool), None
    for k, s in enumerate(samples):
        try:
            with np.errstate(all="ignore"):
                H = find_homography(p1[list(s)], p2[list(s)])
                err = h_errors(H, p1, p2)
            mask = err < threshold
        except Exception:
            continue
        Hs[k], counts[k] = H.ravel(), int(mask.sum())
        if counts[k] > best_count:
            best_count, best_idx, mask_best, err_best = int(counts[k]), k, mask, err
    return dict(H=Hs, counts=counts, best_idx=best_idx, best_count=best_count, mask=mask_best, err=err_best)


def get_homography_inliers(image1_coords_org, image2_coords_org, idx, threshold=30, n_max=1000):
    """GetHomographyInliers.py:85-165 drawing from the global `random` stream (:126)."""
    p1, p2, idx = np.array(image1_coords_org), np.array(image2_coords_org), np.array(idx)
    n = len(p1)
    if n < 4:
        return None, np.array([])
    samples = draw_samples(n, 4, n_max)
    r = h_ransac_detailed(p1, p2, threshold, samples)
    if r["best_idx"] < 0 or r["best_count"] == 0:
        return None, np.array([])
    return r["H"][r["best_idx"]].reshape(3, 3), idx[np.where(r["mask"])[0]]
