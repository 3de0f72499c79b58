"""Host side of PnP RANSAC: mirrors PnPRANSAC.py:6-89 and LinearPnP.py:3-96.

Consensus scoring and the first-maximum argmax run on the GPU (K6 + K3). Hypothesis generation
(LinearPnP on 4-point samples) stays on the host for now: with 4 points the 8x12 DLT system has a
4-dimensional null space and the reference's Vt[-1] is whatever LAPACK's dgesdd returns (SURVEY.md
8.1-Q6), so hypothesis parity is only defined by calling the same LAPACK routine on the same
matrix; this module does exactly that, batched over all hypotheses (numpy's stacked svd/qr/det call
LAPACK once per matrix, so each hypothesis equals the reference's per-sample result)."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, device, sampler


def linear_pnp_batch(Xs, xs, K):
    """LinearPnP.py:3-96 for a stack of problems. Xs [H,m,3], xs [H,m,2] -> (C [H,3], R [H,3,3])."""
    Xs, xs, K = np.asarray(Xs, float), np.asarray(xs, float), np.asarray(K, float)
    H, m = Xs.shape[0], Xs.shape[1]
    if m < 4:
        raise ValueError("At least 4 point correspondences are required for PnP")
    Kinv = np.linalg.inv(K)
    x_hom = np.concatenate([xs, np.ones((H, m, 1))], axis=2)
    xn = np.einsum("ij,hmj->hmi", Kinv, x_hom)                       # :36-38
    Xh = np.concatenate([Xs, np.ones((H, m, 1))], axis=2)
    A = np.zeros((H, 2 * m, 12))
    A[:, 0::2, 0:4] = Xh
    A[:, 0::2, 8:12] = -xn[:, :, 0:1] * Xh
    A[:, 1::2, 4:8] = Xh
    A[:, 1::2, 8:12] = -xn[:, :, 1:2] * Xh                           # :45-53
    _, _, Vt = np.linalg.svd(A)                                      # :56
    P = Vt[:, -1, :].reshape(H, 3, 4)
    M, t = P[:, :, :3].copy(), P[:, :, 3].copy()
    neg = np.linalg.det(M) < 0                                       # :65-67
    M[neg] *= -1
    t[neg] *= -1
    Q, Rqr = np.linalg.qr(np.transpose(M, (0, 2, 1)))                # :73
    R = np.transpose(Q, (0, 2, 1))
    scale = np.mean(np.diagonal(Rqr, axis1=1, axis2=2), axis=1)      # diag(Rqr.T) == diag(Rqr)
    flip = scale < 0
    R[flip] *= -1
    scale = np.where(flip, -scale, scale)
    U, _, Vt2 = np.linalg.svd(R)                                     # :84
    R = U @ Vt2
    bad = np.linalg.det(R) < 0
    if np.any(bad):
        U[bad, :, -1] *= -1
        R[bad] = U[bad] @ Vt2[bad]
    C_ = -np.einsum("hji,hj->hi", R, t / scale[:, None])             # :93-94
    return C_, R


def LinearPnP(Xset, xset, K):
    """LinearPnP.py:3-96 for one problem (any N >= 4)."""
    Xset, xset = np.array(Xset), np.array(xset)
    if len(Xset) < 4:
        raise ValueError("At least 4 point correspondences are required for PnP")
    C_, R = linear_pnp_batch(Xset[None], xset[None], K)
    return C_[0], R[0]


def pnp_score_device(d_P, d_X, d_x, threshold):
    """K6 + K3 on device tensors -> (d_counts [h] i32, d_best [2] i32)."""
    lib = _lib.load()
    h, n = int(d_P.shape[0]), int(d_X.shape[0])
    d_counts = device.empty((h,), torch.int32)
    d_best = device.empty((2,), torch.int32)
    st = device.stream()
    _lib.check(lib.sfm_pnp_score_count(device.ptr(d_P), device.ptr(d_X), device.ptr(d_x), n, h,
                                       float(threshold), device.ptr(d_counts), st), "sfm_pnp_score_count")
    _lib.check(lib.sfm_argmax_first(device.ptr(d_counts), h, 0, device.ptr(d_best), None, st),
               "sfm_argmax_first")
    return d_counts, d_best


def pnp_hypotheses(X, x, K, samples):
    """Poses of every sampled 4-tuple; a failing LAPACK call marks only that hypothesis invalid
    (the reference's per-iteration try/except, PnPRANSAC.py:53,77-79)."""
    with np.errstate(all="ignore"):
        try:
            C_, R = linear_pnp_batch(X[samples], x[samples], K)
        except np.linalg.LinAlgError:
            H = len(samples)
            C_, R = np.full((H, 3), np.nan), np.full((H, 3, 3), np.nan)
            for i, s in enumerate(samples):
                try:
                    C_[i], R[i] = LinearPnP(X[s], x[s], K)
                except Exception:
                    pass
        t = -np.einsum("hij,hj->hi", R, C_)
        P = np.einsum("ij,hjk->hik", K, np.concatenate([R, t[:, :, None]], axis=2))   # :60
    return C_, R, np.ascontiguousarray(P.reshape(len(samples), 12))


def PnPRANSAC(X, x, K, threshold=200, n_max=1000, info=None):
    """Drop-in for PnPRANSAC.py:6-89 -> (C, R)."""
    X = np.array(X)
    x = np.array(x)
    K = np.array(K)
    n_points = len(X)
    if n_points < 4:                                                 # :38-39
        raise ValueError("At least 4 point correspondences are required for PnP")
    best_C = best_R = None
    best_count = 0
    if n_max > 0:
        samples = sampler.sample_stream(n_points, 4, int(n_max))     # one draw per iteration (:49)
        Xf, xf = X.astype(np.float64), x.astype(np.float64)
        C_, R, P = pnp_hypotheses(Xf, xf, K.astype(np.float64), samples)
        d_counts, d_best = pnp_score_device(device.to_device(P), device.to_device(Xf),
                                            device.to_device(xf), threshold)
        best = d_best.cpu().numpy()
        best_count = int(best[1])
        if best_count > 0:
            best_C, best_R = C_[best[0]].copy(), R[best[0]].copy()
        if info is not None:
            info.update(best_idx=int(best[0]), best_count=best_count, counts=d_counts.cpu().numpy(), P=P)
    if best_C is None or best_count < 4:                             # :82-87
        try:
            best_C, best_R = LinearPnP(X, x, K)
            print("Warning: PnP RANSAC failed, using linear PnP on all points")
        except Exception as e:
            raise ValueError(f"PnP estimation failed: {e}")
    return best_C, best_R
