"""Two-view pose front end: host mirror of EssentialMatrixFromFundamentalMatrix.py,
ExtractCameraPose.py and DisambiguateCameraPose.py over twoview.cu, plus the fused device path that
keeps a pair in HBM between F-RANSAC and triangulation (Wrapper_dev.py:156-176).

Order of the four configurations: the reference's list order is LAPACK's sign choice for the singular
vectors of E; ours follows the convention in twoview.cu. The set of configurations, the cheirality
votes and the selected pose are the same (tests/test_gpu_twoview.py)."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, device
from .bundle import _hptr


def EssentialMatrixFromFundamentalMatrix(F, K):
    """EssentialMatrixFromFundamentalMatrix.py:3-19 (one 3x3 product on the host; the fused device
    path below forms E inside its set-up kernel)."""
    return K.T @ F @ K


class TwoViewDevice:
    """Device-resident outputs of sfm_two_view_pose."""
    __slots__ = ("cfg", "P", "Xall", "counts", "best", "pose")


def two_view_pose_device(d_F, K, d_x1, d_x2) -> TwoViewDevice:
    """d_F [9] or [3,3] (device), K host 3x3, d_x1/d_x2 [n,2] (device). No host synchronisation."""
    n = int(d_x1.shape[0])
    r = TwoViewDevice()
    r.cfg = device.empty((4, 12), torch.float64)
    r.P = device.empty((5, 12), torch.float64)
    r.Xall = device.empty((4, max(n, 1), 3), torch.float64)[:, :n]
    r.counts = device.empty((4,), torch.int32)
    r.best = device.empty((2,), torch.int32)
    r.pose = device.empty((12,), torch.float64)
    K = np.ascontiguousarray(K, np.float64).reshape(3, 3)
    rc = _lib.load().sfm_two_view_pose(device.ptr(d_F), _hptr(K), device.ptr(d_x1), device.ptr(d_x2), n,
                                       device.ptr(r.cfg), device.ptr(r.P), device.ptr(r.Xall), device.ptr(r.counts),
                                       device.ptr(r.best), device.ptr(r.pose), device.stream())
    _lib.check(rc, "sfm_two_view_pose")
    return r


def two_view_pose(F, K, x1, x2):
    """Host arrays in -> dict(Cset [4,3], Rset [4,3,3], Xset [4,n,3], counts [4], best, C, R, X):
    E = K^T F K, the four configurations, their DLT triangulations, the cheirality pick."""
    x1 = np.ascontiguousarray(x1, np.float64).reshape(-1, 2)
    x2 = np.ascontiguousarray(x2, np.float64).reshape(-1, 2)
    r = two_view_pose_device(device.to_device(np.ascontiguousarray(F, np.float64).reshape(9)), K,
                             device.to_device(x1), device.to_device(x2))
    cfg = r.cfg.cpu().numpy()
    best = r.best.cpu().numpy()
    Xset = r.Xall.cpu().numpy()
    i = int(best[0])
    return dict(Cset=cfg[:, :3].copy(), Rset=cfg[:, 3:].reshape(4, 3, 3).copy(), Xset=Xset,
                counts=r.counts.cpu().numpy(), best=i, best_count=int(best[1]),
                C=cfg[i, :3].copy(), R=cfg[i, 3:].reshape(3, 3).copy(), X=Xset[i].copy())


def ExtractCameraPose(E):
    """Drop-in for ExtractCameraPose.py:3-81 -> (Cset [4,3], Rset [4,3,3])."""
    E = np.ascontiguousarray(np.array(E), np.float64).reshape(9)
    lib = _lib.load()
    d_cfg = device.empty((4, 12), torch.float64)
    d_P = device.empty((5, 12), torch.float64)
    d_counts = device.empty((4,), torch.int32)
    eye = np.ascontiguousarray(np.eye(3))
    _lib.check(lib.sfm_two_view_configs(device.ptr(device.to_device(E)), _hptr(eye), device.ptr(d_cfg), device.ptr(d_P),
                                        device.ptr(d_counts), device.stream()), "sfm_two_view_configs")
    cfg = d_cfg.cpu().numpy()
    return cfg[:, :3].copy(), cfg[:, 3:].reshape(4, 3, 3).copy()


def DisambiguateCameraPose(Cset, Rset, Xset):
    """Drop-in for DisambiguateCameraPose.py:3-92 -> (best_C, best_R, best_X), same prints."""
    Cset = np.array(Cset)
    Rset = np.array(Rset)
    Xset = [np.array(X) for X in Xset] if isinstance(Xset, list) else [np.array(Xset[i]) for i in range(len(Xset))]
    n_configs = len(Cset)
    counts = np.zeros(n_configs, dtype=np.int64)
    lens = [len(X) for X in Xset]
    if n_configs == 4 and len(set(lens)) == 1 and lens[0] > 0:
        n = lens[0]
        cfg = np.hstack([Cset.reshape(4, 3), Rset.reshape(4, 9)]).astype(np.float64)
        d_counts = device.empty((4,), torch.int32)
        d_X = device.to_device(np.ascontiguousarray(np.stack([X.reshape(n, 3) for X in Xset]), np.float64))
        _lib.check(_lib.load().sfm_cheirality_count(device.ptr(device.to_device(cfg)), device.ptr(d_X), n,
                                                    device.ptr(d_counts), device.stream()), "sfm_cheirality_count")
        counts = d_counts.cpu().numpy().astype(np.int64)
    else:                                   # ragged or non-standard input: one configuration per call
        for i in range(n_configs):
            if lens[i] == 0:
                continue
            cfg = np.tile(np.hstack([Cset[i].reshape(3), Rset[i].reshape(9)]).astype(np.float64), (4, 1))
            d_counts = device.empty((4,), torch.int32)
            X4 = np.ascontiguousarray(np.broadcast_to(Xset[i].reshape(1, lens[i], 3), (4, lens[i], 3)), np.float64)
            _lib.check(_lib.load().sfm_cheirality_count(device.ptr(device.to_device(cfg)), device.ptr(device.to_device(X4)),
                                                        lens[i], device.ptr(d_counts), device.stream()),
                       "sfm_cheirality_count")
            counts[i] = int(d_counts.cpu().numpy()[0])
    best_count, best_idx, best_C, best_R, best_X = -1, 0, None, None, None
    for i in range(n_configs):
        if lens[i] == 0:
            continue
        if counts[i] > best_count:
            best_count, best_idx = int(counts[i]), i
            best_C, best_R, best_X = Cset[i].copy(), Rset[i].copy(), Xset[i].copy()
    if best_C is None:
        best_C, best_R, best_X = Cset[0], Rset[0], Xset[0]
        print("Warning: No valid camera pose found, using first configuration")
    else:
        print(f"Selected camera pose configuration {best_idx} with {best_count} points in front of both cameras")
    return best_C, best_R, best_X
