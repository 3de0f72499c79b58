"""Host side of PnP RANSAC: mirrors PnPRANSAC.py:6-89 and LinearPnP.py:3-96 (names, argument meaning, return order,
error behaviour) and drives the device kernels through the C ABI: K6a LinearPnP hypotheses (csrc/linpnp.cu), K6
consensus scoring, K3 first-maximum argmax. Nothing here calls LAPACK: the reference's library conventions that its
4-point hypotheses depend on (numpy's Vt[-1] of a rank-deficient DLT system, the Householder signs of qr) are replayed
on the device -- see csrc/linpnp_math.cuh for what is pinned and for the one branch that cannot be (flag bit 0)."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, device, sampler

FLAG_REFLECTED = 1      # LinearPnP.py:88-90 branch: the reference's output there is decided by rounding noise
FLAG_NONFINITE = 2


def _hk(K):
    K = np.ascontiguousarray(K, np.float64).reshape(3, 3)
    return K, np.ascontiguousarray(np.linalg.inv(K))        # LinearPnP.py:38 normalises with numpy's inverse


def linear_pnp4_device(d_X, d_x, d_samples, K, d_offsets=None, n_regs=1, want_pose=True, mute_reflected=True):
    """K6a on device tensors: d_samples [n_regs * h, 4] i32 -> (d_C [H,3], d_R [H,9], d_P [H,12], d_flag [H] u8).
    mute_reflected (the drop-in's policy): hypotheses of the noise-decided branch get a NaN row in d_P and cannot win
    the consensus, as in the reference they all but never do (include/sfm_b200.h)."""
    lib = _lib.load()
    K, Kinv = _hk(K)
    H = int(d_samples.shape[0])
    d_P = device.empty((H, 12), torch.float64)
    d_C = device.empty((H, 3), torch.float64) if want_pose else None
    d_R = device.empty((H, 9), torch.float64) if want_pose else None
    d_flag = device.empty((H,), torch.uint8)
    n = int(d_X.shape[0]) if d_offsets is None else 0
    _lib.check(lib.sfm_linear_pnp4_batch(device.ptr(d_X), device.ptr(d_x), device.ptr(d_offsets), int(n_regs), n,
                                         device.ptr(d_samples), H // int(n_regs), K.ctypes.data, Kinv.ctypes.data,
                                         1 if mute_reflected else 0, device.ptr(d_C), device.ptr(d_R), device.ptr(d_P), device.ptr(d_flag),
                                         device.stream()), "sfm_linear_pnp4_batch")
    return d_C, d_R, d_P, d_flag


def linear_pnp_device(d_X, d_x, K):
    """LinearPnP on ALL correspondences of one registration (device tensors) -> (d_CR [12], d_flag [1])."""
    lib = _lib.load()
    K, Kinv = _hk(K)
    n = int(d_X.shape[0])
    nbytes = int(lib.sfm_linear_pnp_npt_workspace_bytes(n))
    ws = device.empty((nbytes,), torch.uint8)
    d_CR = device.empty((12,), torch.float64)
    d_flag = device.empty((1,), torch.uint8)
    _lib.check(lib.sfm_linear_pnp_npt(device.ptr(d_X), device.ptr(d_x), n, K.ctypes.data, Kinv.ctypes.data,
                                      device.ptr(d_CR), None, device.ptr(d_flag), device.ptr(ws), nbytes,
                                      device.stream()), "sfm_linear_pnp_npt")
    return d_CR, d_flag


def LinearPnP(Xset, xset, K):
    """Drop-in for LinearPnP.py:3-96 -> (C (3,), R (3,3))."""
    Xset, xset, K = np.array(Xset), np.array(xset), np.array(K)
    if len(Xset) < 4:                                                # :31-32
        raise ValueError("At least 4 point correspondences are required for PnP")
    d_CR, _ = linear_pnp_device(device.to_device(Xset.astype(np.float64)), device.to_device(xset.astype(np.float64)), K)
    cr = d_CR.cpu().numpy()
    return cr[:3].copy(), cr[3:].reshape(3, 3).copy()


def pnp_score_device(d_P, d_X, d_x, threshold, d_counts=None, d_best=None):
    """K6 + K3 on device tensors -> (d_counts [h] i32, d_best [2] i32)."""
    lib = _lib.load()
    h, n = int(d_P.shape[0]), int(d_X.shape[0])
    d_counts = device.empty((h,), torch.int32) if d_counts is None else d_counts
    d_best = device.empty((2,), torch.int32) if d_best is None else d_best
    st = device.stream()
    _lib.check(lib.sfm_pnp_score_count(device.ptr(d_P), device.ptr(d_X), device.ptr(d_x), n, h,
                                       float(threshold), device.ptr(d_counts), st), "sfm_pnp_score_count")
    _lib.check(lib.sfm_argmax_first(device.ptr(d_counts), h, 0, device.ptr(d_best), None, st),
               "sfm_argmax_first")
    return d_counts, d_best


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
    d_X, d_x = device.to_device(X.astype(np.float64)), device.to_device(x.astype(np.float64))
    if n_max > 0:
        samples = sampler.sample_stream(n_points, 4, int(n_max))     # one draw per iteration (:49)
        d_C, d_R, d_P, d_flag = linear_pnp4_device(d_X, d_x, device.to_device(samples, np.int32), K)
        d_counts, d_best = pnp_score_device(d_P, d_X, d_x, threshold)
        w = d_best[:1].long()
        # one D2H: (winner index, count) | C | R of the winner
        packed = torch.cat([d_best.double(), d_C.index_select(0, w).reshape(-1), d_R.index_select(0, w).reshape(-1)])
        packed = packed.cpu().numpy()
        best_count = int(packed[1])
        if best_count > 0:                                           # strict '>' from zero (:72): count 0 never wins
            best_C, best_R = packed[2:5].copy(), packed[5:14].reshape(3, 3).copy()
        if info is not None:
            info.update(best_idx=int(packed[0]), best_count=best_count, counts=d_counts.cpu().numpy(),
                        P=d_P.cpu().numpy(), flags=d_flag.cpu().numpy(), samples=samples)
    if best_C is None or best_count < 4:                             # :82-87
        try:
            cr = linear_pnp_device(d_X, d_x, K)[0].cpu().numpy()
            best_C, best_R = cr[:3].copy(), cr[3:].reshape(3, 3).copy()
            print("Warning: PnP RANSAC failed, using linear PnP on all points")
        except Exception as e:
            raise ValueError(f"PnP estimation failed: {e}")
    return best_C, best_R


def pnp_ransac_batch(X_list, x_list, K, threshold=200, n_max=1000):
    """Many registrations at once (BASELINE config 4: 500 x 2,000 correspondences x 1,000 hypotheses): the sample
    streams are drawn registration after registration from the global `random` stream exactly as consecutive
    PnPRANSAC calls would (one replayer: the generator state is read and written back once), ONE launch each for the
    hypotheses, the scoring and the argmax of all registrations, one D2H brings all winners back. Returns (C [B,3], R [B,3,3], counts [B]). Registrations whose best
    count is below 4 fall back to LinearPnP on all their points (PnPRANSAC.py:82-87)."""
    B = len(X_list)
    sizes = [len(X) for X in X_list]
    if min(sizes) < 4:
        raise ValueError("At least 4 point correspondences are required for PnP")
    offs = np.zeros(B + 1, np.int32)
    offs[1:] = np.cumsum(sizes)
    samples = np.empty((B * n_max, 4), np.int32)
    rep = sampler.StreamReplayer(min(sizes), 4)           # ONE read and ONE write-back of the generator state
    try:
        for b in range(B):
            rep.draw(int(n_max), samples[b * n_max:(b + 1) * n_max], n=sizes[b])
    finally:
        rep.finish()
    d_X = device.to_device(np.concatenate([np.asarray(X, np.float64) for X in X_list]))
    d_x = device.to_device(np.concatenate([np.asarray(x, np.float64) for x in x_list]))
    d_off = device.to_device(offs, np.int32)
    d_C, d_R, d_P, _ = linear_pnp4_device(d_X, d_x, device.to_device(samples, np.int32), K, d_off, B)
    d_counts = device.empty((B * n_max,), torch.int32)
    d_best = device.empty((B, 2), torch.int32)
    lib, st = _lib.load(), device.stream()
    for b0 in range(0, B, 65535):                                  # gridDim.z
        nb = min(65535, B - b0)
        _lib.check(lib.sfm_pnp_score_count_batch(device.ptr(d_P) + 96 * b0 * n_max, device.ptr(d_X), device.ptr(d_x),
                                                 device.ptr(d_off) + 4 * b0, nb, int(max(sizes[b0:b0 + nb])), int(n_max),
                                                 float(threshold), device.ptr(d_counts) + 4 * b0 * n_max, st),
                   "sfm_pnp_score_count_batch")
    _lib.check(lib.sfm_argmax_first_batch(device.ptr(d_counts), B, int(n_max), device.ptr(d_best), st),
               "sfm_argmax_first_batch")
    w = d_best[:, 0].long() + torch.arange(B, device=d_best.device) * n_max
    packed = torch.cat([d_best[:, 1:2].double(), d_C.index_select(0, w), d_R.index_select(0, w)], dim=1).cpu().numpy()
    counts = packed[:, 0].astype(np.int64)
    C_, R_ = packed[:, 1:4].copy(), packed[:, 4:13].reshape(B, 3, 3).copy()
    for b in np.nonzero(counts < 4)[0]:
        cr = linear_pnp_device(d_X[offs[b]:offs[b + 1]], d_x[offs[b]:offs[b + 1]], K)[0].cpu().numpy()
        C_[b], R_[b] = cr[:3], cr[3:].reshape(3, 3)
    return C_, R_, counts
