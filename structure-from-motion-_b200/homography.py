"""Host side of the homography RANSAC path: mirrors GetHomographyInliers.py (names,
argument meaning, return order, error behaviour) over h_ransac.cu / f_npt.cu."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, device, sampler
from .ransac import FRansacDevice, _alloc_result


def h_ransac_device(d_x1x2: torch.Tensor, d_samples: torch.Tensor, threshold: float, want_err: bool = False) -> FRansacDevice:
    """K1..K4 for homographies on the current stream. d_x1x2 [n,4] f64, d_samples [h,4] i32 (device).
    The result object is the F-RANSAC one: .F holds the [h,9] hypotheses, .F_best the winner."""
    n, h = int(d_x1x2.shape[0]), int(d_samples.shape[0])
    r = _alloc_result(n, h, want_err)
    rc = _lib.load().sfm_h_ransac(device.ptr(d_x1x2), n, device.ptr(d_samples), h, float(threshold), device.ptr(r.F),
                                  device.ptr(r.counts), device.ptr(r.best), device.ptr(r.F_best), device.ptr(r.mask),
                                  device.ptr(r.err), device.stream())
    _lib.check(rc, "sfm_h_ransac")
    return r


def h_ransac_host(points1, points2, samples, threshold, want_err=False, want_all=False):
    """Host arrays in, host results out: dict(best_idx, best_count, H_best[3,3], mask[n] bool, and with
    want_all also H[h,9], counts[h]; with want_err err[n])."""
    x1x2 = np.ascontiguousarray(np.hstack([np.asarray(points1, np.float64), np.asarray(points2, np.float64)]))
    r = h_ransac_device(device.to_device(x1x2), device.to_device(samples, np.int32), threshold, want_err)
    packed = r.packed.cpu().numpy()
    out = dict(H_best=packed[:72].view(np.float64).reshape(3, 3).copy(),
               best_idx=int(packed[72:76].view(np.int32)[0]), best_count=int(packed[76:80].view(np.int32)[0]),
               mask=packed[80:].astype(bool))
    if want_all:
        out["H"] = r.F.cpu().numpy()
        out["counts"] = r.counts.cpu().numpy()
    if want_err:
        out["err"] = r.err.cpu().numpy()
    return out


def find_homography(image1_coords, image2_coords):
    """Drop-in for GetHomographyInliers.py:4-82 -> H (3 x 3), H[2,2] = 1."""
    image1_coords = np.array(image1_coords)
    image2_coords = np.array(image2_coords)
    n_points = len(image1_coords)
    if n_points < 4:
        raise ValueError("At least 4 point correspondences are required for homography estimation")
    lib = _lib.load()
    d_H = device.empty((9,), torch.float64)
    if n_points == 4:
        x = np.ascontiguousarray(np.hstack([image1_coords.reshape(4, 2), image2_coords.reshape(4, 2)]), np.float64)
        d_s = device.to_device(np.arange(4, dtype=np.int32).reshape(1, 4), np.int32)
        _lib.check(lib.sfm_h4pt_batch(device.ptr(device.to_device(x)), 4, device.ptr(d_s), 1, device.ptr(d_H),
                                      device.stream()), "sfm_h4pt_batch")
    else:
        d_x1 = device.to_device(image1_coords.reshape(n_points, 2))
        d_x2 = device.to_device(image2_coords.reshape(n_points, 2))
        nbytes = int(lib.sfm_homography_npt_workspace_bytes(n_points))
        ws = device.empty((nbytes,), torch.uint8)
        _lib.check(lib.sfm_homography_npt(device.ptr(d_x1), device.ptr(d_x2), n_points, device.ptr(d_H), device.ptr(ws),
                                          nbytes, device.stream()), "sfm_homography_npt")
    return d_H.cpu().numpy().reshape(3, 3)


def get_homography_inliers(image1_coords_org, image2_coords_org, idx, threshold=30, n_max=1000):
    """Drop-in for GetHomographyInliers.py:85-165 -> (H_best, inlier_idx). Consumes the global `random`
    stream exactly like the reference: one random.sample(range(N), min(4, N)) per iteration (:126)."""
    image1_coords_org = np.array(image1_coords_org)
    image2_coords_org = np.array(image2_coords_org)
    idx = np.array(idx)
    n_points = len(image1_coords_org)
    if n_points < 4:                                   # :108-110 (no RNG draw)
        return None, np.array([])
    samples = sampler.sample_stream(n_points, 4, int(n_max))
    if n_max <= 0:
        return None, np.array([])
    if image1_coords_org.ndim != 2 or image1_coords_org.shape[1] != 2 or image2_coords_org.shape != image1_coords_org.shape:
        return None, np.array([])                      # every iteration raises inside the try (:131,:159-161)
    res = h_ransac_host(image1_coords_org, image2_coords_org, samples, threshold)
    if res["best_count"] == 0:                         # :164-165
        return None, np.array([])
    return res["H_best"], idx[np.where(res["mask"])[0]]
