"""Bundle Adjustment in the Large (BAL) text problems <-> the sparse arrays of sfm_ba_solve
(SURVEY.md 8f N3). The reference has no on-disk format for problems of BASELINE config 5's size (only
its own matching*.txt, Utils.py:8-64); BAL is the public one.

File layout: `n_cams n_pts n_obs`, then n_obs lines `cam pt x y`, then 9 numbers per camera
(rotation vector 3, translation 3, focal f, radial k1, k2), then 3 per point. BAL's camera looks down
-z: p = -P / P.z, pixel = f * r(p) * p with r = 1 + k1 |p|^2 + k2 |p|^4. Without distortion and with
one shared focal length that is this repo's pinhole model with K = [[-f,0,0],[0,-f,0],[0,0,1]]."""
from __future__ import annotations

import numpy as np


def read_bal(path: str) -> dict:
    """-> dict(n_cams, n_pts, cam_idx[int32], pt_idx[int32], obs[n,2], cams9[n_cams,9], pts[n_pts,3])."""
    with open(path, "r") as f:
        head = f.readline().split()
        if len(head) != 3:
            raise ValueError("not a BAL file: first line must hold three integers")
        nc, npts, nobs = (int(v) for v in head)
        rest = np.array(f.read().split(), dtype=np.float64)
    need = 4 * nobs + 9 * nc + 3 * npts
    if rest.size != need:
        raise ValueError(f"BAL file holds {rest.size} numbers after the header, expected {need}")
    o = rest[:4 * nobs].reshape(nobs, 4)
    cam_idx, pt_idx = o[:, 0].astype(np.int32), o[:, 1].astype(np.int32)
    if np.any(o[:, 0] != cam_idx) or np.any(o[:, 1] != pt_idx) or cam_idx.min(initial=0) < 0 or pt_idx.min(initial=0) < 0 \
            or (nobs and (cam_idx.max() >= nc or pt_idx.max() >= npts)):
        raise ValueError("BAL file: observation indices out of range")
    cams9 = rest[4 * nobs:4 * nobs + 9 * nc].reshape(nc, 9).copy()
    pts = rest[4 * nobs + 9 * nc:].reshape(npts, 3).copy()
    return dict(n_cams=nc, n_pts=npts, cam_idx=cam_idx, pt_idx=pt_idx, obs=o[:, 2:].copy(), cams9=cams9, pts=pts)


def write_bal(path: str, cam_idx, pt_idx, obs, cams9, pts) -> None:
    cam_idx, pt_idx = np.asarray(cam_idx), np.asarray(pt_idx)
    obs, cams9, pts = np.asarray(obs, np.float64), np.asarray(cams9, np.float64), np.asarray(pts, np.float64)
    with open(path, "w") as f:
        f.write(f"{len(cams9)} {len(pts)} {len(cam_idx)}\n")
        for c, p, (x, y) in zip(cam_idx, pt_idx, obs):
            f.write(f"{int(c)} {int(p)} {x!r} {y!r}\n".replace("np.float64(", "").replace(")", ""))
        for v in cams9.reshape(-1):
            f.write(repr(float(v)) + "\n")
        for v in pts.reshape(-1):
            f.write(repr(float(v)) + "\n")


def to_problem(bal: dict, require_pinhole: bool = True) -> dict:
    """Point-sorted (stable; cameras keep their file order within a point) sparse problem for
    bundle.ba_solve_sparse / BAProblem: dict(cams[n,6], pts, cam_idx, pt_idx, obs, K, order).
    Raises when the cameras do not share one focal length or carry distortion (this repo's model is the
    reference's: one K, no distortion), unless require_pinhole is False (then k1, k2 are dropped and the
    median focal length is used)."""
    cams9 = bal["cams9"]
    f, k = cams9[:, 6], cams9[:, 7:9]
    if require_pinhole and (np.ptp(f) > 1e-12 * max(1.0, abs(float(f[0]))) or np.any(k != 0.0)):
        raise ValueError("BAL cameras have individual focal lengths or radial distortion; the reference's model "
                         "is a shared K without distortion")
    fm = float(np.median(f))
    order = np.argsort(bal["pt_idx"], kind="stable")
    K = np.array([[-fm, 0.0, 0.0], [0.0, -fm, 0.0], [0.0, 0.0, 1.0]])
    return dict(cams=cams9[:, :6].copy(), pts=bal["pts"].copy(), cam_idx=bal["cam_idx"][order].copy(),
                pt_idx=bal["pt_idx"][order].copy(), obs=bal["obs"][order].copy(), K=K, order=order)


def from_scene(cams, pts, cam_idx, pt_idx, obs, K) -> dict:
    """Inverse of to_problem for a shared pinhole K = diag(fx, fx, 1) without skew or principal point:
    re-expresses it in BAL's -z convention (f = |fx|, observations and camera axes flipped as needed)."""
    K = np.asarray(K, np.float64)
    if abs(K[0, 0] - K[1, 1]) > 1e-12 * abs(K[0, 0]) or K[0, 1] != 0 or K[0, 2] != 0 or K[1, 2] != 0:
        raise ValueError("BAL holds one focal length per camera and no principal point")
    f = -K[0, 0]                       # K = diag(-f, -f, 1)
    if f <= 0:
        raise ValueError("expected K = diag(-f, -f, 1) with f > 0 (BAL's -z camera)")
    cams9 = np.hstack([np.asarray(cams, np.float64).reshape(-1, 6), np.full((len(cams), 1), f), np.zeros((len(cams), 2))])
    return dict(n_cams=len(cams9), n_pts=len(pts), cam_idx=np.asarray(cam_idx, np.int32), pt_idx=np.asarray(pt_idx, np.int32),
                obs=np.asarray(obs, np.float64), cams9=cams9, pts=np.asarray(pts, np.float64))
