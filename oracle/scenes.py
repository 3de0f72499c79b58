"""TEST AND BENCH INFRASTRUCTURE (lives under oracle/ so that no product module and not the reference arm of
bench.py has to import the product package for its inputs).

Synthetic scenes of the shapes BASELINE.json names. Data generation
only (numpy); shared by tests/ and bench.py so that both sides of every comparison see the
same seeded inputs."""
from __future__ import annotations

import numpy as np
from scipy.spatial.transform import Rotation

# Wrapper_dev.py:143
K_REF = np.array([[531.122155322710, 0.0, 407.192550839899],
                  [0.0, 531.541737503901, 313.308715048366],
                  [0.0, 0.0, 1.0]])


def project(K, C, R, X, eps=1e-8):
    P = K @ np.hstack([R, -R @ np.asarray(C, float).reshape(3, 1)])
    xh = (P @ np.hstack([X, np.ones((len(X), 1))]).T).T
    return xh[:, :2] / (xh[:, 2:3] + eps)


def two_view(n=50_000, outlier_frac=0.30, noise=0.5, seed=20260102):
    """C2: two-view correspondences with gross outliers in view 2."""
    rng = np.random.default_rng(seed)
    X = np.stack([rng.uniform(-3, 3, n), rng.uniform(-2.2, 2.2, n), rng.uniform(4, 8, n)], axis=1)
    R1, C1 = np.eye(3), np.zeros(3)
    R2 = Rotation.from_rotvec([0.05, -0.10, 0.02]).as_matrix()
    C2 = np.array([0.5, 0.05, 0.1])
    x1 = project(K_REF, C1, R1, X) + rng.normal(0, noise, (n, 2))
    x2 = project(K_REF, C2, R2, X) + rng.normal(0, noise, (n, 2))
    n_out = int(round(outlier_frac * n))
    out_idx = rng.choice(n, n_out, replace=False)
    x2[out_idx] = np.stack([rng.uniform(0, 800, n_out), rng.uniform(0, 600, n_out)], axis=1)
    is_outlier = np.zeros(n, bool)
    is_outlier[out_idx] = True
    # C-contiguous, as np.array(list of matches) in the reference's wrapper gives (project() returns a transposed view)
    return dict(K=K_REF.copy(), X=X, x1=np.ascontiguousarray(x1), x2=np.ascontiguousarray(x2), R1=R1, C1=C1, R2=R2, C2=C2,
                is_outlier=is_outlier)


def look_at(C, target):
    """World-to-camera rotation with +z towards target and +y roughly world-down-free."""
    z = np.asarray(target, float) - np.asarray(C, float)
    z /= np.linalg.norm(z)
    up = np.array([0.0, 1.0, 0.0])
    x = np.cross(up, z)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    return np.stack([x, y, z], axis=0)


def ba_scene(n_cams=50, n_pts=20_000, obs_per_pt=10, noise=0.5, seed=20260103, layout="arc",
             window=None, perturb=True):
    """C3 (layout='arc': 120-degree arc of radius 6 around a [-2,2]^3 box at z=6, camera 0 =
    (I,0)) and C5 (layout='ring': radius 8 looking inward; each point seen by `obs_per_pt`
    cameras drawn from a window of `window` consecutive cameras). Observations are sorted
    by point (the reference's order, BundleAdjustment.py:164-169), cameras ascending within a
    point. Returns truth, perturbed initial values and the sparse observation list."""
    rng = np.random.default_rng(seed)
    centre = np.array([0.0, 0.0, 6.0])
    R_true, C_true = [], []
    if layout == "arc":
        pts = rng.uniform(-2, 2, (n_pts, 3)) + centre
        for i in range(n_cams):
            if i == 0:
                C, R = np.zeros(3), np.eye(3)
            else:
                a = np.deg2rad(-60.0 + 120.0 * i / max(n_cams - 1, 1)) - np.deg2rad(-60.0)
                # rotate the camera position about the box centre (y axis), starting at origin
                C = centre + 6.0 * np.array([-np.sin(a), 0.0, -np.cos(a)])
                R = look_at(C, centre)
            R_true.append(R)
            C_true.append(C)
    else:
        pts = rng.uniform(-2, 2, (n_pts, 3))
        for i in range(n_cams):
            a = 2 * np.pi * i / n_cams
            C = 8.0 * np.array([np.sin(a), 0.0, -np.cos(a)])
            R_true.append(look_at(C, np.zeros(3)))
            C_true.append(C)
    R_true, C_true = np.array(R_true), np.array(C_true)
    # visibility: obs_per_pt distinct cameras per point
    if window is None:
        keys = rng.random((n_pts, n_cams))
        cams = np.sort(np.argpartition(keys, obs_per_pt - 1, axis=1)[:, :obs_per_pt], axis=1)
    else:
        start = rng.integers(0, n_cams, n_pts)
        keys = rng.random((n_pts, window))
        off = np.argpartition(keys, obs_per_pt - 1, axis=1)[:, :obs_per_pt]
        cams = np.sort((start[:, None] + off) % n_cams, axis=1)
    cam_idx = cams.reshape(-1).astype(np.int32)
    pt_idx = np.repeat(np.arange(n_pts, dtype=np.int32), obs_per_pt)
    # project every observation with its camera
    Xc = np.einsum("oij,oj->oi", R_true[cam_idx], pts[pt_idx] - C_true[cam_idx])
    xh = Xc @ K_REF.T
    obs = xh[:, :2] / (xh[:, 2:3] + 1e-8) + rng.normal(0, noise, (len(cam_idx), 2))
    if perturb:
        pts0 = pts + rng.normal(0, 0.02, pts.shape)
        C0 = C_true + rng.normal(0, 0.02, C_true.shape)
        dR = Rotation.from_rotvec(rng.normal(0, 0.005, (n_cams, 3))).as_matrix()
        R0 = np.einsum("nij,njk->nik", dR, R_true)
        if layout == "arc":                    # cam 0 unperturbed (but not fixed)
            C0[0], R0[0] = C_true[0], R_true[0]
    else:
        pts0, C0, R0 = pts.copy(), C_true.copy(), R_true.copy()
    return dict(K=K_REF.copy(), pts_true=pts, R_true=R_true, C_true=C_true, pts0=pts0, R0=R0, C0=C0,
                cam_idx=cam_idx, pt_idx=pt_idx, obs=obs)


def pnp_scene(n=2000, noise=0.5, outlier_frac=0.10, seed=20260105):
    """One C4 registration: 3D-2D correspondences for a random pose near identity."""
    rng = np.random.default_rng(seed)
    X = np.stack([rng.uniform(-3, 3, n), rng.uniform(-2.2, 2.2, n), rng.uniform(4, 8, n)], axis=1)
    R = Rotation.from_rotvec(rng.normal(0, 0.2, 3)).as_matrix()
    C = rng.normal(0, 0.5, 3)
    x = project(K_REF, C, R, X) + rng.normal(0, noise, (n, 2))
    n_out = int(round(outlier_frac * n))
    if n_out:
        oi = rng.choice(n, n_out, replace=False)
        x[oi] = np.stack([rng.uniform(0, 800, n_out), rng.uniform(0, 600, n_out)], axis=1)
    return dict(K=K_REF.copy(), X=X, x=x, R=R, C=C)
