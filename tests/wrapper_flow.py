"""The control flow of the reference's sfm_wrapper (Wrapper_dev.py:68-312) on in-memory feature
arrays, with every numerical function taken from a namespace `f`, so the SAME flow can be driven by
the oracle (CPU restatement of the reference) and by the GPU drop-in modules. Drawing, file output and
printing are left out; everything that touches numbers is kept, including the order of the
random.sample draws (homography RANSAC then F-RANSAC per pair, PnPRANSAC per registered image)."""
from itertools import combinations

import numpy as np


def run(f, feature_x, feature_y, feature_flag, K, n_max=300, f_threshold=0.5):
    n_img = feature_x.shape[1]
    filtered = np.zeros_like(feature_flag)
    F_12 = None
    for a, b in combinations(range(1, n_img + 1), 2):                       # :68-126
        _idx = np.where(feature_flag[:, a - 1] & feature_flag[:, b - 1])
        c1 = np.hstack((feature_x[_idx, a - 1].reshape((-1, 1)), feature_y[_idx, a - 1].reshape((-1, 1))))
        c2 = np.hstack((feature_x[_idx, b - 1].reshape((-1, 1)), feature_y[_idx, b - 1].reshape((-1, 1))))
        idx = np.array(_idx).reshape(-1)
        H, h_idx = f.get_homography_inliers(c1, c2, idx, threshold=30, n_max=n_max)
        if H is None or len(h_idx) == 0:
            continue
        p1 = np.hstack((feature_x[h_idx, a - 1].reshape((-1, 1)), feature_y[h_idx, a - 1].reshape((-1, 1))))
        p2 = np.hstack((feature_x[h_idx, b - 1].reshape((-1, 1)), feature_y[h_idx, b - 1].reshape((-1, 1))))
        F, f_idx = f.get_inliers_ransac(p1, p2, h_idx, threshold=f_threshold, n_max=n_max)   # wrapper: 0.06
        if F is None or len(f_idx) == 0:
            continue
        if (a, b) == (1, 2):
            F_12 = F
        filtered[f_idx, a - 1] = 1
        filtered[f_idx, b - 1] = 1
    assert F_12 is not None
    _idx = np.where(filtered[:, 0] & filtered[:, 1])                        # :141-143
    x1 = np.hstack((feature_x[_idx, 0].reshape((-1, 1)), feature_y[_idx, 0].reshape((-1, 1))))
    x2 = np.hstack((feature_x[_idx, 1].reshape((-1, 1)), feature_y[_idx, 1].reshape((-1, 1))))
    E = f.EssentialMatrixFromFundamentalMatrix(F_12, K)                     # :156-176
    Cset, Rset = f.ExtractCameraPose(E)
    C0, R0 = np.zeros(3), np.eye(3)
    Xset = [f.linear_triangulation(K, C0, R0, c, r, x1, x2) for c, r in zip(Cset, Rset)]
    C_corr, R_corr, _ = f.DisambiguateCameraPose(Cset, Rset, Xset)
    X_lin = f.linear_triangulation(K, C0, R0, C_corr, R_corr, x1, x2)
    X_ref = f.nonlinear_triangulation(K, C0, R0, C_corr, R_corr, x1, x2, X_lin)       # :187
    n_feat = feature_x.shape[0]
    world = np.zeros((n_feat, 3))
    have = np.zeros((n_feat, 1), dtype=int)
    world[_idx] = X_ref
    have[_idx] = 1
    have[np.where(world[:, 2] < 0)[0]] = 0                                  # :210
    poses = {1: [C0, R0], 2: [C_corr, R_corr]}
    C_set, R_set = [C0, C_corr], [R0, R_corr]
    log = dict(F_12=F_12, C2=C_corr, R2=R_corr, X12=X_ref.copy(), n12=len(X_ref), ba_costs=[])
    for img in range(3, n_img + 1):                                         # :220-300
        fi = np.where(have[:, 0] & filtered[:, img - 1])
        if len(fi[0]) < 8:
            continue
        in1 = np.hstack((feature_x[fi, 0].reshape(-1, 1), feature_y[fi, 0].reshape(-1, 1)))
        ini = np.hstack((feature_x[fi, img - 1].reshape(-1, 1), feature_y[fi, img - 1].reshape(-1, 1)))
        wc = world[fi, :].reshape(-1, 3)
        C_new, R_new = f.PnPRANSAC(wc, ini, K, threshold=200, n_max=n_max)
        C_new, R_new = f.nonlinear_PnP(K, C_new, R_new, ini, wc)            # :250 (the call that is kept)
        poses[img] = [C_new, R_new]
        R_set.append(R_new)
        C_set.append(C_new)
        for other in range(1, img):
            wi = np.squeeze(np.asarray(np.where(filtered[:, other - 1] & filtered[:, img - 1])))
            if wi.ndim == 0 or len(wi) < 8:
                continue
            C_o, R_o = poses[other]
            a1 = np.hstack((feature_x[wi, other - 1].reshape(-1, 1), feature_y[wi, other - 1].reshape(-1, 1)))
            a2 = np.hstack((feature_x[wi, img - 1].reshape(-1, 1), feature_y[wi, img - 1].reshape(-1, 1)))
            X0 = f.linear_triangulation(K, C_o, R_o, C_new, R_new, a1, a2)
            Xr = f.nonlinear_triangulation(K, C_o, R_o, C_new, R_new, a1, a2, X0)
            world[wi] = Xr
            have[wi] = 1
            R_set, C_set, world = f.perform_bundle_adjustment(world, have, feature_x, feature_y, filtered,
                                                              R_set, C_set, K, img - 1)
            log["ba_costs"].append(f.ba_cost(world, have, feature_x, feature_y, filtered, R_set, C_set, K))
    log.update(R_set=[np.array(r) for r in R_set], C_set=[np.array(c) for c in C_set], world=world, have=have,
               filtered=filtered)
    return log


def synthetic_features(n_img=4, n_pts=90, seed=5, noise=0.05, outlier_frac=0.08):
    """Dense (n_features x n_images) arrays like Utils.extract_features builds (Utils.py:8-64): every
    point seen in every image (the wrapper needs >= 8 common points per registration), some matches
    replaced by gross outliers."""
    from scipy.spatial.transform import Rotation
    rng = np.random.default_rng(seed)
    K = np.array([[531.122155322710, 0, 407.192550839899], [0, 531.541737503901, 313.308715048366], [0, 0, 1]])
    X = np.stack([rng.uniform(-2.5, 2.5, n_pts), rng.uniform(-2, 2, n_pts), rng.uniform(5, 9, n_pts)], axis=1)
    fx, fy = np.zeros((n_pts, n_img)), np.zeros((n_pts, n_img))
    flag = np.ones((n_pts, n_img), dtype=int)
    for i in range(n_img):
        R = Rotation.from_rotvec([0.02 * i, -0.06 * i, 0.01 * i]).as_matrix()
        C = np.array([0.45 * i, 0.05 * i, 0.1 * i])
        x = (K @ (R @ (X - C).T)).T
        x = x[:, :2] / x[:, 2:3] + rng.normal(0, noise, (n_pts, 2))
        bad = rng.random(n_pts) < (outlier_frac if i > 0 else 0.0)
        x[bad] = rng.uniform(50, 750, (int(bad.sum()), 2))
        fx[:, i], fy[:, i] = x[:, 0], x[:, 1]
    return fx, fy, flag, K
