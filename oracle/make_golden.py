"""Generate tests/golden/*.npz by running the UNMODIFIED reference (/root/reference/Phase 1)
in the build container. TEST INFRASTRUCTURE. Run:  python oracle/make_golden.py

Each file stores inputs, the seed, the reference outputs and the library versions that the
reference's arithmetic actually lives in (numpy/LAPACK, scipy/MINPACK). The reference does not
exist on the GPU box; the committed .npz files are what pins the oracle and the CUDA path there.
"""
from __future__ import annotations

import io
import os
import random
import sys
import types
from contextlib import redirect_stdout

import numpy as np
import scipy

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("SFM_REF_DIR", "/root/reference/Phase 1")
OUT = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)  # reference modules import each other by bare name

from oracle import scenes  # noqa: E402  (data generators only)

# matplotlib is absent here; Utils.py imports it at module level (Utils.py:4-5). Stub it.
for name in ("matplotlib", "matplotlib.pyplot", "mpl_toolkits", "mpl_toolkits.mplot3d"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["mpl_toolkits.mplot3d"].Axes3D = object
sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]

import BundleAdjustment as refBA  # noqa: E402
import EstimateFundamentalMatrix as refEFM  # noqa: E402
import GetHomographyInliers as refH  # noqa: E402
import GetInliersRANSAC as refRANSAC  # noqa: E402
import LinearPnP as refLPnP  # noqa: E402
import LinearTriangulation as refLT  # noqa: E402
import NonLinearTriangulation as refNLT  # noqa: E402
import NonlinearPnP as refNLP  # noqa: E402
import PnPRANSAC as refPR  # noqa: E402
import Utils as refUtils  # noqa: E402
import EssentialMatrixFromFundamentalMatrix as refE  # noqa: E402
import ExtractCameraPose as refECP  # noqa: E402
import DisambiguateCameraPose as refDCP  # noqa: E402

VERS = dict(numpy=np.__version__, scipy=scipy.__version__, python=sys.version.split()[0])


def save(name, **kw):
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name), versions=str(VERS), **kw)
    print("wrote", name, {k: getattr(v, "shape", None) for k, v in kw.items()})


def replay_samples(seed, n, k, count):
    random.seed(seed)
    return np.array([random.sample(range(n), min(k, n)) for _ in range(count)], dtype=np.int32)


def gen_f8pt():
    sc = scenes.two_view(n=600, seed=11)
    rng = np.random.default_rng(5)
    idx = np.array([rng.choice(600, 8, replace=False) for _ in range(256)], dtype=np.int32)
    with np.errstate(all="ignore"):
        F = np.array([refEFM.EstimateFundamentalMatrix(sc["x1"][s], sc["x2"][s]).ravel() for s in idx])
    # all-point calls (N > 8), inliers only and with outliers
    good = ~sc["is_outlier"]
    F_all_good = refEFM.EstimateFundamentalMatrix(sc["x1"][good], sc["x2"][good])
    F_all = refEFM.EstimateFundamentalMatrix(sc["x1"], sc["x2"])
    F_20 = refEFM.EstimateFundamentalMatrix(sc["x1"][good][:20], sc["x2"][good][:20])
    save("f8pt.npz", x1=sc["x1"], x2=sc["x2"], good=good, idx=idx, F=F, F_all_good=F_all_good,
         F_all=F_all, F_20=F_20)


def gen_fransac():
    for tag, n, thr, seed, n_max in (("a", 558, 0.06, 3, 400), ("b", 558, 1.0, 4, 400),
                                     ("c", 2000, 0.5, 5, 300), ("d", 90, 1.0, 6, 200)):
        sc = scenes.two_view(n=n, seed=100 + seed)
        index = np.arange(1000, 1000 + n) * 3  # arbitrary, non-trivial index array
        samples = replay_samples(seed, n, 8, n_max)
        random.seed(seed)
        with np.errstate(all="ignore"):
            inl, outl, F = refRANSAC.GetInliersRANSAC(sc["x1"], sc["x2"], index, thr, n_max)
        random.seed(seed)
        with np.errstate(all="ignore"):
            F2, inl_idx = refRANSAC.get_inliers_ransac(sc["x1"], sc["x2"], index, thr, n_max)
        assert F is not None and np.array_equal(F, F2)
        save(f"fransac_{tag}.npz", x1=sc["x1"], x2=sc["x2"], index=index, thr=thr, seed=seed,
             n_max=n_max, samples=samples, inlier_index=inl, outlier_indices=outl, F=F,
             alias_inlier_idx=inl_idx)


def gen_p3data():
    data = os.path.join(REF, "P3Data") + "/"
    with redirect_stdout(io.StringIO()):
        fx, fy, flag = refUtils.get_data(data, 5)[:3]
    out = {}
    for (a, b), seed in (((1, 2), 0), ((2, 3), 1), ((1, 4), 2)):
        idx = np.where(flag[:, a - 1] & flag[:, b - 1])[0]
        c1 = np.hstack((fx[idx, a - 1].reshape(-1, 1), fy[idx, a - 1].reshape(-1, 1)))
        c2 = np.hstack((fx[idx, b - 1].reshape(-1, 1), fy[idx, b - 1].reshape(-1, 1)))
        random.seed(seed)
        with redirect_stdout(io.StringIO()):
            H, h_idx = refH.get_homography_inliers(c1, c2, idx, threshold=30, n_max=1000)
        p1 = np.hstack((fx[h_idx, a - 1].reshape(-1, 1), fy[h_idx, a - 1].reshape(-1, 1)))
        p2 = np.hstack((fx[h_idx, b - 1].reshape(-1, 1), fy[h_idx, b - 1].reshape(-1, 1)))
        state_seed = 1000 + seed
        samples = replay_samples(state_seed, len(p1), 8, 1000)
        random.seed(state_seed)
        with np.errstate(all="ignore"):
            F, f_idx = refRANSAC.get_inliers_ransac(p1, p2, h_idx, threshold=0.06, n_max=1000)
        k = f"{a}{b}"
        out.update({f"p1_{k}": p1, f"p2_{k}": p2, f"hidx_{k}": h_idx, f"seed_{k}": state_seed,
                    f"samples_{k}": samples, f"F_{k}": F, f"fidx_{k}": f_idx})
    save("p3data_fransac.npz", **out)


def gen_tri():
    sc = scenes.two_view(n=300, outlier_frac=0.0, seed=21)
    with redirect_stdout(io.StringIO()):
        x0 = refLT.LinearTriangulation(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"], sc["x1"], sc["x2"])
        X = refNLT.NonLinearTriangulation(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"],
                                          sc["x1"], sc["x2"], x0)
    # a harder start: perturbed initial points
    rng = np.random.default_rng(3)
    x0b = x0 + rng.normal(0, 0.2, x0.shape)
    Xb = refNLT.NonLinearTriangulation(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"],
                                       sc["x1"], sc["x2"], x0b)
    P1 = sc["K"] @ np.hstack([sc["R1"], -sc["R1"] @ sc["C1"].reshape(3, 1)])
    P2 = sc["K"] @ np.hstack([sc["R2"], -sc["R2"] @ sc["C2"].reshape(3, 1)])
    loss0 = np.array([refNLT.Loss(x0[i], sc["x1"][i], sc["x2"][i], P1, P2) for i in range(len(x0))])
    save("nltri.npz", K=sc["K"], C1=sc["C1"], R1=sc["R1"], C2=sc["C2"], R2=sc["R2"], x1=sc["x1"],
         x2=sc["x2"], x0=x0, X=X, x0b=x0b, Xb=Xb, loss0=loss0)


def gen_twoview():
    """Two-view pose front end: E from F, the four (C, R) configurations, their DLT
    triangulations and the cheirality pick, on two seeded scenes (clean and with outliers)."""
    out = {}
    for tag, sc in (("a", scenes.two_view(n=400, outlier_frac=0.0, seed=31)),
                    ("b", scenes.two_view(n=600, outlier_frac=0.2, seed=32))):
        with redirect_stdout(io.StringIO()), np.errstate(all="ignore"):
            F = refEFM.EstimateFundamentalMatrix(sc["x1"], sc["x2"])
            E = refE.EssentialMatrixFromFundamentalMatrix(F, sc["K"])
            Cset, Rset = refECP.ExtractCameraPose(E)
            Xset = [refLT.LinearTriangulation(sc["K"], np.zeros(3), np.eye(3), c, r, sc["x1"], sc["x2"])
                    for c, r in zip(Cset, Rset)]
            C, R, X = refDCP.DisambiguateCameraPose(Cset, Rset, Xset)
        out.update({f"K_{tag}": sc["K"], f"x1_{tag}": sc["x1"], f"x2_{tag}": sc["x2"], f"F_{tag}": F, f"E_{tag}": E,
                    f"Cset_{tag}": Cset, f"Rset_{tag}": Rset, f"Xset_{tag}": np.array(Xset), f"C_{tag}": C,
                    f"R_{tag}": R, f"X_{tag}": X})
    save("twoview.npz", **out)


def gen_homography():
    """Homography RANSAC: find_homography on 4 and on many points, and
    get_homography_inliers on seeded synthetic pairs with the recorded random.sample stream."""
    out = {}
    rng = np.random.default_rng(41)
    for tag, n, out_frac, thr, n_max in (("a", 500, 0.3, 30.0, 400), ("b", 120, 0.5, 5.0, 300)):
        # planar scene: x2 = H_true x1 + noise, a fraction replaced by gross outliers
        Ht = np.array([[0.9, 0.05, 30.0], [-0.04, 1.1, -12.0], [1e-5, -2e-5, 1.0]])
        x1 = rng.uniform(0, 1000, (n, 2))
        t = (Ht @ np.hstack([x1, np.ones((n, 1))]).T).T
        x2 = t[:, :2] / t[:, 2:3] + rng.normal(0, 0.7, (n, 2))
        bad = rng.random(n) < out_frac
        x2[bad] = rng.uniform(0, 1000, (int(bad.sum()), 2))
        idx = np.arange(1000, 1000 + n)
        seed = 500 + n
        samples = replay_samples(seed, n, 4, n_max)
        random.seed(seed)
        with np.errstate(all="ignore"):
            H, inl = refH.get_homography_inliers(x1, x2, idx, threshold=thr, n_max=n_max)
        Hall = refH.find_homography(x1[~bad], x2[~bad])
        H4 = refH.find_homography(x1[:4], x2[:4])
        out.update({f"x1_{tag}": x1, f"x2_{tag}": x2, f"idx_{tag}": idx, f"thr_{tag}": thr, f"seed_{tag}": seed,
                    f"samples_{tag}": samples, f"H_{tag}": H, f"inl_{tag}": inl, f"good_{tag}": ~bad,
                    f"Hall_{tag}": Hall, f"H4_{tag}": H4})
    save("homography.npz", **out)


def gen_pnp():
    out = {}
    for tag, n, ofrac, seed in (("a", 300, 0.0, 31), ("b", 2000, 0.10, 32), ("c", 12, 0.0, 33)):
        sc = scenes.pnp_scene(n=n, outlier_frac=ofrac, seed=seed)
        rng = np.random.default_rng(seed)
        from scipy.spatial.transform import Rotation
        R0 = Rotation.from_rotvec(rng.normal(0, 0.03, 3)).as_matrix() @ sc["R"]
        C0 = sc["C"] + rng.normal(0, 0.1, 3)
        with redirect_stdout(io.StringIO()):
            Cn, Rn = refNLP.NonLinearPnP(sc["X"], sc["x"], sc["K"], C0, R0)
            Ca, Ra = refNLP.nonlinear_PnP(sc["K"], C0, R0, sc["x"], sc["X"])
        assert np.array_equal(Cn, Ca) and np.array_equal(Rn, Ra)
        p0 = np.hstack([Rotation.from_matrix(R0).as_rotvec(), -R0 @ C0])
        loss0 = refNLP.NonLinearPnPLoss(p0, sc["X"], sc["x"], sc["K"])
        Cl, Rl = refLPnP.LinearPnP(sc["X"], sc["x"], sc["K"])
        random.seed(seed)
        samples = replay_samples(seed, n, 4, 200)
        random.seed(seed)
        with redirect_stdout(io.StringIO()), np.errstate(all="ignore"):
            Cr, Rr = refPR.PnPRANSAC(sc["X"], sc["x"], sc["K"], threshold=200, n_max=200)
        out.update({f"X_{tag}": sc["X"], f"x_{tag}": sc["x"], f"C0_{tag}": C0, f"R0_{tag}": R0,
                    f"C_{tag}": Cn, f"R_{tag}": Rn, f"loss0_{tag}": loss0, f"Clin_{tag}": Cl,
                    f"Rlin_{tag}": Rl, f"seed_{tag}": seed, f"samples_{tag}": samples,
                    f"Cransac_{tag}": Cr, f"Rransac_{tag}": Rr})
    save("pnp.npz", K=scenes.K_REF, **out)


def gen_linpnp():
    """The reference's LinearPnP (LinearPnP.py:3-96) on every 4-point sample the PnP goldens recorded, and on small
    and large all-point sets: pins the hypothesis generator of PnPRANSAC (numpy's Vt[-1] of the 8 x 12 system, the QR
    sign conventions) for the device kernel."""
    g = np.load(os.path.join(OUT, "pnp.npz"))
    out = {}
    with np.errstate(all="ignore"):
        for tag in ("a", "b", "c"):
            X, x, samples = g[f"X_{tag}"], g[f"x_{tag}"], g[f"samples_{tag}"]
            C4, R4 = np.full((len(samples), 3), np.nan), np.full((len(samples), 3, 3), np.nan)
            for i, smp in enumerate(samples):
                try:
                    C4[i], R4[i] = refLPnP.LinearPnP(X[smp], x[smp], scenes.K_REF)
                except Exception:
                    pass
            out[f"C4_{tag}"], out[f"R4_{tag}"] = C4, R4
        Xb, xb = g["X_b"], g["x_b"]
        for n in (5, 6, 7, 8, 12, 50, 2000):
            C_, R_ = refLPnP.LinearPnP(Xb[:n], xb[:n], scenes.K_REF)
            out[f"Cn_{n}"], out[f"Rn_{n}"] = C_, R_
    save("linpnp4.npz", K=scenes.K_REF, **out)


def gen_wrapper():
    """BASELINE config 1 as the reference runs it: the UNMODIFIED Wrapper_dev.sfm_wrapper on P3Data (drawing
    stubbed), seeds 0..5, every hot call recorded through the names the wrapper bound (tests/wrapper_harness.py).
    Per pair: homography + inlier index set, F + inlier index set; for pair 1_2: E, the four poses, the chosen pose,
    linear and refined points. PnP / BA are not reached on this data (14-18 points survive)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import wrapper_harness as wh
    assert wh.available()
    out = {}
    for seed in range(6):
        log, _ = wh.run(False, seed)
        pre = f"s{seed}_"
        if seed == 0:
            fx, fy, fl = log["get_data"][0][2]
            out.update(feature_x=fx, feature_y=fy, feature_flag=fl)
        out[pre + "n_pairs"] = len(log["get_homography_inliers"])
        for k, c in enumerate(log["get_homography_inliers"]):
            H, idx = c[2]
            out[pre + f"H_{k}"] = np.full((3, 3), np.nan) if H is None else H
            out[pre + f"hidx_{k}"] = np.asarray(idx, np.int64)
        out[pre + "n_f"] = len(log["get_inliers_ransac"])
        for k, c in enumerate(log["get_inliers_ransac"]):
            F, idx = c[2]
            out[pre + f"F_{k}"] = np.full((3, 3), np.nan) if F is None else F
            out[pre + f"fidx_{k}"] = np.asarray(idx, np.int64)
        out[pre + "E"] = log["EssentialMatrixFromFundamentalMatrix"][0][2]
        Cset, Rset = log["ExtractCameraPose"][0][2]
        out[pre + "Cset"], out[pre + "Rset"] = np.array(Cset), np.array(Rset)
        Cc, Rc = log["get_correct_camera_pose"][0][2]
        out[pre + "C_corr"], out[pre + "R_corr"] = np.array(Cc), np.array(Rc)
        lt = log["linear_triangulation"]
        out[pre + "X_candidates"] = np.array([c[2] for c in lt[:4]])
        out[pre + "X_linear"] = lt[4][2]
        out[pre + "X_refined"] = log["nonlinear_triangulation"][0][2]
        out[pre + "calls"] = np.array([len(log[n]) for n in wh.HOT], np.int64)
        out[pre + "rng_after"] = log["rng_after"]
        print(seed, "F inliers per pair", [len(c[2][1]) for c in log["get_inliers_ransac"]], "points", len(out[pre + "X_refined"]))
    save("p3data_wrapper.npz", **out)


def dense_ba_inputs(sc, n_extra=7):
    """Dense (F x n_img) arrays in the shape perform_bundle_adjustment takes, with some
    untriangulated filler rows so that the valid-point remapping is exercised."""
    n_pts, n_cams = len(sc["pts0"]), len(sc["R0"])
    F = n_pts + n_extra
    rng = np.random.default_rng(1)
    rows = np.sort(rng.choice(F, n_pts, replace=False))
    world = rng.normal(0, 1, (F, 3))
    world[rows] = sc["pts0"]
    valid = np.zeros((F, 1), dtype=int)
    valid[rows] = 1
    fx, fy = np.zeros((F, n_cams)), np.zeros((F, n_cams))
    flags = np.zeros((F, n_cams), dtype=int)
    fx[rows[sc["pt_idx"]], sc["cam_idx"]] = sc["obs"][:, 0]
    fy[rows[sc["pt_idx"]], sc["cam_idx"]] = sc["obs"][:, 1]
    flags[rows[sc["pt_idx"]], sc["cam_idx"]] = 1
    other = np.setdiff1d(np.arange(F), rows)
    flags[other, 0] = 1  # visible but not triangulated: must be ignored
    return world, valid, fx, fy, flags


def gen_ba():
    for tag, nc, npts, k, seed in (("a", 4, 40, 3, 41), ("b", 5, 60, 4, 42)):
        sc = scenes.ba_scene(n_cams=nc, n_pts=npts, obs_per_pt=k, seed=seed)
        world, valid, fx, fy, flags = dense_ba_inputs(sc)
        R_set, C_set = [r for r in sc["R0"]], [c for c in sc["C0"]]
        buf = io.StringIO()
        with redirect_stdout(buf):
            R_opt, C_opt, world_opt = refBA.perform_bundle_adjustment(
                world, valid, fx, fy, flags, R_set, C_set, sc["K"], nc - 1)
        log = buf.getvalue()
        cost = float(log.split("Final cost:")[1].split()[0])
        # residual vector of the reference at the initial parameters
        from scipy.spatial.transform import Rotation
        x0 = []
        for R, C in zip(R_set, C_set):
            x0.extend(Rotation.from_matrix(R).as_rotvec())
            x0.extend(-R @ C)
        x0 = np.array(x0 + list(sc["pts0"].reshape(-1)))
        r0 = refBA.bundle_adjustment_residuals(x0, nc, npts, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
        xf = []
        for R, C in zip(R_opt, C_opt):
            xf.extend(Rotation.from_matrix(R).as_rotvec())
            xf.extend(-R @ C)
        xf = np.array(xf + list(world_opt[valid.flatten() == 1].reshape(-1)))
        rf = refBA.bundle_adjustment_residuals(xf, nc, npts, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
        save(f"ba_{tag}.npz", K=sc["K"], world=world, valid=valid, fx=fx, fy=fy, flags=flags,
             R0=sc["R0"], C0=sc["C0"], cam_idx=sc["cam_idx"], pt_idx=sc["pt_idx"], obs=sc["obs"],
             pts0=sc["pts0"], R_opt=np.array(R_opt), C_opt=np.array(C_opt), world_opt=world_opt,
             printed_cost=cost, x0=x0, r0=r0, rf=rf, log=log)


if __name__ == "__main__":
    which = sys.argv[1:] or ["f8pt", "fransac", "p3data", "tri", "pnp", "linpnp", "ba", "twoview", "homography", "wrapper"]
    for w in which:
        globals()["gen_" + w]()
