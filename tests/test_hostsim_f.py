"""Kernel arithmetic (host build of csrc/f_math.cuh) against the oracle and the reference goldens."""
import ctypes as C

import numpy as np

from conftest import golden, rel_err
from oracle import sfm_oracle as so
import hostsim


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def test_f8pt_device_math_matches_reference():
    hs = hostsim.load()
    g = golden("f8pt.npz")
    worst = 0.0
    for s, Fref in zip(g["idx"], g["F"]):
        rows = np.ascontiguousarray(np.hstack([g["x1"][s], g["x2"][s]]))
        F = np.zeros(9)
        hs.hs_f8pt(_p(rows), _p(F))
        worst = max(worst, rel_err(F, Fref))
    assert worst < 1e-6, worst      # north_star tolerance
    assert worst < 1e-9, worst      # what the QR null vector actually delivers


def test_inlier_test_matches_reference_formula():
    hs = hostsim.load()
    g = golden("fransac_b.npz")
    rows = np.ascontiguousarray(np.hstack([g["x1"], g["x2"]]))
    n = len(rows)
    d = so.f_ransac_detailed(g["x1"], g["x2"], float(g["thr"]), g["samples"][:50])
    for F in d["F"]:
        if not np.all(np.isfinite(F)):
            continue
        mask = np.zeros(n, np.uint8)
        err = np.zeros(n)
        hs.hs_f_score(_p(rows), n, _p(np.ascontiguousarray(F)), C.c_double(float(g["thr"])), _p(mask), _p(err))
        ref = so.f_errors(F.reshape(3, 3), g["x1"], g["x2"])
        assert np.max(np.abs(err - ref) / np.maximum(ref, 1e-12)) < 1e-9
        border = np.abs(ref - float(g["thr"])) < 1e-9
        assert np.array_equal(mask.astype(bool)[~border], (ref < float(g["thr"]))[~border])


# ---- certified FP32 filter of the scoring kernel ------------------------------------------------
def _two_view(rng, n, noise, outliers=0.3, scale=1.0):
    """Synthetic two-view correspondences in pixels (f = 800, 1280 x 960), some gross outliers."""
    K = np.array([[800.0, 0, 640], [0, 800.0, 480], [0, 0, 1]])
    X = np.column_stack([rng.uniform(-4, 4, n), rng.uniform(-3, 3, n), rng.uniform(6, 14, n)])
    ang = 0.15
    R = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    t = np.array([-1.0, 0.05, 0.1])
    a = (K @ X.T).T
    b = (K @ (R @ X.T + t[:, None])).T
    x1 = a[:, :2] / a[:, 2:] + rng.normal(0, noise, (n, 2))
    x2 = b[:, :2] / b[:, 2:] + rng.normal(0, noise, (n, 2))
    bad = rng.random(n) < outliers
    x2[bad] = np.column_stack([rng.uniform(0, 1280, bad.sum()), rng.uniform(0, 960, bad.sum())])
    return np.ascontiguousarray(np.hstack([x1, x2]) * scale)


def _hypotheses(hs, rows, rng, h):
    F = np.zeros((h, 9))
    for k in range(h):
        s = rng.choice(len(rows), 8, replace=False)
        hs.hs_f8pt(_p(np.ascontiguousarray(rows[s])), _p(F[k]))
    return F


def _cert(hs, rows, F, thr, chunk=1024):
    out = np.zeros((len(F), 4), np.int64)
    hs.hs_f_cert(_p(rows), len(rows), _p(np.ascontiguousarray(F)), len(F), C.c_double(thr), chunk, _p(out))
    return out


def test_certified_fp32_filter_agrees_with_fp64_on_two_view_data():
    hs = hostsim.load()
    rng = np.random.default_rng(11)
    for noise, thr in ((0.3, 0.06), (0.05, 0.06), (0.5, 0.5), (1.0, 3.0)):
        rows = _two_view(rng, 1500, noise)
        F = _hypotheses(hs, rows, rng, 200)
        out = _cert(hs, rows, F, thr)
        assert out[:, 1].sum() == 0, (noise, thr, out[:, 1].sum())
        assert np.array_equal(out[:, 2], out[:, 3])
        # and it is a filter worth having: nearly every test is decided in FP32
        assert out[:, 0].sum() > 0.99 * len(rows) * len(F), (noise, thr, out[:, 0].sum())


def test_certified_fp32_filter_on_tests_placed_at_the_threshold():
    """Correspondences moved along the epipolar normal until the error sits at thr (1 +- 10^-k):
    the filter must either decide like FP64 or decline."""
    hs = hostsim.load()
    rng = np.random.default_rng(12)
    base = _two_view(rng, 600, 0.0, outliers=0.0)
    F = _hypotheses(hs, base, rng, 40)
    thr = 0.06
    total_wrong = 0
    declined = 0
    for Fk in F:
        Fm = Fk.reshape(3, 3)
        rows = base.copy()
        x1h = np.column_stack([rows[:, :2], np.ones(len(rows))])
        l2 = x1h @ Fm.T                     # epipolar lines in image 2 (GetInliersRANSAC.py:72)
        nrm = np.linalg.norm(l2[:, :2], axis=1)
        d = (np.sum(l2[:, :2] * rows[:, 2:], axis=1) + l2[:, 2]) / nrm
        rows[:, 2:] -= (d / nrm)[:, None] * l2[:, :2]          # onto the line
        k = rng.integers(1, 13, len(rows))
        sign = rng.choice([-1.0, 1.0], len(rows))
        target = thr * (1.0 + sign * 10.0 ** (-k.astype(float)))
        # error = |e|/2 (1/n1 + 1/n2): move along the line normal by roughly the target distance
        rows[:, 2:] += (target / nrm)[:, None] * l2[:, :2]
        out = _cert(hs, np.ascontiguousarray(rows), Fk[None], thr)
        total_wrong += out[0, 1]
        declined += len(rows) - out[0, 0]
        assert out[0, 2] == out[0, 3]
    assert total_wrong == 0
    assert declined > 0                    # the construction does reach the undecidable band


def test_certified_fp32_filter_scales_and_degenerate_inputs():
    hs = hostsim.load()
    rng = np.random.default_rng(13)
    rows = _two_view(rng, 800, 0.2)
    F = _hypotheses(hs, rows, rng, 30)
    cases = []
    for fs in (1e-25, 1e-8, 1.0, 1e8, 1e25):          # unnormalised hypotheses
        cases.append((rows, F * fs, 0.06))
    for cs in (1e-3, 1e3, 1e7):                        # normalised / huge coordinates
        r2 = np.ascontiguousarray(rows * cs)
        cases.append((r2, _hypotheses(hs, r2, rng, 30), 0.06 * cs))
    Fbad = F.copy()
    Fbad[0, 3] = np.nan
    Fbad[1, 8] = np.inf
    Fbad[2] = 0.0
    cases.append((rows, Fbad, 0.06))
    cases.append((rows, F, 0.0))
    cases.append((rows, F, -1.0))
    rbad = rows.copy()
    rbad[5, 0] = np.nan
    rbad[6, 3] = 1e300
    cases.append((rbad, F, 0.5))
    for r, f, thr in cases:
        for chunk in (128, 1024):
            out = _cert(hs, r, f, thr, chunk)
            assert out[:, 1].sum() == 0, (thr, chunk, out[:, 1])
            assert np.array_equal(out[:, 2], out[:, 3])


def _reject(hs, rows, F, thr, chunk=1024):
    out = np.zeros((len(F), 3), np.int64)
    hs.hs_f_cert_reject(_p(rows), len(rows), _p(np.ascontiguousarray(F)), len(F), C.c_double(thr), chunk, _p(out))
    return out


def test_early_reject_never_rejects_an_inlier_and_catches_the_gross_outliers():
    """K2's sqrt-free early reject (f_math.cuh f_cert_reject): |e| >= T min(N1, N2) certifies an outlier. It must
    never fire on a test the FP64 decision calls an inlier, and it should fire on nearly every gross outlier."""
    hs = hostsim.load()
    rng = np.random.default_rng(21)
    for noise, thr in ((0.3, 0.06), (0.05, 0.06), (0.5, 0.5), (1.0, 3.0)):
        rows = _two_view(rng, 1500, noise)
        F = _hypotheses(hs, rows, rng, 200)
        out = _reject(hs, rows, F, thr)
        assert out[:, 1].sum() == 0, (noise, thr)
        assert out[:, 0].sum() > 0.9 * out[:, 2].sum(), (noise, thr, out[:, 0].sum(), out[:, 2].sum())


def test_early_reject_on_tests_swept_across_its_boundary():
    """Correspondences moved off the epipolar line by thr x 2^[-1, 3]: the sweep crosses both the inlier threshold
    and the reject boundary (about twice the threshold when the two line norms are alike)."""
    hs = hostsim.load()
    rng = np.random.default_rng(22)
    base = _two_view(rng, 600, 0.0, outliers=0.0)
    F = _hypotheses(hs, base, rng, 40)
    thr = 0.06
    rejected = kept_outliers = 0
    for Fk in F:
        Fm = Fk.reshape(3, 3)
        rows = base.copy()
        x1h = np.column_stack([rows[:, :2], np.ones(len(rows))])
        l2 = x1h @ Fm.T
        nrm = np.linalg.norm(l2[:, :2], axis=1)
        d = (np.sum(l2[:, :2] * rows[:, 2:], axis=1) + l2[:, 2]) / nrm
        rows[:, 2:] -= (d / nrm)[:, None] * l2[:, :2]
        target = thr * 2.0 ** rng.uniform(-1.0, 3.0, len(rows)) * rng.choice([-1.0, 1.0], len(rows))
        rows[:, 2:] += (target / nrm)[:, None] * l2[:, :2]
        out = _reject(hs, np.ascontiguousarray(rows), Fk[None], thr)
        assert out[0, 1] == 0
        rejected += out[0, 0]
        kept_outliers += out[0, 2] - out[0, 0]
    assert rejected > 0 and kept_outliers > 0          # the sweep really straddles the boundary


def test_early_reject_scales_and_degenerate_inputs():
    hs = hostsim.load()
    rng = np.random.default_rng(23)
    rows = _two_view(rng, 800, 0.2)
    F = _hypotheses(hs, rows, rng, 30)
    cases = []
    for fs in (1e-25, 1e-8, 1.0, 1e8, 1e25):
        cases.append((rows, F * fs, 0.06))
    for cs in (1e-3, 1e3, 1e7):
        r2 = np.ascontiguousarray(rows * cs)
        cases.append((r2, _hypotheses(hs, r2, rng, 30), 0.06 * cs))
    Fbad = F.copy()
    Fbad[0, 3] = np.nan
    Fbad[1, 8] = np.inf
    Fbad[2] = 0.0
    cases.append((rows, Fbad, 0.06))
    cases.append((rows, F, 0.0))                        # threshold 0: nothing is an inlier
    cases.append((rows, F, 1e6))                        # everything is
    for r, f, thr in cases:
        with np.errstate(all="ignore"):
            out = _reject(hs, r, f, thr)
        assert out[:, 1].sum() == 0


def test_first_stage_of_the_early_reject_implies_the_full_certificate():
    """K2 votes first on f_cert_reject_stage1 (one squared norm, no F^T x2). Whatever it rejects the full certificate
    must reject too -- across noise levels, thresholds, scales of F and of the coordinates, and broken F -- and it
    should catch most of what the full certificate catches."""
    hs = hostsim.load()
    rng = np.random.default_rng(24)

    def stage1(rows, F, thr, chunk=1024):
        out = np.zeros((len(F), 2), np.int64)
        hs.hs_f_cert_stage1(_p(rows), len(rows), _p(np.ascontiguousarray(F)), len(F), C.c_double(thr), chunk, _p(out))
        return out
    rows = _two_view(rng, 1500, 0.3)
    F = _hypotheses(hs, rows, rng, 200)
    for thr in (0.06, 0.5, 3.0):
        s1, full = stage1(rows, F, thr), _reject(hs, rows, F, thr)
        assert s1[:, 1].sum() == 0
        assert s1[:, 0].sum() <= full[:, 0].sum()
        assert s1[:, 0].sum() > 0.95 * full[:, 0].sum()
    Fbad = F[:8].copy()
    Fbad[0, 3], Fbad[1, 8], Fbad[2] = np.nan, np.inf, 0.0
    for r, f, thr in ((rows, F * 1e-25, 0.06), (rows, F * 1e25, 0.06), (np.ascontiguousarray(rows * 1e7), F, 0.06 * 1e7),
                      (rows, Fbad, 0.06), (rows, F, 0.0), (rows, F, 1e6)):
        with np.errstate(all="ignore"):
            assert stage1(r, f, thr)[:, 1].sum() == 0

