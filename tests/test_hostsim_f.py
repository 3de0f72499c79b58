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
