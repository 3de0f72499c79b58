"""CPU-only (X1): the reference's own driver binds the drop-in. The UNMODIFIED Wrapper_dev.py (staged under oracle/_ref by
oracle/stage_ref.py) is imported with structure-from-motion-_b200/dropin ahead of it on sys.path -- the whole integration
(Wrapper_dev.py:5-14 binds by module name) -- and every hot name it holds must BE the product's function, while everything
else (Utils I/O, drawing) stays reference code. Also pins tests/golden/p3data_wrapper.npz: the unmodified wrapper, run here
as shipped, reproduces it."""
import importlib

import numpy as np
import pytest

import wrapper_harness as wh
from conftest import golden

pytestmark = pytest.mark.skipif(not wh.available(), reason="oracle/_ref not staged (needs /root/reference at build time)")


def test_unmodified_wrapper_binds_every_hot_name_to_the_dropin():
    W = wh.import_wrapper(dropin=True)
    for name in wh.HOT:
        home = importlib.import_module(wh.HOME[name])
        assert getattr(W, name) is getattr(home, name), name
    ref_file = W.get_data.__code__.co_filename
    assert ref_file.endswith("oracle/_ref/phase1/Utils.py")                 # I/O stays reference code
    assert W.sfm_wrapper.__code__.co_filename.endswith("oracle/_ref/phase1/Wrapper_dev.py")
    # and without the drop-in directory the same import binds the reference's own modules
    W0 = wh.import_wrapper(dropin=False)
    for name in wh.HOT:
        assert "oracle/_ref/phase1" in getattr(W0, name).__code__.co_filename, name


def test_reference_wrapper_reproduces_the_committed_goldens():
    g = golden("p3data_wrapper.npz")
    log, _ = wh.run(dropin=False, seed=3)
    pre = "s3_"
    fx, fy, fl = log["get_data"][0][2]
    assert np.array_equal(fx, g["feature_x"]) and np.array_equal(fl, g["feature_flag"])
    assert len(log["get_inliers_ransac"]) == int(g[pre + "n_f"]) == 10
    for k, c in enumerate(log["get_inliers_ransac"]):
        assert np.array_equal(np.asarray(c[2][1]), g[pre + f"fidx_{k}"])
        assert np.allclose(c[2][0], g[pre + f"F_{k}"], rtol=1e-9, atol=0)
    assert np.allclose(log["nonlinear_triangulation"][0][2], g[pre + "X_refined"], rtol=1e-9)
    assert log["rng_after"] == float(g[pre + "rng_after"])
    assert [len(log[n]) for n in wh.HOT] == g[pre + "calls"].tolist()
    assert len(log["PnPRANSAC"]) == 0 and len(log["perform_bundle_adjustment"]) == 0     # not reached on P3Data
