"""BAL text I/O (CPU) and the device-side assembly of the BA observation list against the
reference's loop (frozen in tests/golden/ba_*.npz through oracle/make_golden.py)."""
import numpy as np
import pytest

from conftest import golden


def _scene(nc=6, npts=40, seed=3):
    rng = np.random.default_rng(seed)
    cams = np.hstack([rng.normal(0, 0.05, (nc, 3)), rng.normal(0, 0.3, (nc, 3)) + [0, 0, 6.0]])
    pts = rng.uniform(-1, 1, (npts, 3))
    vis = rng.random((npts, nc)) < 0.6
    vis[:, 0] = True
    pt_idx, cam_idx = np.nonzero(vis)
    obs = rng.uniform(-300, 300, (len(pt_idx), 2))
    return cams, pts, cam_idx.astype(np.int32), pt_idx.astype(np.int32), obs


def test_bal_round_trip(tmp_path):
    from sfm_b200 import bal
    cams, pts, cam_idx, pt_idx, obs = _scene()
    K = np.array([[-800.0, 0, 0], [0, -800.0, 0], [0, 0, 1]])
    # shuffle the observation order: BAL files are not necessarily point-sorted
    perm = np.random.default_rng(0).permutation(len(obs))
    d = bal.from_scene(cams, pts, cam_idx[perm], pt_idx[perm], obs[perm], K)
    path = str(tmp_path / "problem.txt")
    bal.write_bal(path, d["cam_idx"], d["pt_idx"], d["obs"], d["cams9"], d["pts"])
    r = bal.read_bal(path)
    assert r["n_cams"] == 6 and r["n_pts"] == 40
    for k in ("cam_idx", "pt_idx", "obs", "cams9", "pts"):
        assert np.array_equal(r[k], d[k]), k                 # repr() round-trips doubles exactly
    prob = bal.to_problem(r)
    assert np.all(np.diff(prob["pt_idx"]) >= 0)
    assert np.array_equal(prob["K"], K)
    # the same multiset of observations
    a = np.lexsort((cam_idx, pt_idx))
    b = np.lexsort((prob["cam_idx"], prob["pt_idx"]))
    assert np.array_equal(prob["obs"][b], obs[a]) and np.array_equal(prob["cam_idx"][b], cam_idx[a])
    assert np.array_equal(prob["cams"], cams) and np.array_equal(prob["pts"], pts)


def test_bal_rejects_what_the_model_cannot_hold(tmp_path):
    from sfm_b200 import bal
    cams, pts, cam_idx, pt_idx, obs = _scene()
    d = bal.from_scene(cams, pts, cam_idx, pt_idx, obs, np.diag([-500.0, -500.0, 1.0]))
    d["cams9"][2, 7] = 1e-3                                   # radial distortion on one camera
    with pytest.raises(ValueError):
        bal.to_problem(d)
    assert bal.to_problem(d, require_pinhole=False)["K"][0, 0] == -500.0
    bad = str(tmp_path / "bad.txt")
    with open(bad, "w") as f:
        f.write("1 1 1\n0 0 1.0 2.0\n")
    with pytest.raises(ValueError):
        bal.read_bal(bad)
    with pytest.raises(ValueError):
        bal.from_scene(cams, pts, cam_idx, pt_idx, obs, np.array([[500.0, 0, 320], [0, 500, 240], [0, 0, 1]]))


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["a", "b"])
def test_gpu_assembly_matches_reference_order(cuda_device, tag):
    from sfm_b200 import bundle
    g = golden(f"ba_{tag}.npz")
    n_cams = len(g["R0"])
    host = bundle.build_problem(g["world"], g["valid"], g["fx"], g["fy"], g["flags"], list(g["R0"]), list(g["C0"]))
    dev = bundle.build_problem_device(g["world"], g["valid"], g["fx"], g["fy"], g["flags"], n_cams)
    assert dev["n_points"] == host["n_points"] and dev["n_obs"] == len(host["cam_idx"])
    assert np.array_equal(dev["cam_idx"].cpu().numpy(), host["cam_idx"])
    assert np.array_equal(dev["pt_idx"].cpu().numpy(), host["pt_idx"])
    assert np.array_equal(dev["obs"].cpu().numpy(), host["obs"])
    assert np.array_equal(dev["pts"].cpu().numpy(), host["pts"])
    assert np.array_equal(dev["valid_rows"].cpu().numpy(), host["valid"])
    # the reference's own lists (BundleAdjustment.py:164-176, frozen by oracle/make_golden.py)
    assert np.array_equal(dev["cam_idx"].cpu().numpy(), g["cam_idx"])
    assert np.array_equal(dev["pt_idx"].cpu().numpy(), g["pt_idx"])
    assert np.array_equal(dev["obs"].cpu().numpy(), g["obs"])


@pytest.mark.gpu
def test_gpu_assembly_edge_cases_and_scale(cuda_device):
    from sfm_b200 import bundle
    rng = np.random.default_rng(5)
    F, n_img = 200_000, 12
    filtered = (rng.random((F, 1)) < 0.4).astype(int)
    flags = (rng.random((F, n_img)) < 0.3).astype(int)
    flags[rng.random(F) < 0.1] = 0                           # valid rows without any observation
    fx, fy = rng.uniform(0, 1000, (F, n_img)), rng.uniform(0, 800, (F, n_img))
    world = rng.normal(size=(F, 3))
    Rs, Cs = [np.eye(3)] * 9, [np.zeros(3)] * 9              # only the first 9 images are cameras (:156)
    host = bundle.build_problem(world, filtered, fx, fy, flags, Rs, Cs)
    dev = bundle.build_problem_device(world, filtered, fx, fy, flags, 9)
    assert dev["n_obs"] == len(host["cam_idx"]) and dev["n_points"] == host["n_points"]
    assert np.array_equal(dev["cam_idx"].cpu().numpy(), host["cam_idx"])
    assert np.array_equal(dev["pt_idx"].cpu().numpy(), host["pt_idx"])
    assert np.array_equal(dev["obs"].cpu().numpy(), host["obs"])
    assert np.array_equal(dev["pts"].cpu().numpy(), host["pts"])
    # nothing valid / nothing observed -> None (BundleAdjustment.py:152-153, :171-172)
    assert bundle.build_problem_device(world, np.zeros((F, 1), int), fx, fy, flags, 9) is None
    assert bundle.build_problem_device(world, filtered, fx, fy, np.zeros_like(flags), 9) is None
    # flags other than exactly 1 are not observations (== 1 in the reference)
    f2 = flags * 2
    assert bundle.build_problem_device(world, filtered, fx, fy, f2, 9) is None
