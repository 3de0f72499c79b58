"""GPU: the threaded pinned staging of large host arrays (sfm_b200.device.upload_via_pinned) and the one-buffer
parameter round trip of bundle.ba_solve_sparse above its threshold."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_upload_via_pinned_small_and_sliced(cuda_device):
    import torch
    from sfm_b200 import device
    rng = np.random.default_rng(1)
    for n in (1000, 700_001):                                   # below / above PAR_MIN_BYTES (rows of 2 doubles)
        src = rng.normal(size=(n, 2))
        pinned = torch.empty((n, 2), dtype=torch.float64, pin_memory=True)
        dst = device.empty((n, 2), torch.float64)
        device.upload_via_pinned(dst, src, pinned)
        assert np.array_equal(dst.cpu().numpy(), src)
        src2 = np.asfortranarray(rng.normal(size=(n, 2)))      # a strided source goes through the same copy
        device.upload_via_pinned(dst, src2, pinned)
        assert np.array_equal(dst.cpu().numpy(), src2)


def test_ba_host_api_with_a_point_set_above_the_staging_threshold(cuda_device):
    """200k points (4.8 MB of parameters): the sliced upload path of ba_solve_sparse against the device-resident
    solve of the same problem."""
    from scipy.spatial.transform import Rotation
    from oracle import scenes
    from sfm_b200 import bundle, device
    sc = scenes.ba_scene(n_cams=20, n_pts=200_000, obs_per_pt=3, seed=5)
    cams0 = np.hstack([Rotation.from_matrix(sc["R0"]).as_rotvec(), -np.einsum("nij,nj->ni", sc["R0"], sc["C0"])])
    pts0 = np.ascontiguousarray(sc["pts0"])
    assert pts0.nbytes >= device.PAR_MIN_BYTES
    cams, pts, info = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"], max_nfev=4)
    cams2, pts2, info2 = bundle.ba_solve_sparse(cams0, pts0, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"], max_nfev=4)
    prob = bundle.BAProblem(20, 200_000, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    d_c, d_p = device.to_device(cams0), device.to_device(pts0)
    ref = prob.solve_device(d_c, d_p, 4)
    prob.close()
    bundle.clear_problem_cache()
    for c, p, i in ((cams, pts, info), (cams2, pts2, info2)):
        assert abs(i["cost"] - ref["cost"]) <= 1e-9 * ref["cost"]
        assert np.allclose(c, d_c.cpu().numpy(), rtol=0, atol=1e-9) and np.allclose(p, d_p.cpu().numpy(), rtol=0, atol=1e-8)
    assert cams0 is not cams and np.array_equal(pts0, sc["pts0"])          # inputs untouched, outputs are copies
