"""Homography RANSAC: oracle against the frozen reference outputs, the device arithmetic
compiled for the host against the oracle, and -- marked gpu -- the kernels through the C ABI.
Tolerances as for F-RANSAC: same winner and counts, masks bit-exact outside a 1e-9 band around the
threshold, H within 1e-6 relative (asserted 1e-8; degenerate samples excepted and counted)."""
import ctypes as C
import random

import numpy as np
import pytest

import hostsim
from conftest import golden, rel_err
from oracle import sfm_oracle as so

BAND = 1e-9


@pytest.mark.parametrize("tag", ["a", "b"])
def test_oracle_matches_reference(tag):
    g = golden("homography.npz")
    x1, x2, thr = g[f"x1_{tag}"], g[f"x2_{tag}"], float(g[f"thr_{tag}"])
    r = so.h_ransac_detailed(x1, x2, thr, g[f"samples_{tag}"])
    assert np.array_equal(r["H"][r["best_idx"]].reshape(3, 3), g[f"H_{tag}"])
    assert np.array_equal(g[f"idx_{tag}"][np.where(r["mask"])[0]], g[f"inl_{tag}"])
    good = g[f"good_{tag}"]
    assert np.array_equal(so.find_homography(x1[good], x2[good]), g[f"Hall_{tag}"])
    assert np.array_equal(so.find_homography(x1[:4], x2[:4]), g[f"H4_{tag}"])
    # the recorded stream is what random.sample draws after seeding
    random.seed(int(g[f"seed_{tag}"]))
    assert np.array_equal(so.draw_samples(len(x1), 4, len(g[f"samples_{tag}"])), g[f"samples_{tag}"])


def _compare(got_H, got_counts, ref, n_hyp):
    ok = np.all(np.isfinite(ref["H"]), axis=1) & np.all(np.isfinite(got_H), axis=1)
    relH = np.linalg.norm(got_H[ok] - ref["H"][ok], axis=1) / np.linalg.norm(ref["H"][ok], axis=1)
    n_bad = int(np.sum(relH > 1e-6))
    mism = int(np.sum(got_counts[ok] != ref["counts"][ok]))
    assert n_bad <= max(1, n_hyp // 100), f"{n_bad} hypotheses differ by > 1e-6 in H"
    assert mism <= max(2, n_hyp // 50), f"{mism} count mismatches"
    return relH


@pytest.mark.parametrize("tag", ["a", "b"])
def test_hostsim_matches_oracle(tag):
    g = golden("homography.npz")
    x1, x2, thr, samples = g[f"x1_{tag}"], g[f"x2_{tag}"], float(g[f"thr_{tag}"]), g[f"samples_{tag}"]
    ref = so.h_ransac_detailed(x1, x2, thr, samples)
    x = np.ascontiguousarray(np.hstack([x1, x2]))
    s = np.ascontiguousarray(samples, np.int32)
    h = len(s)
    H, counts = np.zeros((h, 9)), np.zeros(h, np.int32)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    hs = hostsim.load()
    hs.hs_h4pt(vp(x), vp(s), C.c_int(h), vp(H))
    hs.hs_h_count(vp(x), C.c_int(len(x)), vp(H), C.c_int(h), C.c_double(thr), vp(counts))
    relH = _compare(H, counts, ref, h)
    assert np.median(relH) < 1e-10
    best = int(np.argmax(counts))
    assert best == ref["best_idx"] and counts[best] == ref["best_count"]


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["a", "b"])
def test_gpu_h_ransac_matches_oracle(cuda_device, tag):
    from sfm_b200 import homography
    g = golden("homography.npz")
    x1, x2, thr, samples = g[f"x1_{tag}"], g[f"x2_{tag}"], float(g[f"thr_{tag}"]), g[f"samples_{tag}"]
    ref = so.h_ransac_detailed(x1, x2, thr, samples)
    got = homography.h_ransac_host(x1, x2, samples, thr, want_err=True, want_all=True)
    _compare(got["H"], got["counts"], ref, len(samples))
    assert got["best_idx"] == ref["best_idx"] and got["best_count"] == ref["best_count"]
    assert rel_err(got["H_best"], g[f"H_{tag}"]) < 1e-8
    border = np.abs(ref["err"] - thr) < BAND
    assert np.array_equal(got["mask"][~border], ref["mask"][~border])
    assert got["mask"].sum() == got["best_count"]
    assert np.max(np.abs(got["err"] - ref["err"]) / np.maximum(ref["err"], 1e-9)) < 1e-7


@pytest.mark.gpu
def test_gpu_dropins_and_edge_cases(cuda_device):
    from sfm_b200 import homography
    g = golden("homography.npz")
    x1, x2, idx, thr = g["x1_a"], g["x2_a"], g["idx_a"], float(g["thr_a"])
    n_max = len(g["samples_a"])
    random.seed(int(g["seed_a"]))
    H, inl = homography.get_homography_inliers(x1, x2, idx, threshold=thr, n_max=n_max)
    state = random.getstate()
    random.seed(int(g["seed_a"]))
    so.draw_samples(len(x1), 4, n_max)
    assert random.getstate() == state                       # same draws consumed as the reference
    assert rel_err(H, g["H_a"]) < 1e-8 and np.array_equal(inl, g["inl_a"])
    good = g["good_a"]
    assert rel_err(homography.find_homography(x1[good], x2[good]), g["Hall_a"]) < 1e-8
    assert rel_err(homography.find_homography(x1[:4], x2[:4]), g["H4_a"]) < 1e-8
    with pytest.raises(ValueError):
        homography.find_homography(x1[:3], x2[:3])
    st = random.getstate()
    Hn, inl_n = homography.get_homography_inliers(x1[:3], x2[:3], idx[:3])
    assert Hn is None and len(inl_n) == 0 and random.getstate() == st     # N < 4: no draw (:108-110)
    # nothing within the threshold: (None, [])
    Hn, inl_n = homography.get_homography_inliers(x1[:50], x2[:50] + 1e6 * np.arange(50)[:, None], idx[:50], threshold=1e-9, n_max=20)
    assert Hn is None and len(inl_n) == 0
    # a non-positive threshold never admits an inlier
    Hn, _ = homography.get_homography_inliers(x1[:50], x2[:50], idx[:50], threshold=0.0, n_max=20)
    assert Hn is None


@pytest.mark.gpu
def test_gpu_h_ransac_full_size_properties(cuda_device):
    """C2-sized input (50k correspondences, 100k hypotheses): first maximum, mask = count, order
    independence."""
    import torch
    from sfm_b200 import sampler, device, homography
    rng = np.random.default_rng(9)
    n, h = 50_000, 100_000
    Ht = np.array([[0.95, 0.02, 12.0], [-0.03, 1.05, -7.0], [2e-6, -1e-6, 1.0]])
    x1 = rng.uniform(0, 2000, (n, 2))
    t = (Ht @ np.hstack([x1, np.ones((n, 1))]).T).T
    x2 = t[:, :2] / t[:, 2:3] + rng.normal(0, 0.5, (n, 2))
    bad = rng.random(n) < 0.3
    x2[bad] = rng.uniform(0, 2000, (int(bad.sum()), 2))
    random.seed(4242)
    samples = sampler.sample_stream(n, 4, h)
    d_x, d_s = device.to_device(np.hstack([x1, x2])), device.to_device(samples, np.int32)
    r = homography.h_ransac_device(d_x, d_s, 3.0)
    counts, best = r.counts.cpu().numpy(), r.best.cpu().numpy()
    assert best[1] == counts.max() and best[0] == int(np.argmax(counts))
    assert int(r.mask.sum().item()) == best[1] and best[1] > 0.6 * n
    r2 = homography.h_ransac_device(d_x, torch.flip(d_s, dims=[0]).contiguous(), 3.0)
    assert np.array_equal(r2.counts.cpu().numpy()[::-1], counts)
