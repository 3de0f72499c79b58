"""GPU parity tests of LinearPnP on the device (a8: csrc/linpnp.cu through the C ABI) against outputs of the unmodified
reference frozen in tests/golden/linpnp4.npz / pnp.npz (LinearPnP.py:3-96, PnPRANSAC.py:6-89) and the oracle.
Pose tolerance: north_star's 1e-6 relative. Hypotheses flagged REFLECTED (LinearPnP.py:88-90, decided by rounding
noise in the reference -- csrc/linpnp_math.cuh) are only required to be proper rotations; everything else must match."""
import io
import random
from contextlib import redirect_stdout

import numpy as np
import pytest

from conftest import golden, rel_err
from oracle import sfm_oracle as so

pytestmark = pytest.mark.gpu
POSE_RTOL = 1e-6


def _close(Cg, Rg, Cr, Rr):
    return rel_err(Rg, Rr) < POSE_RTOL and np.linalg.norm(Cg - Cr) <= POSE_RTOL * max(1.0, np.linalg.norm(Cr))


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_four_point_hypotheses_match_reference_goldens(cuda_device, tag):
    from sfm_b200 import device, pnp_ransac
    g, ref = golden("pnp.npz"), golden("linpnp4.npz")
    X, x, K, samples = g[f"X_{tag}"], g[f"x_{tag}"], g["K"], g[f"samples_{tag}"]
    d_X, d_x = device.to_device(X), device.to_device(x)
    d_C, d_R, d_P, d_flag = pnp_ransac.linear_pnp4_device(d_X, d_x, device.to_device(samples, np.int32), K)
    Cg, Rg, Pg, flag = d_C.cpu().numpy(), d_R.cpu().numpy().reshape(-1, 3, 3), d_P.cpu().numpy(), d_flag.cpu().numpy()
    pinned = (flag & 1) == 0
    assert pinned.sum() >= 0.5 * len(samples)
    for i in range(len(samples)):
        assert np.allclose(Rg[i] @ Rg[i].T, np.eye(3), atol=1e-12) and abs(np.linalg.det(Rg[i]) - 1) < 1e-12
        if pinned[i]:
            assert np.allclose(Pg[i].reshape(3, 4), K @ np.hstack([Rg[i], (-Rg[i] @ Cg[i])[:, None]]), rtol=1e-12, atol=1e-9)
        else:
            assert np.all(np.isnan(Pg[i]))                       # parity policy: a reflected hypothesis does not vote
        if pinned[i] and np.all(np.isfinite(ref[f"R4_{tag}"][i])):
            assert _close(Cg[i], Rg[i], ref[f"C4_{tag}"][i], ref[f"R4_{tag}"][i]), (tag, i)
    # consensus counts of the pinned hypotheses equal the reference's; the reference's winner is a pinned one
    d = so.pnp_ransac_detailed(X, x, K, 200, samples)
    d_counts, d_best = pnp_ransac.pnp_score_device(d_P, d_X, d_x, 200)
    counts = d_counts.cpu().numpy()
    assert np.array_equal(counts[pinned], d["counts"][pinned]) and np.all(counts[~pinned] == 0)
    # policy 0: the same hypotheses vote with their documented pose
    _, _, d_P0, _ = pnp_ransac.linear_pnp4_device(d_X, d_x, device.to_device(samples, np.int32), K, mute_reflected=False)
    P0 = d_P0.cpu().numpy()
    assert np.all(np.isfinite(P0[~pinned])) and np.array_equal(P0[pinned], Pg[pinned])
    best = d_best.cpu().numpy()
    assert best[0] == d["best_idx"] and best[1] == d["best_count"] and pinned[d["best_idx"]]


def test_four_point_kernel_equals_its_host_build(cuda_device):
    """The device kernel against the same source compiled for the host (tests/hostsim): identical up to FMA
    contraction, on every sample including the reflected ones and a degenerate one."""
    import hostsim  # noqa: F401
    from test_hostsim_linpnp import hs_small
    from sfm_b200 import device, pnp_ransac
    g = golden("pnp.npz")
    X, x, K = g["X_b"], g["x_b"], g["K"]
    samples = np.array(g["samples_b"][:64], np.int32)
    samples[5] = [7, 7, 7, 7]                                       # degenerate: four times the same correspondence
    d_C, d_R, d_P, d_flag = pnp_ransac.linear_pnp4_device(device.to_device(X), device.to_device(x),
                                                          device.to_device(samples, np.int32), K)
    Cg, Rg, flag = d_C.cpu().numpy(), d_R.cpu().numpy().reshape(-1, 3, 3), d_flag.cpu().numpy()
    for i, s in enumerate(samples):
        Ch, Rh, _, fh = hs_small(K, X[s], x[s], 0)
        if i == 5 or fh & 2 or flag[i] & 2:                          # degenerate: any finite answer is as good as another
            assert i == 5 or (fh & 2) == (flag[i] & 2)
            continue
        assert (fh & 1) == (flag[i] & 1)
        assert np.allclose(Rg[i], Rh, atol=1e-9) and np.allclose(Cg[i], Ch, rtol=1e-8, atol=1e-8), i


@pytest.mark.parametrize("n", [4, 5, 6, 7, 8, 12, 50, 2000])
def test_linear_pnp_all_points_matches_reference(cuda_device, n):
    from sfm_b200 import pnp_ransac
    g, ref = golden("pnp.npz"), golden("linpnp4.npz")
    X, x, K = g["X_b"][:n], g["x_b"][:n], g["K"]
    C, R = pnp_ransac.LinearPnP(X, x, K)
    assert C.shape == (3,) and R.shape == (3, 3)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-10) and abs(np.linalg.det(R) - 1) < 1e-10
    Cr, Rr = (ref[f"Cn_{n}"], ref[f"Rn_{n}"]) if n > 4 else so.linear_pnp(X, x, K)
    from test_hostsim_linpnp import reference_takes_reflection_branch
    if not reference_takes_reflection_branch(X, x, K):
        assert _close(C, R, Cr, Rr), (n, rel_err(R, Rr))


def test_linear_pnp_random_subsets_match_oracle(cuda_device):
    """N-point Jacobi path (6 .. 40 correspondences) wherever the reference is a function of its input."""
    from sfm_b200 import pnp_ransac
    from test_hostsim_linpnp import reference_takes_reflection_branch
    g = golden("pnp.npz")
    X, x, K = g["X_b"], g["x_b"], g["K"]
    rng = np.random.default_rng(8)
    checked = 0
    for _ in range(30):
        idx = rng.choice(2000, int(rng.integers(6, 41)), replace=False)
        if reference_takes_reflection_branch(X[idx], x[idx], K):
            continue
        C, R = pnp_ransac.LinearPnP(X[idx], x[idx], K)
        Cr, Rr = so.linear_pnp(X[idx], x[idx], K)
        assert _close(C, R, Cr, Rr), (len(idx), rel_err(R, Rr))
        checked += 1
    assert checked >= 8


def test_linear_pnp_dropin_module_and_errors(cuda_device):
    import importlib
    import os
    import sys
    dropin = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "structure-from-motion-_b200", "dropin")
    sys.path.insert(0, dropin)
    try:
        sys.modules.pop("LinearPnP", None)
        mod = importlib.import_module("LinearPnP")
    finally:
        sys.path.remove(dropin)
        sys.modules.pop("LinearPnP", None)
    g = golden("pnp.npz")
    C, R = mod.LinearPnP(g["X_c"], g["x_c"], g["K"])
    from test_hostsim_linpnp import reference_takes_reflection_branch
    if not reference_takes_reflection_branch(g["X_c"], g["x_c"], g["K"]):
        assert _close(C, R, g["Clin_c"], g["Rlin_c"])
    with pytest.raises(ValueError, match="At least 4 point correspondences"):
        mod.LinearPnP(g["X_c"][:3], g["x_c"][:3], g["K"])


def test_pnp_ransac_batch_equals_consecutive_calls(cuda_device):
    from oracle import scenes
    from sfm_b200 import pnp_ransac
    regs = [scenes.pnp_scene(n=300 + 40 * b, seed=900 + b) for b in range(6)]
    Xs, xs = [r["X"] for r in regs], [r["x"] for r in regs]
    random.seed(4242)
    seq = [pnp_ransac.PnPRANSAC(X, x, scenes.K_REF, 200, 250) for X, x in zip(Xs, xs)]
    after_seq = random.random()
    random.seed(4242)
    Cb, Rb, counts = pnp_ransac.pnp_ransac_batch(Xs, xs, scenes.K_REF, 200, 250)
    assert random.random() == after_seq                       # the global stream advanced exactly as the calls did
    for b, (C, R) in enumerate(seq):
        assert np.array_equal(C, Cb[b]) and np.array_equal(R, Rb[b])
        assert counts[b] >= 0.8 * len(Xs[b])
    # and the oracle picks the same winners from the same stream
    random.seed(4242)
    for b in range(6):
        Co, Ro = so.pnp_ransac(Xs[b], xs[b], scenes.K_REF, 200, 250)
        assert _close(Cb[b], Rb[b], Co, Ro), b


def test_pnp_ransac_falls_back_to_all_points(cuda_device):
    """PnPRANSAC.py:82-87: fewer than 4 inliers for every hypothesis -> LinearPnP on all points + the warning."""
    from sfm_b200 import pnp_ransac
    g = golden("pnp.npz")
    X, x, K = g["X_a"], g["x_a"], g["K"]
    buf = io.StringIO()
    random.seed(3)
    with redirect_stdout(buf):
        C, R = pnp_ransac.PnPRANSAC(X, x, K, threshold=1e-9, n_max=20)
    assert "Warning: PnP RANSAC failed, using linear PnP on all points" in buf.getvalue()
    Ca, Ra = pnp_ransac.LinearPnP(X, x, K)
    assert np.array_equal(C, Ca) and np.array_equal(R, Ra)
    random.seed(3)
    with redirect_stdout(io.StringIO()):
        C0, R0 = pnp_ransac.PnPRANSAC(X, x, K, threshold=200, n_max=0)          # no iterations at all: same fallback
    assert np.array_equal(C0, Ca)


def test_batched_scoring_and_argmax_equal_the_per_registration_calls(cuda_device):
    """sfm_pnp_score_count_batch / sfm_argmax_first_batch (one launch for all registrations of a batch) against
    sfm_pnp_score_count + sfm_argmax_first per registration: ragged sizes around the 512-row chunk, a registration
    without any inlier, hypothesis counts that are not a multiple of the block."""
    import torch
    from oracle import scenes
    from sfm_b200 import _lib, device, pnp_ransac
    lib = _lib.load()
    sizes = [4, 511, 512, 513, 37, 2000, 1025]
    h = 300
    regs = [scenes.pnp_scene(n=max(n, 8), seed=70 + i) for i, n in enumerate(sizes)]
    Xs = [r["X"][:n] for r, n in zip(regs, sizes)]
    xs = [r["x"][:n].copy() for r, n in zip(regs, sizes)]
    xs[4] += 1e6                                                     # nothing reprojects within the threshold
    offs = np.zeros(len(sizes) + 1, np.int32)
    offs[1:] = np.cumsum(sizes)
    rng = np.random.default_rng(3)
    samples = np.concatenate([np.stack([rng.choice(n, 4, replace=False) for _ in range(h)]) for n in sizes]).astype(np.int32)
    d_X, d_x = device.to_device(np.concatenate(Xs)), device.to_device(np.concatenate(xs))
    d_off = device.to_device(offs, np.int32)
    _, _, d_P, _ = pnp_ransac.linear_pnp4_device(d_X, d_x, device.to_device(samples, np.int32), scenes.K_REF, d_off, len(sizes))
    d_counts = device.empty((len(sizes) * h,), torch.int32)
    d_best = device.empty((len(sizes), 2), torch.int32)
    _lib.check(lib.sfm_pnp_score_count_batch(device.ptr(d_P), device.ptr(d_X), device.ptr(d_x), device.ptr(d_off),
                                             len(sizes), max(sizes), h, 200.0, device.ptr(d_counts), device.stream()), "batch")
    _lib.check(lib.sfm_argmax_first_batch(device.ptr(d_counts), len(sizes), h, device.ptr(d_best), device.stream()), "argmax")
    counts, best = d_counts.cpu().numpy().reshape(len(sizes), h), d_best.cpu().numpy()
    for b, n in enumerate(sizes):
        c1, b1 = pnp_ransac.pnp_score_device(d_P[b * h:(b + 1) * h], d_X[offs[b]:offs[b + 1]], d_x[offs[b]:offs[b + 1]], 200.0)
        assert np.array_equal(counts[b], c1.cpu().numpy()), b
        assert np.array_equal(best[b], b1.cpu().numpy()), b
    assert best[4, 1] == 0 and counts[5].max() > 1500
