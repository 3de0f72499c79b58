#!/usr/bin/env python
"""BASELINE configs[3] (C4) for the bench line: batched LinearTriangulation + NonLinearTriangulation over
1M points and 500 NonLinearPnP registrations of 2,000 correspondences each, device-resident inputs timed
with CUDA events (L2 flushed between repetitions), next to the CPU oracle port of the reference
(NonLinearTriangulation.py / NonlinearPnP.py: one scipy least_squares per point / per pose) on a bounded
sample. Imported by bench.py (N = 1 only); `python bench_c4.py` prints the object alone."""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

N_PTS, N_REG, N_CORR = 1_000_000, 500, 2000


def _timed(fn, flush, reps=5):
    import torch
    fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    out = None
    for a, b in ev:
        if flush is not None:
            flush.zero_()
        a.record()
        out = fn()
        b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / reps, out


def cpu_baseline(sc, P_pts=3000, n_pnp=20):
    """The oracle port on one core: P_pts points of the same scene, n_pnp registrations."""
    from oracle import sfm_oracle as so
    from oracle import scenes
    idx = np.arange(P_pts)
    t0 = time.perf_counter()
    X0 = so.linear_triangulation(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"], sc["x1"][idx], sc["x2"][idx])
    t_lin = time.perf_counter() - t0
    t0 = time.perf_counter()
    so.nonlinear_triangulation(sc["K"], sc["C1"], sc["R1"], sc["C2"], sc["R2"], sc["x1"][idx], sc["x2"][idx], np.asarray(X0)[:, :3])
    t_nl = time.perf_counter() - t0
    t0 = time.perf_counter()
    for b in range(n_pnp):
        s = scenes.pnp_scene(n=N_CORR, seed=20260105 + b)
        so.nonlinear_pnp(s["X"], s["x"], s["K"] if "K" in s else scenes.K_REF, s["C"] + 0.05, s["R"])
    t_pnp = time.perf_counter() - t0
    return {"lintri_points_s": P_pts / t_lin, "nltri_points_s": P_pts / t_nl, "nlpnp_registrations_s": n_pnp / t_pnp,
            "cores": 1, "kind": "port",
            "sample": f"{P_pts} points of the same scene, {n_pnp} registrations of {N_CORR} correspondences "
                      f"({t_lin + t_nl + t_pnp:.1f} s)"}


def run(flush=None, with_cpu=True):
    import torch
    from scipy.spatial.transform import Rotation
    from sfm_b200 import device, refine
    from oracle import scenes
    sc = scenes.two_view(n=N_PTS, outlier_frac=0.0, noise=0.5, seed=20260104)
    P1 = refine.projection_matrix(sc["K"], sc["C1"], sc["R1"])
    P2 = refine.projection_matrix(sc["K"], sc["C2"], sc["R2"])
    d_x1, d_x2 = device.to_device(sc["x1"]), device.to_device(sc["x2"])
    ms_lin, d_X0 = _timed(lambda: refine.lintri_device(P1, P2, d_x1, d_x2), flush)
    ms_nl, (d_X, d_cost, d_st) = _timed(lambda: refine.nltri_device(P1, P2, d_x1, d_x2, d_X0, want_info=True), flush)
    st = d_st.cpu().numpy()
    X = d_X.cpu().numpy()
    rng = np.random.default_rng(20260105)
    Xs, xs, rt0 = [], [], []
    for b in range(N_REG):
        s = scenes.pnp_scene(n=N_CORR, seed=20260105 + b)
        Xs.append(s["X"])
        xs.append(s["x"])
        R0 = Rotation.from_rotvec(rng.normal(0, 0.02, 3)).as_matrix() @ s["R"]
        C0 = s["C"] + rng.normal(0, 0.05, 3)
        rt0.append(np.hstack([Rotation.from_matrix(R0).as_rotvec(), -R0 @ C0]))
    d_Xp, d_xp = device.to_device(np.concatenate(Xs)), device.to_device(np.concatenate(xs))
    d_off = device.to_device(np.arange(0, (N_REG + 1) * N_CORR, N_CORR, dtype=np.int32), np.int32)
    d_rt0 = device.to_device(np.array(rt0))
    ms_pnp, (d_rt, d_c, d_s) = _timed(lambda: refine.nlpnp_batch_device(scenes.K_REF, d_Xp, d_xp, d_off, d_rt0), flush)
    s_ = d_s.cpu().numpy()
    out = {"workload": f"C4: LinearTriangulation + NonLinearTriangulation over {N_PTS} points, {N_REG} NonLinearPnP "
                       f"registrations of {N_CORR} correspondences (device-resident, CUDA events, L2 flushed)",
           "dtype": "f64",
           "lintri_points_s": N_PTS / (ms_lin * 1e-3), "lintri_ms": ms_lin,
           "nltri_points_s": N_PTS / (ms_nl * 1e-3), "nltri_ms": ms_nl,
           "nltri_mean_nfev": float((st & 0xFFFF).mean()),
           "nltri_rel_err_vs_truth": float(np.linalg.norm(X - sc["X"]) / np.linalg.norm(sc["X"])),
           "nlpnp_registrations_s": N_REG / (ms_pnp * 1e-3), "nlpnp_ms": ms_pnp,
           "nlpnp_mean_nfev": float((s_ & 0xFFFF).mean()),
           "algorithmic_gbs": {"lintri (56 B/point)": 56.0 * N_PTS / (ms_lin * 1e-3) / 1e9,
                               "nltri (80 B/point)": 80.0 * N_PTS / (ms_nl * 1e-3) / 1e9},
           "bound": "FP64 latency / issue (thread per point, register-resident LM); not an HBM stream"}
    # PnPRANSAC (reference defaults threshold=200, n_max=1000), host arrays in / out, ALL registrations: hypotheses
    # (K6a LinearPnP on the 4-point samples), scoring (K6) and argmax (K3) on the GPU; the sample stream is the
    # reference's `random` stream replayed on the host
    import random
    from sfm_b200 import pnp_ransac
    random.seed(777)
    pnp_ransac.pnp_ransac_batch(Xs, xs, scenes.K_REF)                         # warm at full size (allocator, pinned pages)
    torch.cuda.synchronize()
    random.seed(777)
    t0 = time.perf_counter()
    Cb, Rb, counts = pnp_ransac.pnp_ransac_batch(Xs, xs, scenes.K_REF)
    torch.cuda.synchronize()
    out["pnp_ransac_registrations_s"] = N_REG / (time.perf_counter() - t0)
    out["pnp_ransac_mean_inliers"] = float(counts.mean())
    n_seq = 40
    random.seed(777)
    t0 = time.perf_counter()
    for b in range(n_seq):
        pnp_ransac.PnPRANSAC(Xs[b], xs[b], scenes.K_REF)
    torch.cuda.synchronize()
    out["pnp_ransac_one_call_per_registration_s"] = n_seq / (time.perf_counter() - t0)
    out["pnp_ransac_note"] = (f"all {N_REG} registrations of {N_CORR} correspondences x 1000 hypotheses end to end from host "
                              f"arrays through pnp_ransac_batch (one hypothesis launch for all registrations); "
                              f"'one_call_per_registration': the drop-in PnPRANSAC signature, {n_seq} calls")
    if with_cpu:
        out["cpu_baseline"] = cpu_baseline(sc)
        from oracle import sfm_oracle as so
        random.seed(777)
        t0 = time.perf_counter()
        for b in range(4):
            so.pnp_ransac(Xs[b], xs[b], scenes.K_REF, 200, 1000)
        out["cpu_baseline"]["pnp_ransac_registrations_s"] = 4 / (time.perf_counter() - t0)
    return out


if __name__ == "__main__":
    print(json.dumps(run()))
