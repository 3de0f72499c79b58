#!/usr/bin/env python
"""BASELINE configs[3] throughput probe: 1M-point linear + non-linear triangulation and 500
NonLinearPnP registrations of 2,000 correspondences each (device-resident inputs, CUDA events)."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from scipy.spatial.transform import Rotation
from sfm_b200 import device, refine
from oracle import scenes


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        out = fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out


n = 1_000_000
sc = scenes.two_view(n=n, outlier_frac=0.0, noise=0.5, seed=20260104)
P1 = refine.projection_matrix(sc["K"], sc["C1"], sc["R1"]); P2 = refine.projection_matrix(sc["K"], sc["C2"], sc["R2"])
d_x1, d_x2 = device.to_device(sc["x1"]), device.to_device(sc["x2"])
ms_lin, d_X0 = timed(lambda: refine.lintri_device(P1, P2, d_x1, d_x2))
ms_nl, (d_X, d_cost, d_st) = timed(lambda: refine.nltri_device(P1, P2, d_x1, d_x2, d_X0, want_info=True))
st = d_st.cpu().numpy()
X = d_X.cpu().numpy()
print(json.dumps({"lintri_Mpts_s": n / ms_lin / 1e3, "lintri_ms": ms_lin, "nltri_Mpts_s": n / ms_nl / 1e3, "nltri_ms": ms_nl,
                  "nltri_mean_nfev": float((st & 0xFFFF).mean()), "status_hist": np.bincount(st >> 16).tolist(),
                  "rel_err_vs_truth": float(np.linalg.norm(X - sc["X"]) / np.linalg.norm(sc["X"]))}), flush=True)
B, M = 500, 2000
rng = np.random.default_rng(20260105)
Xs, xs, rt0 = [], [], []
for b in range(B):
    s = scenes.pnp_scene(n=M, seed=20260105 + b)
    Xs.append(s["X"]); xs.append(s["x"])
    R0 = Rotation.from_rotvec(rng.normal(0, 0.02, 3)).as_matrix() @ s["R"]
    C0 = s["C"] + rng.normal(0, 0.05, 3)
    rt0.append(np.hstack([Rotation.from_matrix(R0).as_rotvec(), -R0 @ C0]))
d_X = device.to_device(np.concatenate(Xs)); d_x = device.to_device(np.concatenate(xs))
d_off = device.to_device(np.arange(0, (B + 1) * M, M, dtype=np.int32), np.int32); d_rt0 = device.to_device(np.array(rt0))
ms_pnp, (d_rt, d_c, d_s) = timed(lambda: refine.nlpnp_batch_device(scenes.K_REF, d_X, d_x, d_off, d_rt0))
s_ = d_s.cpu().numpy()
print(json.dumps({"nlpnp_problems_s": B / ms_pnp * 1e3, "nlpnp_ms": ms_pnp, "mean_nfev": float((s_ & 0xFFFF).mean()),
                  "status_hist": np.bincount(s_ >> 16).tolist()}), flush=True)
