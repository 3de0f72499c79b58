#!/usr/bin/env python
"""Timing probe for the BA chain on BASELINE configs 3 / 5 (single GPU): solves the synthetic
scene, prints iterations, time, LM iterations/s. Used under ncu for the launch lists in profiles/."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c3", choices=["c3", "c5", "small"])
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--tol", type=float, default=1e-8)
    a = ap.parse_args()
    import torch
    from scipy.spatial.transform import Rotation
    from sfm_b200 import bundle, device
    from oracle import scenes
    t0 = time.time()
    if a.config == "c3":
        sc = scenes.ba_scene()
    elif a.config == "c5":
        sc = scenes.ba_scene(n_cams=1000, n_pts=500_000, obs_per_pt=8, seed=20260106, layout="ring", window=40)
    else:
        sc = scenes.ba_scene(n_cams=12, n_pts=300, obs_per_pt=6, seed=9)
    nc, npts, nobs = len(sc["R0"]), len(sc["pts0"]), len(sc["obs"])
    cams0 = np.hstack([Rotation.from_matrix(sc["R0"]).as_rotvec(), -np.einsum("nij,nj->ni", sc["R0"], sc["C0"])])
    print(f"scene {a.config}: {nc} cams {npts} pts {nobs} obs, generated in {time.time()-t0:.1f}s", flush=True)
    prob = bundle.BAProblem(nc, npts, sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    print(f"workspace {prob.ws_bytes/2**20:.1f} MiB", flush=True)
    for rep in range(a.reps):
        d_cam, d_pts = device.to_device(cams0), device.to_device(sc["pts0"])
        torch.cuda.synchronize()
        t = time.perf_counter()
        info = prob.solve_device(d_cam, d_pts, tol=a.tol, verbose=a.verbose and rep == 0)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t
        info.update(seconds=dt, lm_iters_per_s=info["n_linearisations"] / dt, trials_per_s=info["n_trials"] / dt,
                    rmse=float(np.sqrt(2 * info["cost"] / nobs)))
        print(json.dumps(info), flush=True)


if __name__ == "__main__":
    main()
