#!/usr/bin/env python
"""End-to-end time of bundle.ba_solve_sparse at C3 (host arrays in/out), call by call (diagnostic)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from scipy.spatial.transform import Rotation
from sfm_b200 import bundle
from oracle import scenes
sc = scenes.ba_scene()
cams0 = np.hstack([Rotation.from_matrix(sc["R0"]).as_rotvec(), -np.einsum("nij,nj->ni", sc["R0"], sc["C0"])])
for i in range(8):
    torch.cuda.synchronize(); t = time.perf_counter()
    cams, pts, info = bundle.ba_solve_sparse(cams0, sc["pts0"], sc["cam_idx"], sc["pt_idx"], sc["obs"], sc["K"])
    dt = time.perf_counter() - t
    print(f"call {i}: {dt*1e3:.2f} ms, {info['n_linearisations']} LM iterations -> {info['n_linearisations']/dt:.0f} it/s e2e")
