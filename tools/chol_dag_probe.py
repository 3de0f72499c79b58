#!/usr/bin/env python
"""Times the structure-aware K11 (sfm_chol_solve_structured) against the dense kernel (sfm_chol_solve) on
cyclic-banded systems of the C5 shape and on dense ones. CUDA events around `reps` launches on one plan."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from sfm_b200 import _lib, device
from test_chol_plan import NB, ring_pattern, spd_from_pattern

lib = _lib.load()
cases = [("dense300", None), ("dense1200", None), ("ring125", ring_pattern(125, 5)), ("ring40", ring_pattern(40, 3)),
         ("band125", ring_pattern(125, 5, cyclic=False))]
only = sys.argv[1:]
for name, pat in cases:
    if only and name not in only:
        continue
    rng = np.random.default_rng(1)
    if pat is None:
        n = int(name[5:])
        A = rng.normal(size=(n, n + 8))
        S = A @ A.T + 0.5 * np.eye(n)
    else:
        S = spd_from_pattern(pat, rng)
        n = S.shape[0]
    g = rng.normal(size=n)
    d_S, d_g = device.to_device(np.tril(S)), device.to_device(g)
    nbytes = int(lib.sfm_chol_solve_workspace_bytes(n))
    ws = device.empty((nbytes,), torch.uint8)
    x = device.empty((n,), torch.float64)
    fail = device.empty((1,), torch.int32)
    st = device.stream()
    ref = np.linalg.solve(S, g)
    out = {"case": name, "n": n}
    for _ in range(2):
        _lib.check(lib.sfm_chol_solve(device.ptr(d_S), device.ptr(d_g), n, device.ptr(x), device.ptr(fail), device.ptr(ws), nbytes, st), "chol")
    reps = 10
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        _lib.check(lib.sfm_chol_solve(device.ptr(d_S), device.ptr(d_g), n, device.ptr(x), device.ptr(fail), device.ptr(ws), nbytes, st), "chol")
    e1.record(); torch.cuda.synchronize()
    out["dense_ms_incl_pack"] = round(e0.elapsed_time(e1) / reps, 4)
    out["dense_err"] = float(np.linalg.norm(x.cpu().numpy() - ref) / np.linalg.norm(ref))
    for mode in (1, 2):
        info = np.zeros(6)
        for reps in (2, 20):
            _lib.check(lib.sfm_chol_solve_structured(device.ptr(d_S), device.ptr(d_g), n, mode, reps, device.ptr(x), device.ptr(fail),
                                                     device.ptr(ws), nbytes, info.ctypes.data, st), "dag")
        tag = "nat" if mode == 1 else "nd"
        out[f"dag_{tag}_ms"] = round(info[4], 4)
        out[f"dag_{tag}_err"] = float(np.linalg.norm(x.cpu().numpy() - ref) / np.linalg.norm(ref))
        out[f"dag_{tag}_info"] = {"order": int(info[0]), "depth": int(info[1]), "nnzL": int(info[2]), "sim_us": round(info[3], 1),
                                  "partial_tasks": int(info[5])}
        out[f"dag_{tag}_fail"] = int(fail.cpu()[0])
    print(json.dumps(out), flush=True)
