#!/usr/bin/env python
"""Repeats the wrapper flow through the drop-in modules (the test_gpu_pipeline scenario) and prints how the
bundle-adjustment costs vary from run to run; then solves one banded SPD system many times through K11s and counts
results that are not bit-identical (K11s has no atomics: any difference is a race)."""
import io
import os
import random
import sys
from contextlib import redirect_stdout

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import wrapper_flow
import test_gpu_pipeline as tp
from sfm_b200 import _lib, device
from test_chol_plan import ring_pattern, spd_from_pattern

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
fx, fy, flag, K = wrapper_flow.synthetic_features()
ns = tp._dropin_ns()
random.seed(11)
with redirect_stdout(io.StringIO()), np.errstate(all="ignore"):
    ref = wrapper_flow.run(tp._oracle_ns(), fx, fy, flag, K)
print("oracle costs", ref["ba_costs"])
costs, starts = [], []
from oracle import ba_oracle as bo
_orig_ba = ns.perform_bundle_adjustment


def _spy(*a):
    prob = bo.ba_build_problem(*a[:7])
    r0 = bo.ba_residuals(prob["x0"], prob["n_cameras"], prob["n_points"], prob["cam_idx"], prob["pt_idx"], prob["pts2d"], K)
    starts[-1].append((len(r0) // 2, 0.5 * float(r0 @ r0), float(np.abs(prob["x0"]).sum())))
    return _orig_ba(*a)


ns.perform_bundle_adjustment = _spy
for r in range(reps):
    random.seed(11)
    starts.append([])
    with redirect_stdout(io.StringIO()), np.errstate(all="ignore"):
        got = wrapper_flow.run(ns, fx, fy, flag, K)
    costs.append(np.array(got["ba_costs"]))
for r in range(min(reps, 12)):
    print("run", r, "final", ["%.4f" % c for c in costs[r]], "start (nobs, cost0, |x0|_1)", [(n, "%.6f" % c, "%.9f" % x) for n, c, x in starts[r]])
costs = np.array(costs)
spread = (costs.max(0) - costs.min(0)) / costs.min(0)
print("rel diff to oracle, worst run per call:", (np.abs(costs - np.array(ref["ba_costs"])) / np.array(ref["ba_costs"])).max(0))
print("ba calls", costs.shape[1], "max relative spread over", reps, "runs:", float(spread.max()))
for k in np.argsort(-spread)[:3]:
    print("  call", int(k), "costs", sorted(set(costs[:, k].tolist()))[:6])

lib = _lib.load()
rng = np.random.default_rng(3)
for name, S in (("ring40", spd_from_pattern(ring_pattern(40, 3), rng)), ("dense240", None)):
    if S is None:
        A = rng.normal(size=(240, 250))
        S = A @ A.T + 0.5 * np.eye(240)
    n = S.shape[0]
    g = rng.normal(size=n)
    d_S, d_g = device.to_device(np.tril(S)), device.to_device(g)
    nbytes = int(lib.sfm_chol_solve_workspace_bytes(n))
    ws = device.empty((nbytes,), torch.uint8)
    x = device.empty((n,), torch.float64)
    fail = device.empty((1,), torch.int32)
    first, bad = None, 0
    for r in range(reps * 10):
        _lib.check(lib.sfm_chol_solve_structured(device.ptr(d_S), device.ptr(d_g), n, 0, 3, device.ptr(x), device.ptr(fail),
                                                 device.ptr(ws), nbytes, None, device.stream()), "dag")
        xv = x.cpu().numpy().copy()
        if first is None:
            first = xv
        elif not np.array_equal(xv, first):
            bad += 1
    print(name, "K11s solves", reps * 10, "not bit-identical to the first:", bad,
          "rel err vs LAPACK", float(np.linalg.norm(first - np.linalg.solve(S, g)) / np.linalg.norm(first)))
