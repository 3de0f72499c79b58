#!/usr/bin/env python
"""Times K11 (sfm_chol_solve) alone with CUDA events and prints its in-kernel phase profile."""
import os, sys, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sfm_b200 import _lib, device

lib = _lib.load()
for n in [int(a) for a in sys.argv[1:]] or [300, 1200, 3000, 6000]:
    g = torch.Generator(device="cuda").manual_seed(n)
    A = torch.randn(n, n + 8, dtype=torch.float64, device="cuda", generator=g)
    S = A @ A.T + 0.5 * torch.eye(n, dtype=torch.float64, device="cuda")
    rhs = torch.randn(n, dtype=torch.float64, device="cuda", generator=g)
    nbytes = int(lib.sfm_chol_solve_workspace_bytes(n))
    ws = device.empty((nbytes,), torch.uint8)
    x = device.empty((n,), torch.float64)
    fail = device.empty((1,), torch.int32)
    prof = torch.zeros(8, dtype=torch.int64, device="cuda")
    st = device.stream()
    for _ in range(3):
        _lib.check(lib.sfm_chol_solve(device.ptr(S), device.ptr(rhs), n, device.ptr(x), device.ptr(fail), device.ptr(ws), nbytes, st), "chol")
    reps = 10
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        _lib.check(lib.sfm_chol_solve(device.ptr(S), device.ptr(rhs), n, device.ptr(x), device.ptr(fail), device.ptr(ws), nbytes, st), "chol")
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    _lib.check(lib.sfm_chol_solve_profiled(device.ptr(S), device.ptr(rhs), n, device.ptr(x), device.ptr(fail), device.ptr(ws), nbytes, device.ptr(prof), st), "chol")
    torch.cuda.synchronize()
    pc = prof.cpu().numpy()
    ph = pc[:6] / 1e3
    ref = torch.linalg.solve(S, rhs)
    err = float((x - ref).norm() / ref.norm())
    print(json.dumps({"n": n, "ms_incl_pack": round(ms, 4), "gflops": round(n ** 3 / 3 / ms / 1e6, 1), "rel_err": err, "fail": int(fail.cpu()[0]),
                      "phase_us": dict(zip(["diag", "trsm", "sync1", "update", "sync2", "back"], [round(float(v), 1) for v in ph])),
                      "diag_cycles_per_block": {"sweep": int(pc[6]) // max(1, -(-n // 48)), "factor": int(pc[7]) // max(1, -(-n // 48))}}), flush=True)
