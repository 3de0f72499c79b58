#!/usr/bin/env python
"""Cycle breakdown of the K11 diagonal-block factor (diagnostic)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sfm_b200 import _lib, device
lib = _lib.load()
A = torch.randn(48, 60, dtype=torch.float64, device="cuda")
S = (A @ A.T + torch.eye(48, dtype=torch.float64, device="cuda")).contiguous()
for v, name in [(0, "full"), (1, "no rsqrt"), (2, "no L stores"), (3, "bar over 128"), (4, "no publish"), (5, "syncwarp only")]:
    out = torch.zeros(8, dtype=torch.int64, device="cuda")
    for _ in range(3):
        out.zero_()
        _lib.check(lib.sfm_chol_diag_bench(device.ptr(S), v, device.ptr(out), device.stream()), "bench")
    torch.cuda.synchronize()
    o = out.cpu().numpy()
    print(f"{name:20s} sweep {int(o[6]):7d} cycles ({int(o[6])/48:7.1f}/column)  factor {int(o[7]):7d}")
