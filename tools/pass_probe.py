#!/usr/bin/env python
"""Times the isolated BA linearisation passes on C5 (one GPU) with CUDA events."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench_ba
from sfm_b200 import bundle
peaks = {"hbm_gbs": 6539.5}
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
sc5, cams5, pts5 = bench_ba._scene(bench_ba.C5)
prob5 = bundle.BAProblem(1000, 500000, sc5["cam_idx"], sc5["pt_idx"], sc5["obs"], sc5["K"])
n5, p5 = len(sc5["obs"]), 500000
for name, which, nbytes in (("K9p", 1, bench_ba.k9p_bytes(n5, p5)), ("K9c", 2, bench_ba.k9c_bytes(n5)), ("K8", 4, bench_ba.k8_bytes(n5))):
    t_ms = bench_ba.time_pass(prob5, cams5, pts5, which, 10, flush)
    print(name, f"{t_ms*1e3:.1f} us", f"{nbytes/(t_ms*1e-3)/1e9:.0f} GB/s", f"{nbytes/(t_ms*1e-3)/1e9/6539.5:.3f} of measured HBM", flush=True)
