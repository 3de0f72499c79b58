#!/usr/bin/env python
"""Throughput of the homography RANSAC kernels at the C2 shape (50k correspondences, 100k hypotheses)."""
import json, os, random, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from sfm_b200 import sampler, device, homography, _lib
rng = np.random.default_rng(9)
n, h = 50_000, 100_000
Ht = np.array([[0.95, 0.02, 12.0], [-0.03, 1.05, -7.0], [2e-6, -1e-6, 1.0]])
x1 = rng.uniform(0, 2000, (n, 2))
t = (Ht @ np.hstack([x1, np.ones((n, 1))]).T).T
x2 = t[:, :2] / t[:, 2:3] + rng.normal(0, 0.5, (n, 2))
bad = rng.random(n) < 0.3
x2[bad] = rng.uniform(0, 2000, (int(bad.sum()), 2))
random.seed(4242)
samples = sampler.sample_stream(n, 4, h)
d_x, d_s = device.to_device(np.hstack([x1, x2])), device.to_device(samples, np.int32)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    r = homography.h_ransac_device(d_x, d_s, 3.0)
torch.cuda.synchronize()
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
for a, b in ev:
    flush.zero_()
    a.record()
    r = homography.h_ransac_device(d_x, d_s, 3.0)
    b.record()
torch.cuda.synchronize()
ms = sum(a.elapsed_time(b) for a, b in ev) / len(ev)
print(json.dumps({"ms_per_100k_hypotheses": ms, "hypotheses_per_s": h / (ms * 1e-3), "tests_per_s": h * n / (ms * 1e-3),
                  "winner": r.best.cpu().numpy().tolist()}))
