#!/usr/bin/env python
"""One C4-sized batch of LinearPnP hypotheses (500 registrations x 1000 four-point samples) + scoring, timed with
CUDA events; used under ncu for the K6a capture in profiles/."""
import os
import random
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import scenes
from sfm_b200 import device, pnp_ransac, sampler

B, M, H = 500, 2000, 1000
regs = [scenes.pnp_scene(n=M, seed=20260105 + b) for b in range(B)]
d_X = device.to_device(np.concatenate([r["X"] for r in regs]))
d_x = device.to_device(np.concatenate([r["x"] for r in regs]))
offs = np.arange(0, (B + 1) * M, M, dtype=np.int32)
random.seed(777)
samples = np.concatenate([sampler.sample_stream(M, 4, H) for _ in range(B)])
d_s, d_off = device.to_device(samples, np.int32), device.to_device(offs, np.int32)
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    d_C, d_R, d_P, d_flag = pnp_ransac.linear_pnp4_device(d_X, d_x, d_s, scenes.K_REF, d_off, B)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
print(f"linear_pnp4: {B * H} hypotheses in {ms:.3f} ms = {B * H / ms / 1e3:.1f} M hypotheses/s; reflected fraction "
      f"{float((d_flag & 1).float().mean()):.3f}")
