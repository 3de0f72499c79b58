"""One resident F-RANSAC step at C2 (for ncu: capture f_score_packed_kernel)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import bench_config as cfg
from sfm_b200 import ransac, device, sampler
import random
sc = cfg.two_view_scene()
N, H = len(sc["x1"]), 100000
random.seed(1)
s = sampler.sample_stream(N, 8, H)
d_x = torch.cat([device.to_device(sc["x1"]), device.to_device(sc["x2"])], dim=1)
d_s = device.to_device(s, np.int32)
ctx = ransac.FRansacPlan(N, H)
for _ in range(2):
    ctx.run(d_x, d_s, 0.06)
torch.cuda.synchronize()
print("ok")
